/* oracle/shim: minimal stand-in for the autotools-generated bitcoin-config.h so the
 * reference's hot-path sources compile IN PLACE from /root/reference (never copied).
 * TEST INFRASTRUCTURE ONLY (see oracle/README.md). */
#ifndef BCP_ORACLE_SHIM_CONFIG_H
#define BCP_ORACLE_SHIM_CONFIG_H
#define HAVE_ENDIAN_H 1
#define HAVE_BYTESWAP_H 1
#define HAVE_DECL_HTOBE16 1
#define HAVE_DECL_HTOLE16 1
#define HAVE_DECL_BE16TOH 1
#define HAVE_DECL_LE16TOH 1
#define HAVE_DECL_HTOBE32 1
#define HAVE_DECL_HTOLE32 1
#define HAVE_DECL_BE32TOH 1
#define HAVE_DECL_LE32TOH 1
#define HAVE_DECL_HTOBE64 1
#define HAVE_DECL_HTOLE64 1
#define HAVE_DECL_BE64TOH 1
#define HAVE_DECL_LE64TOH 1
#define HAVE_DECL_BSWAP_16 1
#define HAVE_DECL_BSWAP_32 1
#define HAVE_DECL_BSWAP_64 1
#define HAVE_DECL___BUILTIN_CLZL 1
#define HAVE_DECL___BUILTIN_CLZLL 1
#define HAVE_DECL_STRNLEN 1
#define CLIENT_VERSION_MAJOR 0
#define CLIENT_VERSION_MINOR 17
#define CLIENT_VERSION_REVISION 0
#define CLIENT_VERSION_BUILD 0
#define CLIENT_VERSION_IS_RELEASE true
#define COPYRIGHT_YEAR 2018
#endif
