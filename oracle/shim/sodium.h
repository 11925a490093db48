/* oracle/shim/sodium.h — libsodium 1.0.15 is an EXTERNAL dependency of the reference
 * (depends/packages/libsodium.mk:6-9) and is not vendored under /root/reference, nor installed.
 * This header declares only the three generichash calls the Equihash path uses
 * (crypto/equihash.cpp:44-59, pow.cpp:304-315); they are implemented by the oracle's own
 * RFC 7693 restatement in oracle/bcp_oracle.c.  TEST INFRASTRUCTURE ONLY. */
#ifndef BCP_ORACLE_SHIM_SODIUM_H
#define BCP_ORACLE_SHIM_SODIUM_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define crypto_generichash_blake2b_PERSONALBYTES 16U
#define crypto_generichash_blake2b_SALTBYTES 16U
typedef struct crypto_generichash_blake2b_state {
    uint64_t h[8];
    uint64_t t[2];
    uint64_t f[2];
    uint8_t buf[256];
    size_t buflen;
    uint8_t outlen;
} crypto_generichash_blake2b_state;
int crypto_generichash_blake2b_init_salt_personal(crypto_generichash_blake2b_state *state,
        const unsigned char *key, const size_t keylen, const size_t outlen,
        const unsigned char *salt, const unsigned char *personal);
int crypto_generichash_blake2b_update(crypto_generichash_blake2b_state *state,
        const unsigned char *in, unsigned long long inlen);
int crypto_generichash_blake2b_final(crypto_generichash_blake2b_state *state,
        unsigned char *out, const size_t outlen);
static inline int sodium_init(void) { return 0; }
/* crypto/common.h:97-134 defines an (uncalled here) init_and_check_sodium() that names these; they only
 * need to compile.  Ed25519 is not on the hot path and is NOT implemented. */
#define crypto_sign_PUBLICKEYBYTES 32U
#define crypto_sign_SECRETKEYBYTES 64U
#define crypto_sign_BYTES 64U
static inline int crypto_sign_keypair(unsigned char *pk, unsigned char *sk) { (void)pk; (void)sk; return -1; }
static inline int crypto_sign_detached(unsigned char *sig, unsigned long long *siglen, const unsigned char *m,
                                       unsigned long long mlen, const unsigned char *sk) {
    (void)sig; (void)siglen; (void)m; (void)mlen; (void)sk; return -1;
}
static inline int crypto_sign_verify_detached(const unsigned char *sig, const unsigned char *m,
                                              unsigned long long mlen, const unsigned char *pk) {
    (void)sig; (void)m; (void)mlen; (void)pk; return -1;
}
#ifdef __cplusplus
}
#endif
#endif
