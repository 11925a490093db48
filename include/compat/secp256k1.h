/* include/compat/secp256k1.h — level B3 of the boundary: the verify-side subset of the
 * libsecp256k1 C API that the node's src/pubkey.cpp uses, with the reference's exact names, argument
 * meaning and return conventions (src/secp256k1/include/secp256k1.h:46-70, 176-610), implemented on top of
 * libbcp_b200.so (bitcoincashplus_b200/host/secp256k1_compat.cpp).  Put this directory in front of
 * src/secp256k1/include on the include path and src/pubkey.cpp compiles unchanged and links against
 * libbcp_secp256k1_compat.so alone (tests/test_compat_cpp.py builds exactly that: the reference's pubkey.cpp over these
 * headers, CPubKey::Verify / RecoverCompact / Derive / Decompress / CheckLowS running on the GPU).  Everything
 * pubkey.cpp calls is here and in secp256k1_recovery.h next to this file.  The signing side (src/key.cpp:
 * secp256k1_ecdsa_sign, ec_pubkey_create, privkey tweaks) is NOT: a wallet keeps the stock library for that, built
 * with a symbol prefix or as a separate shared object, since both define the names below.
 *
 * All curve arithmetic runs on the GPU; there is no CPU fallback.  A single secp256k1_ecdsa_verify is a
 * batch of one (correct, but latency-bound): block validation should go through
 * secp256k1_ecdsa_verify_batch (an extension, below) or the CCheckQueue drop-in (compat/checkqueue.h).
 *
 * Opaque types stay caller-owned 64-byte PODs, freely copyable, layout implementation-defined exactly as in
 * the reference (:46-70): here secp256k1_pubkey holds affine x || y big-endian, secp256k1_ecdsa_signature
 * holds r || s big-endian.
 */
#ifndef BCP_COMPAT_SECP256K1_H
#define BCP_COMPAT_SECP256K1_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct secp256k1_context_struct secp256k1_context;
typedef struct { unsigned char data[64]; } secp256k1_pubkey;
typedef struct { unsigned char data[64]; } secp256k1_ecdsa_signature;

#define SECP256K1_FLAGS_TYPE_CONTEXT (1 << 0)
#define SECP256K1_FLAGS_TYPE_COMPRESSION (1 << 1)
#define SECP256K1_FLAGS_BIT_CONTEXT_VERIFY (1 << 8)
#define SECP256K1_FLAGS_BIT_CONTEXT_SIGN (1 << 9)
#define SECP256K1_FLAGS_BIT_COMPRESSION (1 << 8)
#define SECP256K1_CONTEXT_VERIFY (SECP256K1_FLAGS_TYPE_CONTEXT | SECP256K1_FLAGS_BIT_CONTEXT_VERIFY)
#define SECP256K1_CONTEXT_NONE (SECP256K1_FLAGS_TYPE_CONTEXT)
#define SECP256K1_EC_COMPRESSED (SECP256K1_FLAGS_TYPE_COMPRESSION | SECP256K1_FLAGS_BIT_COMPRESSION)
#define SECP256K1_EC_UNCOMPRESSED (SECP256K1_FLAGS_TYPE_COMPRESSION)

#if defined(__GNUC__)
#define SECP256K1_API __attribute__((visibility("default")))
#else
#define SECP256K1_API
#endif

/* secp256k1.c:59-105.  The context selects the CUDA device (environment BCP_DEVICE, default 0) and owns
 * nothing else; creating it initialises the device tables (the analogue of ecmult_context_build).
 * Returns NULL when no CUDA device is available. */
SECP256K1_API secp256k1_context *secp256k1_context_create(unsigned int flags);
SECP256K1_API void secp256k1_context_destroy(secp256k1_context *ctx);
/* secp256k1.c:107-119: callback invoked on NULL/illegal arguments before the call returns 0
 * (default: print to stderr and abort, like secp256k1.c:28-37). */
SECP256K1_API void secp256k1_context_set_illegal_callback(secp256k1_context *ctx, void (*fun)(const char *message, void *data),
                                                          const void *data);

/* secp256k1.c:150-163 */
SECP256K1_API int secp256k1_ec_pubkey_parse(const secp256k1_context *ctx, secp256k1_pubkey *pubkey, const unsigned char *input,
                                            size_t inputlen);
/* secp256k1.c:165-186 */
SECP256K1_API int secp256k1_ec_pubkey_serialize(const secp256k1_context *ctx, unsigned char *output, size_t *outputlen,
                                                const secp256k1_pubkey *pubkey, unsigned int flags);
/* secp256k1.c:228-247: returns 0 and zeroes the signature when r or s >= n */
SECP256K1_API int secp256k1_ecdsa_signature_parse_compact(const secp256k1_context *ctx, secp256k1_ecdsa_signature *sig,
                                                          const unsigned char *input64);
/* secp256k1.c:261-272 */
SECP256K1_API int secp256k1_ecdsa_signature_serialize_compact(const secp256k1_context *ctx, unsigned char *output64,
                                                              const secp256k1_ecdsa_signature *sig);
/* secp256k1.c:274-291: returns 1 if sigin was high-S; sigout (may be NULL) receives the low-S form */
SECP256K1_API int secp256k1_ecdsa_signature_normalize(const secp256k1_context *ctx, secp256k1_ecdsa_signature *sigout,
                                                      const secp256k1_ecdsa_signature *sigin);
/* secp256k1.c:293-308 */
SECP256K1_API int secp256k1_ecdsa_verify(const secp256k1_context *ctx, const secp256k1_ecdsa_signature *sig,
                                         const unsigned char *msg32, const secp256k1_pubkey *pubkey);

/* secp256k1.c:477-499 (declared src/secp256k1/include/secp256k1.h:544): pubkey := pubkey + tweak*G.  Returns 0 (and
 * zeroes the key) when the tweak is >= n or the result is the point at infinity. */
SECP256K1_API int secp256k1_ec_pubkey_tweak_add(const secp256k1_context *ctx, secp256k1_pubkey *pubkey, const unsigned char *tweak);

/* Extension (no reference counterpart): n independent secp256k1_ecdsa_verify calls in one launch.
 * results[i] receives 1/0.  Returns 1 on success, 0 if the device call failed (then results are untouched). */
SECP256K1_API int secp256k1_ecdsa_verify_batch(const secp256k1_context *ctx, const secp256k1_ecdsa_signature *sigs,
                                               const unsigned char *msgs32, const secp256k1_pubkey *pubkeys, size_t n,
                                               unsigned char *results);

#ifdef __cplusplus
}
#endif
#endif
