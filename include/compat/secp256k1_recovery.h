/* include/compat/secp256k1_recovery.h — the recovery-module calls of level B3: the subset of
 * src/secp256k1/include/secp256k1_recovery.h that src/pubkey.cpp:8,195-218 uses (CPubKey::RecoverCompact), same names,
 * argument meaning and return conventions, over libbcp_b200.so.  Signing (secp256k1_ecdsa_sign_recoverable) is the
 * wallet's side and stays with the stock library.
 */
#ifndef BCP_COMPAT_SECP256K1_RECOVERY_H
#define BCP_COMPAT_SECP256K1_RECOVERY_H
#include "secp256k1.h"
#ifdef __cplusplus
extern "C" {
#endif

/* 65 bytes, freely copyable, identical signatures compare equal with memcmp (secp256k1_recovery.h:12-27);
 * here: r || s big-endian || recovery id */
typedef struct { unsigned char data[65]; } secp256k1_ecdsa_recoverable_signature;

/* modules/recovery/main_impl.h:38-58 (declared :36): returns 0 and zeroes the signature when r or s >= n;
 * recid outside 0..3 is an illegal argument */
SECP256K1_API int secp256k1_ecdsa_recoverable_signature_parse_compact(const secp256k1_context *ctx,
                                                                      secp256k1_ecdsa_recoverable_signature *sig,
                                                                      const unsigned char *input64, int recid);
/* main_impl.h:73-85 */
SECP256K1_API int secp256k1_ecdsa_recoverable_signature_convert(const secp256k1_context *ctx, secp256k1_ecdsa_signature *sig,
                                                                const secp256k1_ecdsa_recoverable_signature *sigin);
/* main_impl.h:60-71 */
SECP256K1_API int secp256k1_ecdsa_recoverable_signature_serialize_compact(const secp256k1_context *ctx, unsigned char *output64,
                                                                          int *recid,
                                                                          const secp256k1_ecdsa_recoverable_signature *sig);
/* main_impl.h:170-190 (declared :99): 1 and the key, or 0 and a zeroed key */
SECP256K1_API int secp256k1_ecdsa_recover(const secp256k1_context *ctx, secp256k1_pubkey *pubkey,
                                          const secp256k1_ecdsa_recoverable_signature *sig, const unsigned char *msg32);

#ifdef __cplusplus
}
#endif
#endif
