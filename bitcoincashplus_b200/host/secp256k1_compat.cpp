// secp256k1_compat.cpp — host side of level B3: the libsecp256k1 verify-side C API (include/compat/secp256k1.h)
// implemented over the C ABI of libbcp_b200.so.  Byte-level bookkeeping (range checks against n, low-S
// normalisation, serialisation) happens here; every curve operation — key decompression / on-curve test and
// the verification itself — is a call into the CUDA library.  No CPU fallback: if the device call fails the
// function reports failure exactly like an invalid input (return 0) after setting bcp_gpu_last_error().
//
// Reference behaviour mirrored: src/secp256k1/src/secp256k1.c (ARG_CHECK :21-26, illegal callback :28-37,
// context :59-119, pubkey parse/serialize :150-186, signature parse_compact/serialize/normalize :228-291,
// verify :293-308).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/bcp_gpu.h"
#include "../../include/compat/secp256k1.h"
#include "../../include/compat/secp256k1_recovery.h"

struct secp256k1_context_struct {
    int device;
    void (*illegal)(const char *, void *);
    const void *illegal_data;
};

static void default_illegal(const char *str, void *) {
    fprintf(stderr, "[libsecp256k1] illegal argument: %s\n", str);
    abort();
}
#define ARG_CHECK(cond)                                                                \
    do {                                                                               \
        if (!(cond)) {                                                                 \
            if (ctx) ctx->illegal(#cond, const_cast<void *>(ctx->illegal_data));      \
            return 0;                                                                  \
        }                                                                              \
    } while (0)

static const unsigned char N_BE[32] = {0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF,
                                       0xFF, 0xFF, 0xFF, 0xFF, 0xFE, 0xBA, 0xAE, 0xDC, 0xE6, 0xAF, 0x48,
                                       0xA0, 0x3B, 0xBF, 0xD2, 0x5E, 0x8C, 0xD0, 0x36, 0x41, 0x41};
static const unsigned char NHALF_BE[32] = {0x7F, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF,
                                           0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0x5D, 0x57, 0x6E, 0x73, 0x57, 0xA4,
                                           0x50, 0x1D, 0xDF, 0xE9, 0x2F, 0x46, 0x68, 0x1B, 0x20, 0xA0};

extern "C" {

secp256k1_context *secp256k1_context_create(unsigned int flags) {
    (void)flags;
    const char *e = getenv("BCP_DEVICE");
    int dev = e ? atoi(e) : 0;
    if (bcp_gpu_init(dev) != BCP_OK) return nullptr;
    secp256k1_context *c = new secp256k1_context_struct{dev, default_illegal, nullptr};
    return c;
}
void secp256k1_context_destroy(secp256k1_context *ctx) { delete ctx; }
void secp256k1_context_set_illegal_callback(secp256k1_context *ctx, void (*fun)(const char *, void *), const void *data) {
    if (!ctx) return;
    ctx->illegal = fun ? fun : default_illegal;
    ctx->illegal_data = data;
}

int secp256k1_ec_pubkey_parse(const secp256k1_context *ctx, secp256k1_pubkey *pubkey, const unsigned char *input,
                              size_t inputlen) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(pubkey != nullptr);
    memset(pubkey, 0, sizeof(*pubkey));
    ARG_CHECK(input != nullptr);
    // length/prefix agreement as secp256k1_eckey_pubkey_parse (eckey_impl.h:17-35)
    bool shape = (inputlen == 33 && (input[0] == 2 || input[0] == 3)) ||
                 (inputlen == 65 && (input[0] == 4 || input[0] == 6 || input[0] == 7));
    if (!shape) return 0;
    unsigned char slot[65] = {0}, xy[64], ok = 0;
    memcpy(slot, input, inputlen);
    if (bcp_gpu_pubkey_parse_batch(ctx->device, slot, 1, xy, &ok) != BCP_OK || !ok) return 0;
    memcpy(pubkey->data, xy, 64);
    return 1;
}

int secp256k1_ec_pubkey_serialize(const secp256k1_context *ctx, unsigned char *output, size_t *outputlen,
                                  const secp256k1_pubkey *pubkey, unsigned int flags) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(outputlen != nullptr);
    ARG_CHECK(*outputlen >= ((flags & SECP256K1_FLAGS_BIT_COMPRESSION) ? 33u : 65u));
    size_t len = *outputlen;
    *outputlen = 0;
    ARG_CHECK(output != nullptr);
    memset(output, 0, len);
    ARG_CHECK(pubkey != nullptr);
    ARG_CHECK((flags & 0xFF) == SECP256K1_FLAGS_TYPE_COMPRESSION);
    bool zero = true;
    for (int i = 0; i < 32; i++) zero = zero && pubkey->data[i] == 0;
    ARG_CHECK(!zero); /* secp256k1_pubkey_load: x == 0 is an illegal argument (secp256k1.c:132) */
    if (flags & SECP256K1_FLAGS_BIT_COMPRESSION) {
        output[0] = 0x02 | (pubkey->data[63] & 1);
        memcpy(output + 1, pubkey->data, 32);
        *outputlen = 33;
    } else {
        output[0] = 0x04;
        memcpy(output + 1, pubkey->data, 64);
        *outputlen = 65;
    }
    return 1;
}

int secp256k1_ecdsa_signature_parse_compact(const secp256k1_context *ctx, secp256k1_ecdsa_signature *sig,
                                            const unsigned char *input64) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(sig != nullptr);
    ARG_CHECK(input64 != nullptr);
    bool overflow = memcmp(input64, N_BE, 32) >= 0 || memcmp(input64 + 32, N_BE, 32) >= 0;
    if (overflow) {
        memset(sig, 0, sizeof(*sig));
        return 0;
    }
    memcpy(sig->data, input64, 64);
    return 1;
}

int secp256k1_ecdsa_signature_serialize_compact(const secp256k1_context *ctx, unsigned char *output64,
                                                const secp256k1_ecdsa_signature *sig) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(output64 != nullptr);
    ARG_CHECK(sig != nullptr);
    memcpy(output64, sig->data, 64);
    return 1;
}

int secp256k1_ecdsa_signature_normalize(const secp256k1_context *ctx, secp256k1_ecdsa_signature *sigout,
                                        const secp256k1_ecdsa_signature *sigin) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(sigin != nullptr);
    const unsigned char *s = sigin->data + 32;
    int high = memcmp(s, NHALF_BE, 32) > 0;
    if (sigout) {
        secp256k1_ecdsa_signature tmp = *sigin;
        if (high) { /* s := n - s, byte-wise big-endian subtraction */
            int borrow = 0;
            for (int i = 31; i >= 0; i--) {
                int d = (int)N_BE[i] - (int)s[i] - borrow;
                borrow = d < 0;
                tmp.data[32 + i] = (unsigned char)(d + (borrow ? 256 : 0));
            }
        }
        *sigout = tmp;
    }
    return high;
}

int secp256k1_ecdsa_verify_batch(const secp256k1_context *ctx, const secp256k1_ecdsa_signature *sigs,
                                 const unsigned char *msgs32, const secp256k1_pubkey *pubkeys, size_t n,
                                 unsigned char *results) {
    ARG_CHECK(ctx != nullptr);
    if (n == 0) return 1;
    ARG_CHECK(sigs != nullptr);
    ARG_CHECK(msgs32 != nullptr);
    ARG_CHECK(pubkeys != nullptr);
    ARG_CHECK(results != nullptr);
    std::vector<unsigned char> pub(65 * n), ok(n);
    for (size_t i = 0; i < n; i++) {
        pub[65 * i] = 0x04; /* parsed keys are affine points: hand them over uncompressed */
        memcpy(&pub[65 * i + 1], pubkeys[i].data, 64);
    }
    static_assert(sizeof(secp256k1_ecdsa_signature) == 64, "signature blob is r||s");
    if (bcp_gpu_ecdsa_verify_batch(ctx->device, msgs32, pub.data(), sigs[0].data, n, 0, ok.data()) != BCP_OK) return 0;
    memcpy(results, ok.data(), n);
    return 1;
}

int secp256k1_ec_pubkey_tweak_add(const secp256k1_context *ctx, secp256k1_pubkey *pubkey, const unsigned char *tweak) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(pubkey != nullptr);
    ARG_CHECK(tweak != nullptr);
    bool zero = true;
    for (int i = 0; i < 32; i++) zero = zero && pubkey->data[i] == 0;
    unsigned char slot[65], xy[64], ok = 0;
    slot[0] = 0x04;
    memcpy(slot + 1, pubkey->data, 64);
    memset(pubkey, 0, sizeof(*pubkey)); /* secp256k1.c:489: cleared before the outcome is known */
    ARG_CHECK(!zero);                   /* secp256k1_pubkey_load (secp256k1.c:132) */
    if (bcp_gpu_pubkey_tweak_add_batch(ctx->device, slot, tweak, 1, xy, &ok) != BCP_OK || !ok) return 0;
    memcpy(pubkey->data, xy, 64);
    return 1;
}

int secp256k1_ecdsa_recoverable_signature_parse_compact(const secp256k1_context *ctx, secp256k1_ecdsa_recoverable_signature *sig,
                                                        const unsigned char *input64, int recid) {
    ARG_CHECK(sig != nullptr);
    ARG_CHECK(input64 != nullptr);
    ARG_CHECK(recid >= 0 && recid <= 3);
    const bool overflow = memcmp(input64, N_BE, 32) >= 0 || memcmp(input64 + 32, N_BE, 32) >= 0;
    if (overflow) {
        memset(sig, 0, sizeof(*sig));
        return 0;
    }
    memcpy(sig->data, input64, 64);
    sig->data[64] = (unsigned char)recid;
    return 1;
}

int secp256k1_ecdsa_recoverable_signature_convert(const secp256k1_context *ctx, secp256k1_ecdsa_signature *sig,
                                                  const secp256k1_ecdsa_recoverable_signature *sigin) {
    ARG_CHECK(sig != nullptr);
    ARG_CHECK(sigin != nullptr);
    memcpy(sig->data, sigin->data, 64);
    return 1;
}

int secp256k1_ecdsa_recoverable_signature_serialize_compact(const secp256k1_context *ctx, unsigned char *output64, int *recid,
                                                            const secp256k1_ecdsa_recoverable_signature *sig) {
    ARG_CHECK(output64 != nullptr);
    ARG_CHECK(sig != nullptr);
    ARG_CHECK(recid != nullptr);
    memcpy(output64, sig->data, 64);
    *recid = sig->data[64];
    return 1;
}

int secp256k1_ecdsa_recover(const secp256k1_context *ctx, secp256k1_pubkey *pubkey, const secp256k1_ecdsa_recoverable_signature *sig,
                            const unsigned char *msg32) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(msg32 != nullptr);
    ARG_CHECK(sig != nullptr);
    ARG_CHECK(pubkey != nullptr);
    unsigned char xy[64], ok = 0;
    const unsigned char rid = sig->data[64];
    if (bcp_gpu_ecdsa_recover_batch(ctx->device, msg32, sig->data, &rid, 1, xy, &ok) != BCP_OK || !ok) {
        memset(pubkey, 0, sizeof(*pubkey));
        return 0;
    }
    memcpy(pubkey->data, xy, 64);
    return 1;
}

int secp256k1_ecdsa_verify(const secp256k1_context *ctx, const secp256k1_ecdsa_signature *sig, const unsigned char *msg32,
                           const secp256k1_pubkey *pubkey) {
    ARG_CHECK(ctx != nullptr);
    ARG_CHECK(msg32 != nullptr);
    ARG_CHECK(sig != nullptr);
    ARG_CHECK(pubkey != nullptr);
    bool zero = true;
    for (int i = 0; i < 32; i++) zero = zero && pubkey->data[i] == 0;
    ARG_CHECK(!zero); /* secp256k1_pubkey_load (secp256k1.c:132) */
    unsigned char r = 0;
    if (!secp256k1_ecdsa_verify_batch(ctx, sig, msg32, pubkey, 1, &r)) return 0;
    return r;
}

} // extern "C"
