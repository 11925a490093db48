// blake2b.cuh — Blake2b compression (RFC 7693) for sm_100a.
//
// The reference reaches Blake2b through libsodium's crypto_generichash_blake2b_* (EXTERNAL,
// depends/packages/libsodium.mk:6-9; call sites src/crypto/equihash.cpp:44-59, src/pow.cpp:304-315).
// B200 has no 64-bit integer pipe: uint64_t adds become IADD3 + IADD3.X pairs and the rotations by
// 32/24/16 become byte permutes, the rotation by 63 a funnel shift; nvcc does that lowering from the
// plain C below.
#pragma once
#include "bcp_common.cuh"

BCP_CONST uint64_t BLAKE2B_IV[8] = {0x6a09e667f3bcc908ULL, 0xbb67ae8584caa73bULL, 0x3c6ef372fe94f82bULL,
                                    0xa54ff53a5f1d36f1ULL, 0x510e527fade682d1ULL, 0x9b05688c2b3e6c1fULL,
                                    0x1f83d9abfb41bd6bULL, 0x5be0cd19137e2179ULL};
// message schedule (RFC 7693 sigma, rounds 10 and 11 repeat rounds 0 and 1) is spelled out as literal
// indices below so that m[] stays in registers
BCP_DEV uint64_t rotr64(uint64_t x, int c) { return (x >> c) | (x << (64 - c)); }

#define BCP_B2G(a, b, c, d, x, y)        \
    a = a + b + (x);                     \
    d = rotr64(d ^ a, 32);               \
    c = c + d;                           \
    b = rotr64(b ^ c, 24);               \
    a = a + b + (y);                     \
    d = rotr64(d ^ a, 16);               \
    c = c + d;                           \
    b = rotr64(b ^ c, 63);

// h: chaining value (in/out); m: 16 little-endian message words; t0: byte counter (low word; the high
// word is zero for every message on this path); last: final-block flag
BCP_DEV void blake2b_compress(uint64_t *h, const uint64_t *m, uint64_t t0, bool last) {
    uint64_t v[16];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { v[i] = h[i]; v[i + 8] = BLAKE2B_IV[i]; }
    v[12] ^= t0;
    if (last) v[14] = ~v[14];
#define BCP_B2ROUND(s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, s10, s11, s12, s13, s14, s15) \
    BCP_B2G(v[0], v[4], v[8], v[12], m[s0], m[s1]);   \
    BCP_B2G(v[1], v[5], v[9], v[13], m[s2], m[s3]);   \
    BCP_B2G(v[2], v[6], v[10], v[14], m[s4], m[s5]);  \
    BCP_B2G(v[3], v[7], v[11], v[15], m[s6], m[s7]);  \
    BCP_B2G(v[0], v[5], v[10], v[15], m[s8], m[s9]);  \
    BCP_B2G(v[1], v[6], v[11], v[12], m[s10], m[s11]); \
    BCP_B2G(v[2], v[7], v[8], v[13], m[s12], m[s13]); \
    BCP_B2G(v[3], v[4], v[9], v[14], m[s14], m[s15])
    BCP_B2ROUND(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
    BCP_B2ROUND(14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3);
    BCP_B2ROUND(11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4);
    BCP_B2ROUND(7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8);
    BCP_B2ROUND(9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13);
    BCP_B2ROUND(2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9);
    BCP_B2ROUND(12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11);
    BCP_B2ROUND(13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10);
    BCP_B2ROUND(6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5);
    BCP_B2ROUND(10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0);
    BCP_B2ROUND(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
    BCP_B2ROUND(14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3);
#undef BCP_B2ROUND
    BCP_UNROLL
    for (int i = 0; i < 8; i++) h[i] ^= v[i] ^ v[i + 8];
}
// parameter block -> initial chaining value: digest length, fanout = depth = 1, 16-byte personalization
BCP_DEV void blake2b_init_personal(uint64_t *h, uint32_t outlen, uint64_t pers_lo, uint64_t pers_hi) {
    BCP_UNROLL
    for (int i = 0; i < 8; i++) h[i] = BLAKE2B_IV[i];
    h[0] ^= 0x01010000ULL ^ (uint64_t)outlen;
    h[6] ^= pers_lo;
    h[7] ^= pers_hi;
}
