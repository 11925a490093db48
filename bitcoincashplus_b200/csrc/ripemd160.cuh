// ripemd160.cuh — RIPEMD-160 compression for sm_100a (published algorithm of Dobbertin, Bosselaers, Preneel).
// The reference's is src/crypto/ripemd160.cpp; it is reached through CHash160 (src/hash.h:46-68) by OP_HASH160
// (src/script/interpreter.cpp:994-997).  Only what the P2PKH path needs: one 64-byte block over a 32-byte SHA-256
// digest.  Rounds are spelled out through macros so that the message words stay in registers.
#pragma once
#include "bcp_common.cuh"

BCP_DEV uint32_t rol32(uint32_t x, int c) {
#if BCP_PTX
    return __funnelshift_l(x, x, c);
#else
    return (x << c) | (x >> (32 - c));
#endif
}
#define RMD_F1(x, y, z) ((x) ^ (y) ^ (z))
#define RMD_F2(x, y, z) (((x) & (y)) | (~(x) & (z)))
#define RMD_F3(x, y, z) (((x) | ~(y)) ^ (z))
#define RMD_F4(x, y, z) (((x) & (z)) | ((y) & ~(z)))
#define RMD_F5(x, y, z) ((x) ^ ((y) | ~(z)))
#define RMD_STEP(f, a, b, c, d, e, w, k, s) \
    a = rol32(a + f(b, c, d) + (w) + (k), s) + e; \
    c = rol32(c, 10);

// s: 5-word chaining value (in/out); w: 16 little-endian message words
BCP_DEV void ripemd160_compress(uint32_t *s, const uint32_t *w) {
    uint32_t a1 = s[0], b1 = s[1], c1 = s[2], d1 = s[3], e1 = s[4];
    uint32_t a2 = a1, b2 = b1, c2 = c1, d2 = d1, e2 = e1;
#define L1(a, b, c, d, e, i, sh) RMD_STEP(RMD_F1, a, b, c, d, e, w[i], 0u, sh)
#define L2(a, b, c, d, e, i, sh) RMD_STEP(RMD_F2, a, b, c, d, e, w[i], 0x5A827999u, sh)
#define L3(a, b, c, d, e, i, sh) RMD_STEP(RMD_F3, a, b, c, d, e, w[i], 0x6ED9EBA1u, sh)
#define L4(a, b, c, d, e, i, sh) RMD_STEP(RMD_F4, a, b, c, d, e, w[i], 0x8F1BBCDCu, sh)
#define L5(a, b, c, d, e, i, sh) RMD_STEP(RMD_F5, a, b, c, d, e, w[i], 0xA953FD4Eu, sh)
#define R1(a, b, c, d, e, i, sh) RMD_STEP(RMD_F5, a, b, c, d, e, w[i], 0x50A28BE6u, sh)
#define R2(a, b, c, d, e, i, sh) RMD_STEP(RMD_F4, a, b, c, d, e, w[i], 0x5C4DD124u, sh)
#define R3(a, b, c, d, e, i, sh) RMD_STEP(RMD_F3, a, b, c, d, e, w[i], 0x6D703EF3u, sh)
#define R4(a, b, c, d, e, i, sh) RMD_STEP(RMD_F2, a, b, c, d, e, w[i], 0x7A6D76E9u, sh)
#define R5(a, b, c, d, e, i, sh) RMD_STEP(RMD_F1, a, b, c, d, e, w[i], 0u, sh)
    // left line
    L1(a1, b1, c1, d1, e1, 0, 11) L1(e1, a1, b1, c1, d1, 1, 14) L1(d1, e1, a1, b1, c1, 2, 15) L1(c1, d1, e1, a1, b1, 3, 12)
    L1(b1, c1, d1, e1, a1, 4, 5) L1(a1, b1, c1, d1, e1, 5, 8) L1(e1, a1, b1, c1, d1, 6, 7) L1(d1, e1, a1, b1, c1, 7, 9)
    L1(c1, d1, e1, a1, b1, 8, 11) L1(b1, c1, d1, e1, a1, 9, 13) L1(a1, b1, c1, d1, e1, 10, 14) L1(e1, a1, b1, c1, d1, 11, 15)
    L1(d1, e1, a1, b1, c1, 12, 6) L1(c1, d1, e1, a1, b1, 13, 7) L1(b1, c1, d1, e1, a1, 14, 9) L1(a1, b1, c1, d1, e1, 15, 8)
    L2(e1, a1, b1, c1, d1, 7, 7) L2(d1, e1, a1, b1, c1, 4, 6) L2(c1, d1, e1, a1, b1, 13, 8) L2(b1, c1, d1, e1, a1, 1, 13)
    L2(a1, b1, c1, d1, e1, 10, 11) L2(e1, a1, b1, c1, d1, 6, 9) L2(d1, e1, a1, b1, c1, 15, 7) L2(c1, d1, e1, a1, b1, 3, 15)
    L2(b1, c1, d1, e1, a1, 12, 7) L2(a1, b1, c1, d1, e1, 0, 12) L2(e1, a1, b1, c1, d1, 9, 15) L2(d1, e1, a1, b1, c1, 5, 9)
    L2(c1, d1, e1, a1, b1, 2, 11) L2(b1, c1, d1, e1, a1, 14, 7) L2(a1, b1, c1, d1, e1, 11, 13) L2(e1, a1, b1, c1, d1, 8, 12)
    L3(d1, e1, a1, b1, c1, 3, 11) L3(c1, d1, e1, a1, b1, 10, 13) L3(b1, c1, d1, e1, a1, 14, 6) L3(a1, b1, c1, d1, e1, 4, 7)
    L3(e1, a1, b1, c1, d1, 9, 14) L3(d1, e1, a1, b1, c1, 15, 9) L3(c1, d1, e1, a1, b1, 8, 13) L3(b1, c1, d1, e1, a1, 1, 15)
    L3(a1, b1, c1, d1, e1, 2, 14) L3(e1, a1, b1, c1, d1, 7, 8) L3(d1, e1, a1, b1, c1, 0, 13) L3(c1, d1, e1, a1, b1, 6, 6)
    L3(b1, c1, d1, e1, a1, 13, 5) L3(a1, b1, c1, d1, e1, 11, 12) L3(e1, a1, b1, c1, d1, 5, 7) L3(d1, e1, a1, b1, c1, 12, 5)
    L4(c1, d1, e1, a1, b1, 1, 11) L4(b1, c1, d1, e1, a1, 9, 12) L4(a1, b1, c1, d1, e1, 11, 14) L4(e1, a1, b1, c1, d1, 10, 15)
    L4(d1, e1, a1, b1, c1, 0, 14) L4(c1, d1, e1, a1, b1, 8, 15) L4(b1, c1, d1, e1, a1, 12, 9) L4(a1, b1, c1, d1, e1, 4, 8)
    L4(e1, a1, b1, c1, d1, 13, 9) L4(d1, e1, a1, b1, c1, 3, 14) L4(c1, d1, e1, a1, b1, 7, 5) L4(b1, c1, d1, e1, a1, 15, 6)
    L4(a1, b1, c1, d1, e1, 14, 8) L4(e1, a1, b1, c1, d1, 5, 6) L4(d1, e1, a1, b1, c1, 6, 5) L4(c1, d1, e1, a1, b1, 2, 12)
    L5(b1, c1, d1, e1, a1, 4, 9) L5(a1, b1, c1, d1, e1, 0, 15) L5(e1, a1, b1, c1, d1, 5, 5) L5(d1, e1, a1, b1, c1, 9, 11)
    L5(c1, d1, e1, a1, b1, 7, 6) L5(b1, c1, d1, e1, a1, 12, 8) L5(a1, b1, c1, d1, e1, 2, 13) L5(e1, a1, b1, c1, d1, 10, 12)
    L5(d1, e1, a1, b1, c1, 14, 5) L5(c1, d1, e1, a1, b1, 1, 12) L5(b1, c1, d1, e1, a1, 3, 13) L5(a1, b1, c1, d1, e1, 8, 14)
    L5(e1, a1, b1, c1, d1, 11, 11) L5(d1, e1, a1, b1, c1, 6, 8) L5(c1, d1, e1, a1, b1, 15, 5) L5(b1, c1, d1, e1, a1, 13, 6)
    // right line
    R1(a2, b2, c2, d2, e2, 5, 8) R1(e2, a2, b2, c2, d2, 14, 9) R1(d2, e2, a2, b2, c2, 7, 9) R1(c2, d2, e2, a2, b2, 0, 11)
    R1(b2, c2, d2, e2, a2, 9, 13) R1(a2, b2, c2, d2, e2, 2, 15) R1(e2, a2, b2, c2, d2, 11, 15) R1(d2, e2, a2, b2, c2, 4, 5)
    R1(c2, d2, e2, a2, b2, 13, 7) R1(b2, c2, d2, e2, a2, 6, 7) R1(a2, b2, c2, d2, e2, 15, 8) R1(e2, a2, b2, c2, d2, 8, 11)
    R1(d2, e2, a2, b2, c2, 1, 14) R1(c2, d2, e2, a2, b2, 10, 14) R1(b2, c2, d2, e2, a2, 3, 12) R1(a2, b2, c2, d2, e2, 12, 6)
    R2(e2, a2, b2, c2, d2, 6, 9) R2(d2, e2, a2, b2, c2, 11, 13) R2(c2, d2, e2, a2, b2, 3, 15) R2(b2, c2, d2, e2, a2, 7, 7)
    R2(a2, b2, c2, d2, e2, 0, 12) R2(e2, a2, b2, c2, d2, 13, 8) R2(d2, e2, a2, b2, c2, 5, 9) R2(c2, d2, e2, a2, b2, 10, 11)
    R2(b2, c2, d2, e2, a2, 14, 7) R2(a2, b2, c2, d2, e2, 15, 7) R2(e2, a2, b2, c2, d2, 8, 12) R2(d2, e2, a2, b2, c2, 12, 7)
    R2(c2, d2, e2, a2, b2, 4, 6) R2(b2, c2, d2, e2, a2, 9, 15) R2(a2, b2, c2, d2, e2, 1, 13) R2(e2, a2, b2, c2, d2, 2, 11)
    R3(d2, e2, a2, b2, c2, 15, 9) R3(c2, d2, e2, a2, b2, 5, 7) R3(b2, c2, d2, e2, a2, 1, 15) R3(a2, b2, c2, d2, e2, 3, 11)
    R3(e2, a2, b2, c2, d2, 7, 8) R3(d2, e2, a2, b2, c2, 14, 6) R3(c2, d2, e2, a2, b2, 6, 6) R3(b2, c2, d2, e2, a2, 9, 14)
    R3(a2, b2, c2, d2, e2, 11, 12) R3(e2, a2, b2, c2, d2, 8, 13) R3(d2, e2, a2, b2, c2, 12, 5) R3(c2, d2, e2, a2, b2, 2, 14)
    R3(b2, c2, d2, e2, a2, 10, 13) R3(a2, b2, c2, d2, e2, 0, 13) R3(e2, a2, b2, c2, d2, 4, 7) R3(d2, e2, a2, b2, c2, 13, 5)
    R4(c2, d2, e2, a2, b2, 8, 15) R4(b2, c2, d2, e2, a2, 6, 5) R4(a2, b2, c2, d2, e2, 4, 8) R4(e2, a2, b2, c2, d2, 1, 11)
    R4(d2, e2, a2, b2, c2, 3, 14) R4(c2, d2, e2, a2, b2, 11, 14) R4(b2, c2, d2, e2, a2, 15, 6) R4(a2, b2, c2, d2, e2, 0, 14)
    R4(e2, a2, b2, c2, d2, 5, 6) R4(d2, e2, a2, b2, c2, 12, 9) R4(c2, d2, e2, a2, b2, 2, 12) R4(b2, c2, d2, e2, a2, 13, 9)
    R4(a2, b2, c2, d2, e2, 9, 12) R4(e2, a2, b2, c2, d2, 7, 5) R4(d2, e2, a2, b2, c2, 10, 15) R4(c2, d2, e2, a2, b2, 14, 8)
    R5(b2, c2, d2, e2, a2, 12, 8) R5(a2, b2, c2, d2, e2, 15, 5) R5(e2, a2, b2, c2, d2, 10, 12) R5(d2, e2, a2, b2, c2, 4, 9)
    R5(c2, d2, e2, a2, b2, 1, 12) R5(b2, c2, d2, e2, a2, 5, 5) R5(a2, b2, c2, d2, e2, 8, 14) R5(e2, a2, b2, c2, d2, 7, 6)
    R5(d2, e2, a2, b2, c2, 6, 8) R5(c2, d2, e2, a2, b2, 2, 13) R5(b2, c2, d2, e2, a2, 13, 6) R5(a2, b2, c2, d2, e2, 14, 5)
    R5(e2, a2, b2, c2, d2, 0, 15) R5(d2, e2, a2, b2, c2, 3, 13) R5(c2, d2, e2, a2, b2, 9, 11) R5(b2, c2, d2, e2, a2, 11, 11)
#undef L1
#undef L2
#undef L3
#undef L4
#undef L5
#undef R1
#undef R2
#undef R3
#undef R4
#undef R5
    uint32_t t = s[1] + c1 + d2;
    s[1] = s[2] + d1 + e2;
    s[2] = s[3] + e1 + a2;
    s[3] = s[4] + a1 + b2;
    s[4] = s[0] + b1 + c2;
    s[0] = t;
}
// RIPEMD160 of a 32-byte SHA-256 digest given as 8 big-endian words (the state of sha256_compress): out = 5
// little-endian words (memory order of the 20 digest bytes)
BCP_DEV void ripemd160_of_sha256(uint32_t *out5, const uint32_t *sha_be8) {
    uint32_t w[16];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) {
        uint32_t v = sha_be8[i];
        w[i] = (v >> 24) | ((v >> 8) & 0xFF00u) | ((v << 8) & 0xFF0000u) | (v << 24);
    }
    w[8] = 0x80u;
    BCP_UNROLL
    for (int i = 9; i < 14; i++) w[i] = 0;
    w[14] = 256;
    w[15] = 0;
    out5[0] = 0x67452301u; out5[1] = 0xEFCDAB89u; out5[2] = 0x98BADCFEu; out5[3] = 0x10325476u; out5[4] = 0xC3D2E1F0u;
    ripemd160_compress(out5, w);
}
