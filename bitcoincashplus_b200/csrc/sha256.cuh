// sha256.cuh — SHA-256 compression and SHA-256d helpers for sm_100a.
//
// Functional counterpart of the reference's scalar hasher: sha256::Transform
// (src/crypto/sha256.cpp:57-187), CSHA256::Write/Finalize (:198-237) and CHash256 (src/hash.h:21-43).
// Written from FIPS 180-4; the message schedule is kept as a rolling 16-word window in registers.
#pragma once
#include "bcp_common.cuh"

BCP_CONST uint32_t SHA256_K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98,
    0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786,
    0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8,
    0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
    0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819,
    0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a,
    0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7,
    0xc67178f2};

BCP_DEV uint32_t rotr32(uint32_t x, int c) {
#if BCP_PTX
    return __funnelshift_r(x, x, c);
#else
    return (x >> c) | (x << (32 - c));
#endif
}
BCP_DEV void sha256_init(uint32_t *s) {
    s[0] = 0x6a09e667u; s[1] = 0xbb67ae85u; s[2] = 0x3c6ef372u; s[3] = 0xa54ff53au;
    s[4] = 0x510e527fu; s[5] = 0x9b05688cu; s[6] = 0x1f83d9abu; s[7] = 0x5be0cd19u;
}
// one 64-byte block; w[16] = big-endian message words (clobbered)
BCP_DEV void sha256_compress(uint32_t *s, uint32_t *w) {
    uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4], f = s[5], g = s[6], h = s[7];
    BCP_UNROLL
    for (int i = 0; i < 64; i++) {
        if (i >= 16) {
            uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
            uint32_t s0 = rotr32(w15, 7) ^ rotr32(w15, 18) ^ (w15 >> 3);
            uint32_t s1 = rotr32(w2, 17) ^ rotr32(w2, 19) ^ (w2 >> 10);
            w[i & 15] = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
        }
        uint32_t S1 = rotr32(e, 6) ^ rotr32(e, 11) ^ rotr32(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = h + S1 + ch + SHA256_K[i] + w[i & 15];
        uint32_t S0 = rotr32(a, 2) ^ rotr32(a, 13) ^ rotr32(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e; s[5] += f; s[6] += g; s[7] += h;
}
// second pass of SHA-256d: hash the 32-byte digest held in s[8] (as words); result left in s
BCP_DEV void sha256_second_pass(uint32_t *s) {
    uint32_t w[16];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) w[i] = s[i];
    w[8] = 0x80000000u;
    BCP_UNROLL
    for (int i = 9; i < 15; i++) w[i] = 0;
    w[15] = 256;
    sha256_init(s);
    sha256_compress(s, w);
}
// SHA256d of a 64-byte message given as 16 big-endian words (a Merkle inner node: L || R)
BCP_DEV void sha256d_64(uint32_t *s /*out 8*/, uint32_t *w /*16, clobbered*/) {
    sha256_init(s);
    sha256_compress(s, w);
    uint32_t p[16];
    p[0] = 0x80000000u;
    BCP_UNROLL
    for (int i = 1; i < 15; i++) p[i] = 0;
    p[15] = 512;
    sha256_compress(s, p);
    sha256_second_pass(s);
}
// big-endian word j (bytes 4j..4j+3) of a message starting at an arbitrary byte address, with
// zero / 0x80 / length padding applied past `len`.  Generic (byte loads); the txid kernel has a
// faster aligned-word path for full blocks.
BCP_DEV uint32_t sha256_msg_word(const uint8_t *data, size_t len, size_t total_padded, size_t pos) {
    uint32_t w = 0;
    BCP_UNROLL
    for (int b = 0; b < 4; b++) {
        size_t q = pos + b;
        uint32_t byte;
        if (q < len) byte = data[q];
        else if (q == len) byte = 0x80;
        else if (q >= total_padded - 8) byte = (uint32_t)(((uint64_t)len * 8) >> (8 * (total_padded - 1 - q))) & 0xff;
        else byte = 0;
        w = (w << 8) | byte;
    }
    return w;
}
BCP_DEV void dev_sha256_words(uint32_t *s, const uint8_t *data, size_t len) {
    size_t total = ((len + 9 + 63) / 64) * 64;
    sha256_init(s);
    for (size_t off = 0; off < total; off += 64) {
        uint32_t w[16];
        BCP_UNROLL
        for (int j = 0; j < 16; j++) w[j] = sha256_msg_word(data, len, total, off + 4 * j);
        sha256_compress(s, w);
    }
}
BCP_DEV void sha256_store_be(uint8_t *out, const uint32_t *s) {
    BCP_UNROLL
    for (int i = 0; i < 8; i++) {
        out[4 * i] = (uint8_t)(s[i] >> 24); out[4 * i + 1] = (uint8_t)(s[i] >> 16);
        out[4 * i + 2] = (uint8_t)(s[i] >> 8); out[4 * i + 3] = (uint8_t)s[i];
    }
}
BCP_DEV void dev_sha256_bytes(uint8_t *out32, const uint8_t *data, size_t len) {
    uint32_t s[8];
    dev_sha256_words(s, data, len);
    sha256_store_be(out32, s);
}
BCP_DEV void dev_sha256d_bytes(uint8_t *out32, const uint8_t *data, size_t len) {
    uint32_t s[8];
    dev_sha256_words(s, data, len);
    sha256_second_pass(s);
    sha256_store_be(out32, s);
}
