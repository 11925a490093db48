/* oracle/bcp_oracle.c — CPU ORACLE, TEST INFRASTRUCTURE ONLY (see bcp_oracle.h).
 *
 * Plain C11 restatement of the reference's CPU algorithms for the hot path.  Written from the
 * published algorithms (FIPS 180-4, RFC 7693, SEC1/SEC2, RFC 6979, the Equihash paper) and the
 * decision chains of the reference functions cited at each routine (paths relative to
 * /root/reference).  Verdicts are mathematical functions of the input bytes; where the reference
 * uses a particular evaluation strategy (wNAF Strauss ecmult, 5x52 field limbs) the oracle uses
 * the simplest strategy with the same result (fixed 4-bit windows, 4x64 limbs) and the tests pin
 * equality of RESULTS against the reference's vectors and against oracle/_ref.
 */
#include "bcp_oracle.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 u128;

/* ============================================================== SHA-256 (FIPS 180-4) ======= */
/* follows src/crypto/sha256.cpp:57-187 (Transform), :198-237 (Write/Finalize) */
static const uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98,
    0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786,
    0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8,
    0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
    0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819,
    0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a,
    0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7,
    0xc67178f2};
typedef struct {
    uint32_t s[8];
    uint8_t buf[64];
    uint64_t bytes;
} sha256_ctx;
static inline uint32_t ror32(uint32_t x, int c) { return (x >> c) | (x << (32 - c)); }
static void sha256_init(sha256_ctx *c) {
    static const uint32_t iv[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a,
                                   0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    memcpy(c->s, iv, 32);
    c->bytes = 0;
}
static void sha256_compress(uint32_t s[8], const uint8_t *blk) {
    uint32_t w[64], a, b, c, d, e, f, g, h;
    for (int i = 0; i < 16; i++)
        w[i] = ((uint32_t)blk[4 * i] << 24) | ((uint32_t)blk[4 * i + 1] << 16) | ((uint32_t)blk[4 * i + 2] << 8) |
               blk[4 * i + 3];
    for (int i = 16; i < 64; i++) {
        uint32_t s0 = ror32(w[i - 15], 7) ^ ror32(w[i - 15], 18) ^ (w[i - 15] >> 3);
        uint32_t s1 = ror32(w[i - 2], 17) ^ ror32(w[i - 2], 19) ^ (w[i - 2] >> 10);
        w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    a = s[0]; b = s[1]; c = s[2]; d = s[3]; e = s[4]; f = s[5]; g = s[6]; h = s[7];
    for (int i = 0; i < 64; i++) {
        uint32_t S1 = ror32(e, 6) ^ ror32(e, 11) ^ ror32(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = h + S1 + ch + K256[i] + w[i];
        uint32_t S0 = ror32(a, 2) ^ ror32(a, 13) ^ ror32(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e; s[5] += f; s[6] += g; s[7] += h;
}
static void sha256_update(sha256_ctx *c, const uint8_t *d, size_t len) {
    size_t fill = c->bytes % 64;
    c->bytes += len;
    if (fill) {
        size_t take = 64 - fill < len ? 64 - fill : len;
        memcpy(c->buf + fill, d, take);
        d += take; len -= take;
        if (fill + take < 64) return;
        sha256_compress(c->s, c->buf);
    }
    while (len >= 64) { sha256_compress(c->s, d); d += 64; len -= 64; }
    if (len) memcpy(c->buf, d, len);
}
static void sha256_final(sha256_ctx *c, uint8_t out[32]) {
    uint64_t bits = c->bytes * 8;
    uint8_t pad[72] = {0x80};
    size_t fill = c->bytes % 64;
    size_t padlen = (fill < 56) ? 56 - fill : 120 - fill;
    uint8_t lenb[8];
    for (int i = 0; i < 8; i++) lenb[i] = (uint8_t)(bits >> (56 - 8 * i));
    sha256_update(c, pad, padlen);
    sha256_update(c, lenb, 8);
    for (int i = 0; i < 8; i++) {
        out[4 * i] = c->s[i] >> 24; out[4 * i + 1] = c->s[i] >> 16; out[4 * i + 2] = c->s[i] >> 8; out[4 * i + 3] = c->s[i];
    }
}
void orc_sha256(const uint8_t *data, size_t len, uint8_t out[32]) {
    sha256_ctx c;
    sha256_init(&c); sha256_update(&c, data, len); sha256_final(&c, out);
}
/* CHash256 (src/hash.h:21-43): SHA256(SHA256(x)) */
void orc_sha256d(const uint8_t *data, size_t len, uint8_t out[32]) {
    uint8_t t[32];
    orc_sha256(data, len, t);
    orc_sha256(t, 32, out);
}

/* ============================================================== RIPEMD-160 ================== */
/* published algorithm (Dobbertin/Bosselaers/Preneel); the reference's is src/crypto/ripemd160.cpp */
static inline uint32_t rol32(uint32_t x, int c) { return (x << c) | (x >> (32 - c)); }
static void rmd_compress(uint32_t s[5], const uint8_t *blk) {
    static const uint8_t rl[80] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 7, 4, 13, 1, 10, 6, 15, 3, 12, 0, 9,
                                   5, 2, 14, 11, 8, 3, 10, 14, 4, 9, 15, 8, 1, 2, 7, 0, 6, 13, 11, 5, 12, 1, 9, 11, 10, 0, 8,
                                   12, 4, 13, 3, 7, 15, 14, 5, 6, 2, 4, 0, 5, 9, 7, 12, 2, 10, 14, 1, 3, 8, 11, 6, 15, 13};
    static const uint8_t rr[80] = {5, 14, 7, 0, 9, 2, 11, 4, 13, 6, 15, 8, 1, 10, 3, 12, 6, 11, 3, 7, 0, 13, 5, 10, 14, 15, 8,
                                   12, 4, 9, 1, 2, 15, 5, 1, 3, 7, 14, 6, 9, 11, 8, 12, 2, 10, 0, 4, 13, 8, 6, 4, 1, 3, 11,
                                   15, 0, 5, 12, 2, 13, 9, 7, 10, 14, 12, 15, 10, 4, 1, 5, 8, 7, 6, 2, 13, 14, 0, 3, 9, 11};
    static const uint8_t sl[80] = {11, 14, 15, 12, 5, 8, 7, 9, 11, 13, 14, 15, 6, 7, 9, 8, 7, 6, 8, 13, 11, 9, 7, 15, 7, 12, 15,
                                   9, 11, 7, 13, 12, 11, 13, 6, 7, 14, 9, 13, 15, 14, 8, 13, 6, 5, 12, 7, 5, 11, 12, 14, 15, 14,
                                   15, 9, 8, 9, 14, 5, 6, 8, 6, 5, 12, 9, 15, 5, 11, 6, 8, 13, 12, 5, 12, 13, 14, 11, 8, 5, 6};
    static const uint8_t sr[80] = {8, 9, 9, 11, 13, 15, 15, 5, 7, 7, 8, 11, 14, 14, 12, 6, 9, 13, 15, 7, 12, 8, 9, 11, 7, 7, 12,
                                   7, 6, 15, 13, 11, 9, 7, 15, 11, 8, 6, 6, 14, 12, 13, 5, 14, 13, 13, 7, 5, 15, 5, 8, 11, 14,
                                   14, 6, 14, 6, 9, 12, 9, 12, 5, 15, 8, 8, 5, 12, 9, 12, 5, 14, 6, 8, 13, 6, 5, 15, 13, 11, 11};
    static const uint32_t kl[5] = {0, 0x5A827999, 0x6ED9EBA1, 0x8F1BBCDC, 0xA953FD4E};
    static const uint32_t kr[5] = {0x50A28BE6, 0x5C4DD124, 0x6D703EF3, 0x7A6D76E9, 0};
    uint32_t w[16];
    for (int i = 0; i < 16; i++)
        w[i] = blk[4 * i] | ((uint32_t)blk[4 * i + 1] << 8) | ((uint32_t)blk[4 * i + 2] << 16) | ((uint32_t)blk[4 * i + 3] << 24);
    uint32_t al = s[0], bl = s[1], cl = s[2], dl = s[3], el = s[4];
    uint32_t ar = al, br = bl, cr = cl, dr = dl, er = el;
    for (int j = 0; j < 80; j++) {
        int rnd = j / 16;
        uint32_t f, g;
        switch (rnd) {
            case 0: f = bl ^ cl ^ dl; g = br ^ (cr | ~dr); break;
            case 1: f = (bl & cl) | (~bl & dl); g = (br & dr) | (cr & ~dr); break;
            case 2: f = (bl | ~cl) ^ dl; g = (br | ~cr) ^ dr; break;
            case 3: f = (bl & dl) | (cl & ~dl); g = (br & cr) | (~br & dr); break;
            default: f = bl ^ (cl | ~dl); g = br ^ cr ^ dr; break;
        }
        uint32_t t = rol32(al + f + w[rl[j]] + kl[rnd], sl[j]) + el;
        al = el; el = dl; dl = rol32(cl, 10); cl = bl; bl = t;
        t = rol32(ar + g + w[rr[j]] + kr[rnd], sr[j]) + er;
        ar = er; er = dr; dr = rol32(cr, 10); cr = br; br = t;
    }
    uint32_t t = s[1] + cl + dr;
    s[1] = s[2] + dl + er; s[2] = s[3] + el + ar; s[3] = s[4] + al + br; s[4] = s[0] + bl + cr; s[0] = t;
}
void orc_ripemd160(const uint8_t *data, size_t len, uint8_t out[20]) {
    uint32_t s[5] = {0x67452301, 0xEFCDAB89, 0x98BADCFE, 0x10325476, 0xC3D2E1F0};
    size_t full = len / 64;
    for (size_t i = 0; i < full; i++) rmd_compress(s, data + 64 * i);
    uint8_t buf[128] = {0};
    size_t rem = len % 64;
    memcpy(buf, data + 64 * full, rem);
    buf[rem] = 0x80;
    size_t tot = (rem < 56) ? 64 : 128;
    uint64_t bits = (uint64_t)len * 8;
    for (int i = 0; i < 8; i++) buf[tot - 8 + i] = (uint8_t)(bits >> (8 * i));
    rmd_compress(s, buf);
    if (tot == 128) rmd_compress(s, buf + 64);
    for (int i = 0; i < 5; i++) {
        out[4 * i] = s[i]; out[4 * i + 1] = s[i] >> 8; out[4 * i + 2] = s[i] >> 16; out[4 * i + 3] = s[i] >> 24;
    }
}
/* CHash160 (src/hash.h:46-68) */
void orc_hash160(const uint8_t *data, size_t len, uint8_t out[20]) {
    uint8_t t[32];
    orc_sha256(data, len, t);
    orc_ripemd160(t, 32, out);
}

/* ============================================================== Merkle ====================== */
/* ComputeMerkleRoot (src/consensus/merkle.cpp:47-150), restated level by level.
 * Level l holds ceil(n / 2^l) nodes.  A sibling pair (2j, 2j+1) is hashed; an odd tail is
 * paired with itself.  `mutated` is raised for a pair L==R only when the RIGHT node's leaf range
 * lies wholly inside the real leaves — exactly the pairs the reference's first (eager) loop
 * combines (:86); pairs produced by its final sweep (:109-140) are never tested. */
void orc_merkle_root(const uint8_t *leaves, size_t n, uint8_t root[32], int *mutated) {
    int mut = 0;
    if (n == 0) {
        memset(root, 0, 32);
        if (mutated) *mutated = 0;
        return;
    }
    uint8_t *cur = (uint8_t *)malloc(32 * n);
    memcpy(cur, leaves, 32 * n);
    size_t cnt = n, span = 1; /* span = leaves per node at this level */
    while (cnt > 1) {
        size_t nxt = (cnt + 1) / 2;
        for (size_t j = 0; j < nxt; j++) {
            const uint8_t *L = cur + 64 * j;
            const uint8_t *R = (2 * j + 1 < cnt) ? L + 32 : L;
            if (2 * j + 1 < cnt && (2 * j + 2) * span <= n && memcmp(L, R, 32) == 0) mut = 1;
            uint8_t cat[64];
            memcpy(cat, L, 32);
            memcpy(cat + 32, R, 32);
            orc_sha256d(cat, 64, cur + 32 * j);
        }
        cnt = nxt;
        span *= 2;
    }
    memcpy(root, cur, 32);
    free(cur);
    if (mutated) *mutated = mut;
}

/* ============================================================== block framing =============== */
/* compact size (src/serialize.h:260-300): canonical-encoding check as ReadCompactSize does */
static int read_compact(const uint8_t *p, size_t len, size_t *pos, uint64_t *v) {
    if (*pos >= len) return 0;
    uint8_t c = p[(*pos)++];
    if (c < 253) { *v = c; return 1; }
    int nb = c == 253 ? 2 : c == 254 ? 4 : 8;
    if (*pos + nb > len) return 0;
    uint64_t x = 0;
    for (int i = 0; i < nb; i++) x |= (uint64_t)p[*pos + i] << (8 * i);
    *pos += nb;
    if ((nb == 2 && x < 253) || (nb == 4 && x < 0x10000ULL) || (nb == 8 && x < 0x100000000ULL)) return 0;
    if (x > 0x02000000ULL) return 0; /* MAX_SIZE */
    *v = x;
    return 1;
}
/* one serialized tx (src/primitives/transaction.h:223-243): returns end offset or 0 */
static size_t skip_tx(const uint8_t *p, size_t len, size_t pos) {
    uint64_t nin, nout, sl;
    if (pos + 4 > len) return 0;
    pos += 4;
    if (!read_compact(p, len, &pos, &nin)) return 0;
    for (uint64_t i = 0; i < nin; i++) {
        if (pos + 36 > len) return 0;
        pos += 36;
        if (!read_compact(p, len, &pos, &sl)) return 0;
        if (pos + sl + 4 > len) return 0;
        pos += sl + 4;
    }
    if (!read_compact(p, len, &pos, &nout)) return 0;
    for (uint64_t i = 0; i < nout; i++) {
        if (pos + 8 > len) return 0;
        pos += 8;
        if (!read_compact(p, len, &pos, &sl)) return 0;
        if (pos + sl > len) return 0;
        pos += sl;
    }
    if (pos + 4 > len) return 0;
    return pos + 4;
}
long orc_block_index(const uint8_t *blk, size_t len, int legacy, uint32_t *tx_off, size_t max_tx) {
    size_t pos;
    uint64_t v;
    if (legacy) {
        pos = 80;
    } else {
        pos = 140;
        if (len < 140 || !read_compact(blk, len, &pos, &v)) return -1;
        pos += v;
    }
    if (pos > len) return -1;
    uint64_t ntx;
    if (!read_compact(blk, len, &pos, &ntx)) return -1;
    if (ntx > max_tx) return -2;
    for (uint64_t i = 0; i < ntx; i++) {
        tx_off[i] = (uint32_t)pos;
        pos = skip_tx(blk, len, pos);
        if (!pos) return -1;
    }
    tx_off[ntx] = (uint32_t)pos;
    return (long)ntx;
}
/* txid = SHA256d(serialized tx) (src/primitives/transaction.cpp:73-75); BlockMerkleRoot (merkle.cpp:175-182) */
long orc_block_txids(const uint8_t *blk, size_t len, int legacy, uint8_t *txids, size_t max_tx, uint8_t root[32],
                     int *mutated) {
    uint32_t *off = (uint32_t *)malloc(sizeof(uint32_t) * (max_tx + 1));
    long ntx = orc_block_index(blk, len, legacy, off, max_tx);
    if (ntx >= 0) {
        for (long i = 0; i < ntx; i++) orc_sha256d(blk + off[i], off[i + 1] - off[i], txids + 32 * i);
        if (root) orc_merkle_root(txids, (size_t)ntx, root, mutated);
    }
    free(off);
    return ntx;
}
void orc_header_hash(const uint8_t *hdr, size_t len, uint8_t out[32]) { orc_sha256d(hdr, len, out); }

/* CheckProofOfWork (src/pow.cpp:141-161) with arith_uint256::SetCompact (src/arith_uint256.cpp) */
int orc_check_pow(const uint8_t hash32[32], uint32_t nbits, const uint8_t pow_limit32[32]) {
    uint8_t target[32] = {0}; /* little-endian like uint256 */
    int size = nbits >> 24;
    uint32_t word = nbits & 0x007fffff;
    int negative = word != 0 && (nbits & 0x00800000) != 0;
    int overflow = word != 0 && ((size > 34) || (word > 0xff && size > 33) || (word > 0xffff && size > 32));
    if (size <= 3) {
        word >>= 8 * (3 - size);
        target[0] = word; target[1] = word >> 8; target[2] = word >> 16;
    } else {
        int sh = size - 3;
        for (int i = 0; i < 3; i++)
            if (sh + i < 32) target[sh + i] = (uint8_t)(word >> (8 * i));
    }
    int zero = 1;
    for (int i = 0; i < 32; i++) if (target[i]) zero = 0;
    if (negative || zero || overflow) return 0;
    for (int i = 31; i >= 0; i--) { /* target > powLimit ? */
        if (target[i] != pow_limit32[i]) { if (target[i] > pow_limit32[i]) return 0; break; }
    }
    for (int i = 31; i >= 0; i--) { /* hash > target ? */
        if (hash32[i] != target[i]) return hash32[i] < target[i];
    }
    return 1;
}

/* ============================================================== Blake2b (RFC 7693) ========== */
/* Stands in for libsodium 1.0.15 crypto_generichash_blake2b_{init_salt_personal,update,final}
 * (EXTERNAL dependency, depends/packages/libsodium.mk:6-9); reference call sites
 * src/crypto/equihash.cpp:44-59, src/pow.cpp:304-315. */
static const uint64_t B2B_IV[8] = {0x6a09e667f3bcc908ULL, 0xbb67ae8584caa73bULL, 0x3c6ef372fe94f82bULL,
                                   0xa54ff53a5f1d36f1ULL, 0x510e527fade682d1ULL, 0x9b05688c2b3e6c1fULL,
                                   0x1f83d9abfb41bd6bULL, 0x5be0cd19137e2179ULL};
static const uint8_t B2B_SIGMA[12][16] = {
    {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3},
    {11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4}, {7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8},
    {9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13}, {2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9},
    {12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11}, {13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10},
    {6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5}, {10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0},
    {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3}};
typedef struct {
    uint64_t h[8];
    uint64_t t[2];
    uint64_t f[2];
    uint8_t buf[256];
    size_t buflen;
    uint8_t outlen;
} b2b_state; /* layout mirrored by oracle/shim/sodium.h */
static inline uint64_t ror64(uint64_t x, int c) { return (x >> c) | (x << (64 - c)); }
static inline uint64_t ld64(const uint8_t *p) {
    uint64_t v = 0;
    for (int i = 0; i < 8; i++) v |= (uint64_t)p[i] << (8 * i);
    return v;
}
static void b2b_compress(b2b_state *S, const uint8_t blk[128]) {
    uint64_t m[16], v[16];
    for (int i = 0; i < 16; i++) m[i] = ld64(blk + 8 * i);
    for (int i = 0; i < 8; i++) { v[i] = S->h[i]; v[i + 8] = B2B_IV[i]; }
    v[12] ^= S->t[0]; v[13] ^= S->t[1]; v[14] ^= S->f[0]; v[15] ^= S->f[1];
#define B2G(r, i, a, b, c, d)                        \
    a = a + b + m[B2B_SIGMA[r][2 * i]];              \
    d = ror64(d ^ a, 32); c = c + d; b = ror64(b ^ c, 24); \
    a = a + b + m[B2B_SIGMA[r][2 * i + 1]];          \
    d = ror64(d ^ a, 16); c = c + d; b = ror64(b ^ c, 63);
    for (int r = 0; r < 12; r++) {
        B2G(r, 0, v[0], v[4], v[8], v[12]); B2G(r, 1, v[1], v[5], v[9], v[13]);
        B2G(r, 2, v[2], v[6], v[10], v[14]); B2G(r, 3, v[3], v[7], v[11], v[15]);
        B2G(r, 4, v[0], v[5], v[10], v[15]); B2G(r, 5, v[1], v[6], v[11], v[12]);
        B2G(r, 6, v[2], v[7], v[8], v[13]); B2G(r, 7, v[3], v[4], v[9], v[14]);
    }
#undef B2G
    for (int i = 0; i < 8; i++) S->h[i] ^= v[i] ^ v[i + 8];
}
static int b2b_init(b2b_state *S, size_t outlen, const uint8_t *salt, const uint8_t *personal) {
    if (outlen == 0 || outlen > 64) return -1;
    uint8_t P[64] = {0};
    P[0] = (uint8_t)outlen; P[1] = 0 /*keylen*/; P[2] = 1 /*fanout*/; P[3] = 1 /*depth*/;
    if (salt) memcpy(P + 32, salt, 16);
    if (personal) memcpy(P + 48, personal, 16);
    for (int i = 0; i < 8; i++) S->h[i] = B2B_IV[i] ^ ld64(P + 8 * i);
    S->t[0] = S->t[1] = S->f[0] = S->f[1] = 0;
    S->buflen = 0;
    S->outlen = (uint8_t)outlen;
    memset(S->buf, 0, sizeof S->buf);
    return 0;
}
static void b2b_update(b2b_state *S, const uint8_t *in, size_t inlen) {
    while (inlen > 0) {
        /* keep the last (possibly full) block buffered: it must be compressed with the final flag */
        if (S->buflen == 128) {
            S->t[0] += 128;
            if (S->t[0] < 128) S->t[1]++;
            b2b_compress(S, S->buf);
            S->buflen = 0;
        }
        size_t take = 128 - S->buflen < inlen ? 128 - S->buflen : inlen;
        memcpy(S->buf + S->buflen, in, take);
        S->buflen += take; in += take; inlen -= take;
    }
}
static void b2b_final(b2b_state *S, uint8_t *out, size_t outlen) {
    S->t[0] += S->buflen;
    if (S->t[0] < S->buflen) S->t[1]++;
    S->f[0] = ~0ULL;
    memset(S->buf + S->buflen, 0, 128 - S->buflen);
    b2b_compress(S, S->buf);
    uint8_t full[64];
    for (int i = 0; i < 8; i++)
        for (int j = 0; j < 8; j++) full[8 * i + j] = (uint8_t)(S->h[i] >> (8 * j));
    memcpy(out, full, outlen < S->outlen ? outlen : S->outlen);
}
/* libsodium-named entry points so the reference's equihash.cpp links against this restatement
 * when oracle/_ref is built (declared in oracle/shim/sodium.h) */
int crypto_generichash_blake2b_init_salt_personal(b2b_state *state, const unsigned char *key, const size_t keylen,
                                                  const size_t outlen, const unsigned char *salt,
                                                  const unsigned char *personal) {
    if (key != NULL && keylen != 0) return -1; /* keyed mode is not used by the reference's Equihash path */
    return b2b_init(state, outlen, salt, personal);
}
int crypto_generichash_blake2b_update(b2b_state *state, const unsigned char *in, unsigned long long inlen) {
    b2b_update(state, in, (size_t)inlen);
    return 0;
}
int crypto_generichash_blake2b_final(b2b_state *state, unsigned char *out, const size_t outlen) {
    b2b_state tmp = *state; /* libsodium finalises in place; callers here always work on a copy */
    b2b_final(&tmp, out, outlen);
    return 0;
}
void orc_blake2b(const uint8_t *in, size_t inlen, const uint8_t *personal16, size_t outlen, uint8_t *out) {
    b2b_state S;
    b2b_init(&S, outlen, NULL, personal16);
    b2b_update(&S, in, inlen);
    b2b_final(&S, out, outlen);
}

/* ============================================================== Equihash verify ============= */
/* IsValidSolution (src/crypto/equihash.cpp:724-770) + CheckEquihashSolution (src/pow.cpp:298-323) */
int orc_equihash_verify(unsigned N, unsigned K, const uint8_t *hdr, size_t len) {
    if (!((N == 200 && K == 9) || (N == 48 && K == 5) || (N == 96 && K == 5) || (N == 96 && K == 3))) return 0;
    const unsigned cbits = N / (K + 1);            /* CollisionBitLength   equihash.h:174 */
    const unsigned cbytes = (cbits + 7) / 8;       /* CollisionByteLength */
    const unsigned hash_len = (K + 1) * cbytes;    /* HashLength */
    const unsigned ipo = 512 / N;                  /* IndicesPerHashOutput */
    const unsigned hout = ipo * N / 8;             /* HashOutput */
    const size_t nidx = (size_t)1 << K;
    const size_t soln_w = nidx * (cbits + 1) / 8;  /* SolutionWidth */
    if (len < 140) return 0;
    size_t pos = 140;
    uint64_t sl;
    if (!read_compact(hdr, len, &pos, &sl)) return 0;
    if (pos + sl > len) return 0;
    if (sl != soln_w) return 0; /* :726-730 */
    const uint8_t *soln = hdr + pos;
    /* base state: personalization "ZcashPoW"||LE32(N)||LE32(K) (:36-48), then I||nonce = first 140 bytes */
    uint8_t pers[16];
    memcpy(pers, "ZcashPoW", 8);
    for (int i = 0; i < 4; i++) { pers[8 + i] = (uint8_t)(N >> (8 * i)); pers[12 + i] = (uint8_t)(K >> (8 * i)); }
    b2b_state base;
    b2b_init(&base, hout, NULL, pers);
    b2b_update(&base, hdr, 140);
    /* GetIndicesFromMinimal (:176-190): (cbits+1)-bit big-endian bit-packed indices */
    uint32_t *idx = (uint32_t *)malloc(sizeof(uint32_t) * nidx);
    {
        uint64_t acc = 0;
        unsigned accbits = 0;
        size_t j = 0;
        for (size_t i = 0; i < soln_w; i++) {
            acc = (acc << 8) | soln[i];
            accbits += 8;
            if (accbits >= cbits + 1) {
                accbits -= cbits + 1;
                idx[j++] = (uint32_t)((acc >> accbits) & ((1u << (cbits + 1)) - 1));
            }
        }
    }
    /* leaves: GenerateHash (:51-60) on index/ipo, slice, ExpandArray (:62-103) cbits-bit groups -> cbytes each */
    uint8_t *rows = (uint8_t *)malloc(nidx * hash_len);
    for (size_t i = 0; i < nidx; i++) {
        b2b_state st = base;
        uint32_t g = idx[i] / ipo;
        uint8_t le[4] = {(uint8_t)g, (uint8_t)(g >> 8), (uint8_t)(g >> 16), (uint8_t)(g >> 24)};
        uint8_t h[64];
        b2b_update(&st, le, 4);
        b2b_final(&st, h, hout);
        const uint8_t *in = h + (idx[i] % ipo) * (N / 8);
        uint8_t *out = rows + i * hash_len;
        uint32_t acc = 0;
        unsigned accbits = 0;
        size_t j = 0;
        for (unsigned b = 0; b < N / 8; b++) {
            acc = (acc << 8) | in[b];
            accbits += 8;
            if (accbits >= cbits) {
                accbits -= cbits;
                uint32_t v = (acc >> accbits) & ((1u << cbits) - 1);
                for (unsigned x = 0; x < cbytes; x++) out[j + x] = (uint8_t)(v >> (8 * (cbytes - 1 - x)));
                j += cbytes;
            }
        }
    }
    int ok = 1;
    size_t cnt = nidx, span = 1; /* span = indices per row */
    unsigned hl = hash_len;
    while (cnt > 1 && ok) {
        for (size_t i = 0; i < cnt && ok; i += 2) {
            uint8_t *a = rows + i * hash_len, *b = rows + (i + 1) * hash_len;
            if (memcmp(a, b, cbytes) != 0) { ok = 0; break; }           /* HasCollision :747 */
            const uint32_t *ia = idx + i * span, *ib = idx + (i + 1) * span;
            /* IndicesBefore(X[i+1], X[i]) :753 — lexicographic over big-endian index arrays */
            for (size_t t = 0; t < span; t++) {
                if (ib[t] != ia[t]) { if (ib[t] < ia[t]) ok = 0; break; }
            }
            if (!ok) break;
            for (size_t x = 0; x < span && ok; x++)                       /* DistinctIndices :757 */
                for (size_t y = 0; y < span; y++)
                    if (ia[x] == ib[y]) { ok = 0; break; }
            if (!ok) break;
            /* merged row (:224-247 ctor): XOR, drop cbytes; stored at slot i/2; indices stay in place */
            uint8_t *o = rows + (i / 2) * hash_len;
            for (unsigned t = cbytes; t < hl; t++) o[t - cbytes] = a[t] ^ b[t];
        }
        cnt /= 2;
        span *= 2;
        hl -= cbytes;
    }
    if (ok)
        for (unsigned t = 0; t < hl; t++) if (rows[t]) ok = 0; /* IsZero(hashLen) :769 */
    free(rows);
    free(idx);
    return ok;
}

/* ============================================================== 256-bit arithmetic ========== */
typedef struct { uint64_t v[4]; } u256; /* little-endian limbs */
static const u256 FE_P = {{0xFFFFFFFEFFFFFC2FULL, 0xFFFFFFFFFFFFFFFFULL, 0xFFFFFFFFFFFFFFFFULL, 0xFFFFFFFFFFFFFFFFULL}};
static const u256 SC_N = {{0xBFD25E8CD0364141ULL, 0xBAAEDCE6AF48A03BULL, 0xFFFFFFFFFFFFFFFEULL, 0xFFFFFFFFFFFFFFFFULL}};
static const u256 SC_NHALF = {{0xDFE92F46681B20A0ULL, 0x5D576E7357A4501DULL, 0xFFFFFFFFFFFFFFFFULL, 0x7FFFFFFFFFFFFFFFULL}};
/* 2^256 - p and 2^256 - n */
static const uint64_t FE_C[3] = {0x1000003D1ULL, 0, 0};
static const uint64_t SC_C[3] = {0x402DA1732FC9BEBFULL, 0x4551231950B75FC4ULL, 1};
/* p - n (second-chance bound, src/secp256k1/src/ecdsa_impl.h:45-47,258-266) */
static const u256 P_MINUS_N = {{0x402DA1722FC9BAEEULL, 0x4551231950B75FC4ULL, 1, 0}};

static int u256_cmp(const u256 *a, const u256 *b) {
    for (int i = 3; i >= 0; i--)
        if (a->v[i] != b->v[i]) return a->v[i] < b->v[i] ? -1 : 1;
    return 0;
}
static int u256_is_zero(const u256 *a) { return (a->v[0] | a->v[1] | a->v[2] | a->v[3]) == 0; }
static uint64_t u256_add(u256 *r, const u256 *a, const u256 *b) {
    u128 c = 0;
    for (int i = 0; i < 4; i++) { c += (u128)a->v[i] + b->v[i]; r->v[i] = (uint64_t)c; c >>= 64; }
    return (uint64_t)c;
}
static uint64_t u256_sub(u256 *r, const u256 *a, const u256 *b) {
    uint64_t br = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a->v[i] - b->v[i] - br;
        r->v[i] = (uint64_t)d;
        br = (uint64_t)(d >> 64) & 1;
    }
    return br;
}
static void u256_from_be(u256 *r, const uint8_t b[32]) {
    for (int i = 0; i < 4; i++) {
        uint64_t v = 0;
        for (int j = 0; j < 8; j++) v = (v << 8) | b[(3 - i) * 8 + j];
        r->v[i] = v;
    }
}
static void u256_to_be(uint8_t b[32], const u256 *a) {
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 8; j++) b[(3 - i) * 8 + j] = (uint8_t)(a->v[i] >> (56 - 8 * j));
}
/* r = t (512 bit) mod m, where m = 2^256 - c (c up to 3 limbs): fold the high half while it is non-zero */
static void mod_reduce512(u256 *r, const uint64_t t_in[8], const uint64_t c[3], const u256 *m) {
    uint64_t t[8];
    memcpy(t, t_in, sizeof t);
    for (;;) {
        if ((t[4] | t[5] | t[6] | t[7]) == 0) break;
        uint64_t acc[8] = {t[0], t[1], t[2], t[3], 0, 0, 0, 0};
        for (int i = 0; i < 4; i++) {
            uint64_t carry = 0;
            for (int j = 0; j < 3; j++) {
                u128 x = (u128)t[4 + i] * c[j] + acc[i + j] + carry;
                acc[i + j] = (uint64_t)x;
                carry = (uint64_t)(x >> 64);
            }
            for (int k = i + 3; carry && k < 8; k++) {
                u128 x = (u128)acc[k] + carry;
                acc[k] = (uint64_t)x;
                carry = (uint64_t)(x >> 64);
            }
        }
        memcpy(t, acc, sizeof t);
    }
    u256 x = {{t[0], t[1], t[2], t[3]}};
    while (u256_cmp(&x, m) >= 0) u256_sub(&x, &x, m);
    *r = x;
}
static void mod_mul(u256 *r, const u256 *a, const u256 *b, const uint64_t c[3], const u256 *m) {
    uint64_t t[8] = {0};
    for (int i = 0; i < 4; i++) {
        uint64_t carry = 0;
        for (int j = 0; j < 4; j++) {
            u128 x = (u128)a->v[i] * b->v[j] + t[i + j] + carry;
            t[i + j] = (uint64_t)x;
            carry = (uint64_t)(x >> 64);
        }
        t[i + 4] = carry;
    }
    mod_reduce512(r, t, c, m);
}
static void mod_add(u256 *r, const u256 *a, const u256 *b, const u256 *m) {
    uint64_t c = u256_add(r, a, b);
    if (c || u256_cmp(r, m) >= 0) u256_sub(r, r, m);
}
static void mod_sub(u256 *r, const u256 *a, const u256 *b, const u256 *m) {
    if (u256_sub(r, a, b)) u256_add(r, r, m);
}
static void mod_neg(u256 *r, const u256 *a, const u256 *m) {
    if (u256_is_zero(a)) { *r = *a; return; }
    u256_sub(r, m, a);
}
static void mod_pow(u256 *r, const u256 *a, const u256 *e, const uint64_t c[3], const u256 *m) {
    u256 acc = {{1, 0, 0, 0}}, base = *a;
    for (int i = 0; i < 256; i++) {
        if ((e->v[i / 64] >> (i % 64)) & 1) mod_mul(&acc, &acc, &base, c, m);
        mod_mul(&base, &base, &base, c, m);
    }
    *r = acc;
}
/* field helpers mod p.  fe_mul is the one specialised routine (p = 2^256 - 0x1000003D1 folds with a
 * single-limb constant); its result equals mod_mul(a, b, FE_C, FE_P), which the self-test checks. */
static void fe_mul(u256 *r, const u256 *a, const u256 *b) {
    uint64_t t[8] = {0};
    for (int i = 0; i < 4; i++) {
        uint64_t carry = 0;
        for (int j = 0; j < 4; j++) {
            u128 x = (u128)a->v[i] * b->v[j] + t[i + j] + carry;
            t[i + j] = (uint64_t)x;
            carry = (uint64_t)(x >> 64);
        }
        t[i + 4] = carry;
    }
    const uint64_t Cc = 0x1000003D1ULL;
    u128 acc = 0;
    u256 o;
    for (int i = 0; i < 4; i++) {
        acc += (u128)t[4 + i] * Cc + t[i];
        o.v[i] = (uint64_t)acc;
        acc >>= 64;
    }
    u128 x = (u128)(uint64_t)acc * Cc + o.v[0];
    o.v[0] = (uint64_t)x;
    uint64_t carry = (uint64_t)(x >> 64);
    for (int i = 1; i < 4; i++) {
        x = (u128)o.v[i] + carry;
        o.v[i] = (uint64_t)x;
        carry = (uint64_t)(x >> 64);
    }
    if (carry) { /* wrapped past 2^256 == Cc (mod p); the remainder is tiny so this cannot wrap again */
        x = (u128)o.v[0] + Cc;
        o.v[0] = (uint64_t)x;
        carry = (uint64_t)(x >> 64);
        for (int i = 1; i < 4 && carry; i++) {
            x = (u128)o.v[i] + carry;
            o.v[i] = (uint64_t)x;
            carry = (uint64_t)(x >> 64);
        }
    }
    if (u256_cmp(&o, &FE_P) >= 0) u256_sub(&o, &o, &FE_P);
    *r = o;
}
#define fe_sqr(r, a) fe_mul(r, a, a)
#define fe_add(r, a, b) mod_add(r, a, b, &FE_P)
#define fe_sub(r, a, b) mod_sub(r, a, b, &FE_P)
#define fe_neg(r, a) mod_neg(r, a, &FE_P)
static void fe_inv(u256 *r, const u256 *a) { /* a^(p-2) */
    u256 e = FE_P;
    e.v[0] -= 2;
    mod_pow(r, a, &e, FE_C, &FE_P);
}
/* sqrt: a^((p+1)/4), valid iff result^2 == a  (src/secp256k1/src/field_impl.h:38-134) */
static int fe_sqrt(u256 *r, const u256 *a) {
    u256 e = {{0xFFFFFFFFBFFFFF0CULL, 0xFFFFFFFFFFFFFFFFULL, 0xFFFFFFFFFFFFFFFFULL, 0x3FFFFFFFFFFFFFFFULL}};
    u256 t, s;
    mod_pow(&t, a, &e, FE_C, &FE_P);
    fe_sqr(&s, &t);
    *r = t;
    return u256_cmp(&s, a) == 0;
}
#define sc_mul(r, a, b) mod_mul(r, a, b, SC_C, &SC_N)
static void sc_inv(u256 *r, const u256 *a) { /* a^(n-2)  (src/secp256k1/src/scalar_impl.h:55-218 computes the same power) */
    u256 e = SC_N;
    e.v[0] -= 2;
    mod_pow(r, a, &e, SC_C, &SC_N);
}
/* secp256k1_scalar_set_b32 (scalar_4x64_impl.h / scalar_8x32_impl.h:165-179): reduce once, report overflow */
static int sc_set_b32(u256 *r, const uint8_t b[32]) {
    u256_from_be(r, b);
    if (u256_cmp(r, &SC_N) >= 0) { u256_sub(r, r, &SC_N); return 1; }
    return 0;
}

/* ============================================================== group ======================= */
typedef struct { u256 x, y; int inf; } ge;
typedef struct { u256 x, y, z; int inf; } gej;
static const ge GE_G = {{{0x59F2815B16F81798ULL, 0x029BFCDB2DCE28D9ULL, 0x55A06295CE870B07ULL, 0x79BE667EF9DCBBACULL}},
                        {{0x9C47D08FFB10D4B8ULL, 0xFD17B448A6855419ULL, 0x5DA4FBFC0E1108A8ULL, 0x483ADA7726A3C465ULL}},
                        0};
static void gej_set_ge(gej *r, const ge *a) {
    r->x = a->x; r->y = a->y; r->inf = a->inf;
    r->z = (u256){{1, 0, 0, 0}};
}
/* doubling on y^2 = x^3 + 7 (a = 0); same map as secp256k1_gej_double_var (group_impl.h:301-354) */
static void gej_double(gej *r, const gej *a) {
    if (a->inf) { *r = *a; return; } /* y = 0 is impossible on this curve (no order-2 points) */
    u256 A, B, C, D, E, F, t, x3, y3, z3;
    fe_sqr(&A, &a->x); fe_sqr(&B, &a->y); fe_sqr(&C, &B);
    fe_add(&t, &a->x, &B); fe_sqr(&t, &t); fe_sub(&t, &t, &A); fe_sub(&t, &t, &C); fe_add(&D, &t, &t);
    fe_add(&E, &A, &A); fe_add(&E, &E, &A);
    fe_sqr(&F, &E);
    fe_sub(&x3, &F, &D); fe_sub(&x3, &x3, &D);
    fe_sub(&t, &D, &x3); fe_mul(&t, &E, &t);
    u256 c8; fe_add(&c8, &C, &C); fe_add(&c8, &c8, &c8); fe_add(&c8, &c8, &c8);
    fe_sub(&y3, &t, &c8);
    fe_mul(&z3, &a->y, &a->z); fe_add(&z3, &z3, &z3);
    r->x = x3; r->y = y3; r->z = z3; r->inf = 0;
}
/* general addition with every degenerate branch of secp256k1_gej_add_var / _add_ge_var (group_impl.h:356-461) */
static void gej_add(gej *r, const gej *a, const gej *b) {
    if (a->inf) { *r = *b; return; }
    if (b->inf) { *r = *a; return; }
    u256 z1z1, z2z2, u1, u2, s1, s2, h, rr, t;
    fe_sqr(&z1z1, &a->z); fe_sqr(&z2z2, &b->z);
    fe_mul(&u1, &a->x, &z2z2); fe_mul(&u2, &b->x, &z1z1);
    fe_mul(&s1, &a->y, &z2z2); fe_mul(&s1, &s1, &b->z);
    fe_mul(&s2, &b->y, &z1z1); fe_mul(&s2, &s2, &a->z);
    fe_sub(&h, &u2, &u1); fe_sub(&rr, &s2, &s1);
    if (u256_is_zero(&h)) {
        if (u256_is_zero(&rr)) { gej_double(r, a); return; }
        memset(r, 0, sizeof *r); r->inf = 1; return;
    }
    u256 h2, h3, v, x3, y3, z3;
    fe_sqr(&h2, &h); fe_mul(&h3, &h2, &h); fe_mul(&v, &u1, &h2);
    fe_sqr(&x3, &rr); fe_sub(&x3, &x3, &h3); fe_sub(&x3, &x3, &v); fe_sub(&x3, &x3, &v);
    fe_sub(&t, &v, &x3); fe_mul(&t, &t, &rr); fe_mul(&y3, &s1, &h3); fe_sub(&y3, &t, &y3);
    fe_mul(&z3, &a->z, &b->z); fe_mul(&z3, &z3, &h);
    r->x = x3; r->y = y3; r->z = z3; r->inf = 0;
}
/* mixed addition (b affine), branches as secp256k1_gej_add_ge_var (group_impl.h:414-461) */
static void gej_add_ge(gej *r, const gej *a, const ge *b) {
    if (b->inf) { *r = *a; return; }
    if (a->inf) { gej_set_ge(r, b); return; }
    u256 z1z1, u2, s2, h, rr, t;
    fe_sqr(&z1z1, &a->z);
    fe_mul(&u2, &b->x, &z1z1);
    fe_mul(&s2, &b->y, &z1z1); fe_mul(&s2, &s2, &a->z);
    fe_sub(&h, &u2, &a->x); fe_sub(&rr, &s2, &a->y);
    if (u256_is_zero(&h)) {
        if (u256_is_zero(&rr)) { gej_double(r, a); return; }
        memset(r, 0, sizeof *r); r->inf = 1; return;
    }
    u256 h2, h3, v, x3, y3, z3;
    fe_sqr(&h2, &h); fe_mul(&h3, &h2, &h); fe_mul(&v, &a->x, &h2);
    fe_sqr(&x3, &rr); fe_sub(&x3, &x3, &h3); fe_sub(&x3, &x3, &v); fe_sub(&x3, &x3, &v);
    fe_sub(&t, &v, &x3); fe_mul(&t, &t, &rr); fe_mul(&y3, &a->y, &h3); fe_sub(&y3, &t, &y3);
    fe_mul(&z3, &a->z, &h);
    r->x = x3; r->y = y3; r->z = z3; r->inf = 0;
}
static void ge_set_gej(ge *r, const gej *a) {
    if (a->inf) { memset(r, 0, sizeof *r); r->inf = 1; return; }
    u256 zi, zi2, zi3;
    fe_inv(&zi, &a->z); fe_sqr(&zi2, &zi); fe_mul(&zi3, &zi2, &zi);
    fe_mul(&r->x, &a->x, &zi2); fe_mul(&r->y, &a->y, &zi3);
    r->inf = 0;
}
/* R = na*A + ng*G: same result as secp256k1_ecmult (ecmult_impl.h:286-404), evaluated with fixed
 * 4-bit windows (Shamir).  The G table is built once. */
static ge G_COMB[64][16]; /* G_COMB[w][d] = d * 16^w * G (affine) */
static pthread_once_t g_tab_once = PTHREAD_ONCE_INIT;
static void build_g_tab(void) {
    gej base;
    gej_set_ge(&base, &GE_G);
    for (int w = 0; w < 64; w++) {
        gej acc;
        memset(&acc, 0, sizeof acc); acc.inf = 1;
        memset(&G_COMB[w][0], 0, sizeof(ge)); G_COMB[w][0].inf = 1;
        for (int d = 1; d < 16; d++) {
            gej_add(&acc, &acc, &base);
            ge_set_gej(&G_COMB[w][d], &acc);
        }
        gej_add(&base, &acc, &base); /* 16 * base */
    }
}
static void ecmult_gen_acc(gej *acc, const u256 *ng) {
    for (int w = 0; w < 64; w++) {
        unsigned dg = (ng->v[w / 16] >> ((w % 16) * 4)) & 15;
        if (dg) gej_add_ge(acc, acc, &G_COMB[w][dg]);
    }
}
static void ecmult(gej *r, const gej *a, const u256 *na, const u256 *ng) {
    pthread_once(&g_tab_once, build_g_tab);
    gej acc;
    memset(&acc, 0, sizeof acc); acc.inf = 1;
    if (!a->inf && !u256_is_zero(na)) {
        gej at[16];
        memset(&at[0], 0, sizeof(gej)); at[0].inf = 1;
        at[1] = *a;
        for (int i = 2; i < 16; i++) gej_add(&at[i], &at[i - 1], a);
        for (int w = 63; w >= 0; w--) {
            for (int k = 0; k < 4; k++) gej_double(&acc, &acc);
            unsigned da = (na->v[w / 16] >> ((w % 16) * 4)) & 15;
            if (da) gej_add(&acc, &acc, &at[da]);
        }
    }
    ecmult_gen_acc(&acc, ng);
    *r = acc;
}

/* ============================================================== ECDSA ======================= */
/* secp256k1_eckey_pubkey_parse (eckey_impl.h:17-35) + ge_set_xo_var (group_impl.h:216-237) +
 * ge_is_valid_var (:287-299) + fe_set_b32 range check (field_10x26_impl.h:323-338) */
static int pubkey_parse_ge(ge *q, const uint8_t *in, size_t len) {
    if (len == 33 && (in[0] == 0x02 || in[0] == 0x03)) {
        u256 x, x3, y;
        u256_from_be(&x, in + 1);
        if (u256_cmp(&x, &FE_P) >= 0) return 0;
        fe_sqr(&x3, &x); fe_mul(&x3, &x3, &x);
        u256 seven = {{7, 0, 0, 0}};
        fe_add(&x3, &x3, &seven);
        if (!fe_sqrt(&y, &x3)) return 0;
        if ((y.v[0] & 1) != (uint64_t)(in[0] == 0x03)) fe_neg(&y, &y);
        q->x = x; q->y = y; q->inf = 0;
        return 1;
    } else if (len == 65 && (in[0] == 0x04 || in[0] == 0x06 || in[0] == 0x07)) {
        u256 x, y, l, r;
        u256_from_be(&x, in + 1); u256_from_be(&y, in + 33);
        if (u256_cmp(&x, &FE_P) >= 0 || u256_cmp(&y, &FE_P) >= 0) return 0;
        if ((in[0] == 0x06 || in[0] == 0x07) && ((y.v[0] & 1) != (uint64_t)(in[0] == 0x07))) return 0;
        fe_sqr(&l, &y);
        fe_sqr(&r, &x); fe_mul(&r, &r, &x);
        u256 seven = {{7, 0, 0, 0}};
        fe_add(&r, &r, &seven);
        if (u256_cmp(&l, &r) != 0) return 0;
        q->x = x; q->y = y; q->inf = 0;
        return 1;
    }
    return 0;
}
int orc_pubkey_parse(const uint8_t *in, size_t len, uint8_t xy64[64]) {
    ge q;
    memset(xy64, 0, 64);
    if (!pubkey_parse_ge(&q, in, len)) return 0;
    u256_to_be(xy64, &q.x); u256_to_be(xy64 + 32, &q.y);
    return 1;
}
/* secp256k1_ecdsa_signature_parse_compact (secp256k1.c:228-247): overflow => (0,0), return 0 */
int orc_sig_parse_compact(const uint8_t in64[64], uint8_t out64[64]) {
    u256 r, s;
    int ov = sc_set_b32(&r, in64) | sc_set_b32(&s, in64 + 32);
    if (ov) { memset(out64, 0, 64); return 0; }
    u256_to_be(out64, &r); u256_to_be(out64 + 32, &s);
    return 1;
}
/* secp256k1_ecdsa_signature_normalize (secp256k1.c:274-291); is_high: s > n/2 (scalar_8x32_impl.h:220-236) */
int orc_sig_normalize(uint8_t sig64[64]) {
    u256 s;
    u256_from_be(&s, sig64 + 32);
    if (u256_cmp(&s, &SC_NHALF) > 0) {
        mod_neg(&s, &s, &SC_N);
        u256_to_be(sig64 + 32, &s);
        return 1;
    }
    return 0;
}
/* ecdsa_signature_parse_der_lax (src/pubkey.cpp:26-168): same accept/reject and same (r,s) */
int orc_der_lax_parse(const uint8_t *input, size_t inputlen, uint8_t sig64[64]) {
    size_t rpos, rlen, spos, slen, pos = 0, lenbyte;
    uint8_t tmp[64] = {0};
    int overflow = 0;
    memset(sig64, 0, 64);
    if (pos == inputlen || input[pos] != 0x30) return 0;
    pos++;
    if (pos == inputlen) return 0;
    lenbyte = input[pos++];
    if (lenbyte & 0x80) {
        lenbyte -= 0x80;
        if (pos + lenbyte > inputlen) return 0;
        pos += lenbyte;
    }
    for (int which = 0; which < 2; which++) {
        size_t vlen;
        if (pos == inputlen || input[pos] != 0x02) return 0;
        pos++;
        if (pos == inputlen) return 0;
        lenbyte = input[pos++];
        if (lenbyte & 0x80) {
            lenbyte -= 0x80;
            if (pos + lenbyte > inputlen) return 0;
            while (lenbyte > 0 && input[pos] == 0) { pos++; lenbyte--; }
            if (lenbyte >= sizeof(size_t)) return 0;
            vlen = 0;
            while (lenbyte > 0) { vlen = (vlen << 8) + input[pos]; pos++; lenbyte--; }
        } else {
            vlen = lenbyte;
        }
        if (vlen > inputlen - pos) return 0;
        if (which == 0) { rpos = pos; rlen = vlen; } else { spos = pos; slen = vlen; }
        pos += vlen;
    }
    while (rlen > 0 && input[rpos] == 0) { rlen--; rpos++; }
    if (rlen > 32) overflow = 1; else memcpy(tmp + 32 - rlen, input + rpos, rlen);
    while (slen > 0 && input[spos] == 0) { slen--; spos++; }
    if (slen > 32) overflow = 1; else memcpy(tmp + 64 - slen, input + spos, slen);
    if (!overflow) overflow = !orc_sig_parse_compact(tmp, sig64);
    if (overflow) memset(sig64, 0, 64);
    return 1;
}
/* CPubKey::CheckLowS (src/pubkey.cpp:294-302) */
int orc_check_low_s(const uint8_t *der, size_t len) {
    uint8_t sig[64];
    if (!orc_der_lax_parse(der, len, sig)) return 0;
    u256 s;
    u256_from_be(&s, sig + 32);
    return u256_cmp(&s, &SC_NHALF) <= 0;
}
/* secp256k1_ecdsa_verify (secp256k1.c:293-308) + secp256k1_ecdsa_sig_verify (ecdsa_impl.h:201-269) on parsed inputs */
static int ecdsa_verify_core(const u256 *r, const u256 *s, const uint8_t msg32[32], const ge *q) {
    u256 m, sn, u1, u2;
    if (u256_cmp(s, &SC_NHALF) > 0) return 0;                   /* !is_high(s)            secp256k1.c:305 */
    if (u256_is_zero(r) || u256_is_zero(s)) return 0;           /*                         ecdsa_impl.h:210 */
    sc_set_b32(&m, msg32);                                      /* reduced, overflow ignored secp256k1.c:303 */
    sc_inv(&sn, s);
    sc_mul(&u1, &sn, &m);
    sc_mul(&u2, &sn, r);
    gej qj, R;
    gej_set_ge(&qj, q);
    ecmult(&R, &qj, &u2, &u1);
    if (R.inf) return 0;                                        /* :220 */
    /* gej_eq_x_var: r*Z^2 == X, then the r+n < p retry (:246-266) */
    u256 z2, t;
    fe_sqr(&z2, &R.z);
    fe_mul(&t, r, &z2);
    if (u256_cmp(&t, &R.x) == 0) return 1;
    if (u256_cmp(r, &P_MINUS_N) >= 0) return 0;
    u256 rn;
    u256_add(&rn, r, &SC_N);
    fe_mul(&t, &rn, &z2);
    return u256_cmp(&t, &R.x) == 0;
}
int orc_ecdsa_verify(const uint8_t msg32[32], const uint8_t *pub, size_t publen, const uint8_t sig64[64]) {
    ge q;
    uint8_t sig[64];
    if (!pubkey_parse_ge(&q, pub, publen)) return 0;
    if (!orc_sig_parse_compact(sig64, sig)) return 0;
    u256 r, s;
    u256_from_be(&r, sig); u256_from_be(&s, sig + 32);
    return ecdsa_verify_core(&r, &s, msg32, &q);
}
/* secp256k1_ecdsa_recoverable_signature_parse_compact + secp256k1_ecdsa_recover
 * (src/secp256k1/src/modules/recovery/main_impl.h:38-58, 170-190) with secp256k1_ecdsa_sig_recover (:87-121):
 * x = r (+ n when recid & 2, only if r < p - n), lift with the parity recid & 1, Q = (s/r) X - (m/r) G.
 * xy64 = affine x || y of the recovered key; returns 0 (and zeros) when nothing can be recovered. */
int orc_ecdsa_recover(const uint8_t msg32[32], const uint8_t sig64[64], int recid, uint8_t xy64[64]) {
    memset(xy64, 0, 64);
    if (recid < 0 || recid > 3) return 0;
    u256 r, s, m;
    if (sc_set_b32(&r, sig64) || sc_set_b32(&s, sig64 + 32)) return 0; /* parse_compact: overflow => fail */
    if (u256_is_zero(&r) || u256_is_zero(&s)) return 0;                /* :97-99 */
    u256 fx = r;
    if (recid & 2) {
        if (u256_cmp(&fx, &P_MINUS_N) >= 0) return 0;                  /* :106-108 */
        u256_add(&fx, &fx, &SC_N);
    }
    uint8_t pub[33];
    pub[0] = (recid & 1) ? 0x03 : 0x02;
    u256_to_be(pub + 1, &fx);
    ge x;
    if (!pubkey_parse_ge(&x, pub, 33)) return 0;                       /* ge_set_xo_var :111 */
    sc_set_b32(&m, msg32);
    u256 rn, u1, u2;
    sc_inv(&rn, &r);
    sc_mul(&u1, &rn, &m);
    mod_neg(&u1, &u1, &SC_N);
    sc_mul(&u2, &rn, &s);
    gej xj, q;
    gej_set_ge(&xj, &x);
    ecmult(&q, &xj, &u2, &u1);
    if (q.inf) return 0;
    ge qa;
    ge_set_gej(&qa, &q);
    u256_to_be(xy64, &qa.x); u256_to_be(xy64 + 32, &qa.y);
    return 1;
}
/* secp256k1_ec_pubkey_tweak_add (src/secp256k1/src/secp256k1.c:477-499) + secp256k1_eckey_pubkey_tweak_add
 * (eckey_impl.h:63-75): Q = P + tweak*G; fails when tweak >= n or the sum is the point at infinity. */
int orc_pubkey_tweak_add(const uint8_t *pub, size_t publen, const uint8_t tweak32[32], uint8_t xy64[64]) {
    memset(xy64, 0, 64);
    ge p;
    u256 t, one = {{1, 0, 0, 0}};
    if (!pubkey_parse_ge(&p, pub, publen)) return 0;
    if (sc_set_b32(&t, tweak32)) return 0;
    gej pj, q;
    gej_set_ge(&pj, &p);
    ecmult(&q, &pj, &one, &t);
    if (q.inf) return 0;
    ge qa;
    ge_set_gej(&qa, &q);
    u256_to_be(xy64, &qa.x); u256_to_be(xy64 + 32, &qa.y);
    return 1;
}
/* CPubKey::Verify (src/pubkey.cpp:170-193); CPubKey validity/length by prefix (src/pubkey.h:47-66) */
static size_t publen_by_prefix(uint8_t h) {
    if (h == 2 || h == 3) return 33;
    if (h == 4 || h == 6 || h == 7) return 65;
    return 0;
}
int orc_pubkey_verify(const uint8_t hash32[32], const uint8_t *pub, size_t publen, const uint8_t *der, size_t derlen) {
    ge q;
    uint8_t sig[64];
    if (publen == 0 || publen_by_prefix(pub[0]) != publen) return 0; /* IsValid() */
    if (!pubkey_parse_ge(&q, pub, publen)) return 0;
    if (derlen == 0) return 0;
    if (!orc_der_lax_parse(der, derlen, sig)) return 0;
    orc_sig_normalize(sig);
    u256 r, s;
    u256_from_be(&r, sig); u256_from_be(&s, sig + 32);
    return ecdsa_verify_core(&r, &s, hash32, &q);
}

typedef struct {
    void (*fn)(size_t, size_t, void *);
    void *arg;
    size_t lo, hi;
} par_job;
static void *par_tramp(void *p) {
    par_job *j = (par_job *)p;
    j->fn(j->lo, j->hi, j->arg);
    return NULL;
}
static void parallel_for(size_t n, int nthreads, void (*fn)(size_t, size_t, void *), void *arg) {
    if (nthreads <= 1 || n < 2) { fn(0, n, arg); return; }
    if ((size_t)nthreads > n) nthreads = (int)n;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * nthreads);
    par_job *jobs = (par_job *)malloc(sizeof(par_job) * nthreads);
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; t++) {
        size_t lo = (size_t)t * chunk, hi = lo + chunk;
        if (lo > n) lo = n;
        if (hi > n) hi = n;
        jobs[t] = (par_job){fn, arg, lo, hi};
        pthread_create(&th[t], NULL, par_tramp, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}
typedef struct { const uint8_t *msg, *pub, *sig; uint8_t *ok; } vb_arg;
static void vb_run(size_t lo, size_t hi, void *p) {
    vb_arg *a = (vb_arg *)p;
    for (size_t i = lo; i < hi; i++) {
        size_t len = publen_by_prefix(a->pub[65 * i]);
        a->ok[i] = len ? (uint8_t)orc_ecdsa_verify(a->msg + 32 * i, a->pub + 65 * i, len, a->sig + 64 * i) : 0;
    }
}
void orc_ecdsa_verify_batch(const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64, size_t n, int nthreads,
                            uint8_t *ok) {
    vb_arg a = {msg32, pub65, sig64, ok};
    pthread_once(&g_tab_once, build_g_tab);
    parallel_for(n, nthreads, vb_run, &a);
}
typedef struct { unsigned n, k; const uint8_t *h; size_t stride, len; uint8_t *ok; } eb_arg;
static void eb_run(size_t lo, size_t hi, void *p) {
    eb_arg *a = (eb_arg *)p;
    for (size_t i = lo; i < hi; i++) a->ok[i] = (uint8_t)orc_equihash_verify(a->n, a->k, a->h + i * a->stride, a->len);
}
void orc_equihash_verify_batch(unsigned n, unsigned k, const uint8_t *hdrs, size_t stride, size_t len, size_t count,
                               int nthreads, uint8_t *ok) {
    eb_arg a = {n, k, hdrs, stride, len, ok};
    parallel_for(count, nthreads, eb_run, &a);
}

/* ---- key generation and RFC 6979 signing (test-data generator; mirrors secp256k1_ec_pubkey_create
 *      secp256k1.c:395-420, secp256k1_ecdsa_sign :346-390, nonce_function_rfc6979 :310-341,
 *      hash_impl.h:205-264, secp256k1_ecdsa_sig_sign ecdsa_impl.h:271-311) */
static void ecmult_gen(gej *r, const u256 *k) {
    gej inf;
    memset(&inf, 0, sizeof inf); inf.inf = 1;
    u256 zero = {{0, 0, 0, 0}};
    ecmult(r, &inf, &zero, k);
}
int orc_pubkey_create(const uint8_t seckey32[32], int compressed, uint8_t *out, size_t *outlen) {
    u256 d;
    if (sc_set_b32(&d, seckey32) || u256_is_zero(&d)) return 0;
    gej pj; ge p;
    ecmult_gen(&pj, &d);
    ge_set_gej(&p, &pj);
    if (compressed) {
        out[0] = 0x02 | (uint8_t)(p.y.v[0] & 1);
        u256_to_be(out + 1, &p.x);
        *outlen = 33;
    } else {
        out[0] = 0x04;
        u256_to_be(out + 1, &p.x); u256_to_be(out + 33, &p.y);
        *outlen = 65;
    }
    return 1;
}
static void hmac_sha256(const uint8_t *key, size_t keylen, const uint8_t *const *parts, const size_t *lens, int nparts,
                        uint8_t out[32]) {
    uint8_t k[64] = {0}, pad[64], inner[32];
    if (keylen > 64) orc_sha256(key, keylen, k); else memcpy(k, key, keylen);
    sha256_ctx c;
    for (int i = 0; i < 64; i++) pad[i] = k[i] ^ 0x36;
    sha256_init(&c); sha256_update(&c, pad, 64);
    for (int i = 0; i < nparts; i++) sha256_update(&c, parts[i], lens[i]);
    sha256_final(&c, inner);
    for (int i = 0; i < 64; i++) pad[i] = k[i] ^ 0x5c;
    sha256_init(&c); sha256_update(&c, pad, 64); sha256_update(&c, inner, 32);
    sha256_final(&c, out);
}
int orc_sign_rfc6979(const uint8_t seckey32[32], const uint8_t msg32[32], uint8_t sig64[64]) {
    u256 d, m;
    if (sc_set_b32(&d, seckey32) || u256_is_zero(&d)) { memset(sig64, 0, 64); return 0; }
    sc_set_b32(&m, msg32);
    uint8_t keydata[64], V[32], Kk[32];
    memcpy(keydata, seckey32, 32); memcpy(keydata + 32, msg32, 32);
    memset(V, 1, 32); memset(Kk, 0, 32);
    static const uint8_t zero = 0, one = 1;
    { const uint8_t *p[3] = {V, &zero, keydata}; size_t l[3] = {32, 1, 64}; hmac_sha256(Kk, 32, p, l, 3, Kk); }
    { const uint8_t *p[1] = {V}; size_t l[1] = {32}; hmac_sha256(Kk, 32, p, l, 1, V); }
    { const uint8_t *p[3] = {V, &one, keydata}; size_t l[3] = {32, 1, 64}; hmac_sha256(Kk, 32, p, l, 3, Kk); }
    { const uint8_t *p[1] = {V}; size_t l[1] = {32}; hmac_sha256(Kk, 32, p, l, 1, V); }
    /* the reference re-initialises the generator for every attempt and draws (count+1) outputs; drawing
     * sequentially with the 3.2.h retry step between outputs yields the same sequence */
    for (int attempt = 0; attempt < 1000; attempt++) {
        if (attempt > 0) {
            const uint8_t *p[2] = {V, &zero}; size_t l[2] = {32, 1};
            hmac_sha256(Kk, 32, p, l, 2, Kk);
            const uint8_t *p2[1] = {V}; size_t l2[1] = {32};
            hmac_sha256(Kk, 32, p2, l2, 1, V);
        }
        { const uint8_t *p[1] = {V}; size_t l[1] = {32}; hmac_sha256(Kk, 32, p, l, 1, V); }
        u256 k;
        if (sc_set_b32(&k, V) || u256_is_zero(&k)) continue;
        gej Rj; ge R;
        ecmult_gen(&Rj, &k);
        ge_set_gej(&R, &Rj);
        uint8_t xb[32];
        u256_to_be(xb, &R.x);
        u256 r, s, t, ki;
        sc_set_b32(&r, xb);
        if (u256_is_zero(&r)) continue;
        sc_mul(&t, &r, &d);
        mod_add(&t, &t, &m, &SC_N);
        sc_inv(&ki, &k);
        sc_mul(&s, &ki, &t);
        if (u256_is_zero(&s)) continue;
        if (u256_cmp(&s, &SC_NHALF) > 0) mod_neg(&s, &s, &SC_N);
        u256_to_be(sig64, &r); u256_to_be(sig64 + 32, &s);
        return 1;
    }
    return 0;
}
/* strict DER of (r,s) (secp256k1_ecdsa_sig_serialize, ecdsa_impl.h:171-199) */
int orc_sig_to_der(const uint8_t sig64[64], uint8_t *der, size_t *derlen) {
    uint8_t r[33] = {0}, s[33] = {0};
    memcpy(r + 1, sig64, 32); memcpy(s + 1, sig64 + 32, 32);
    const uint8_t *rp = r, *sp = s;
    size_t lr = 33, ls = 33;
    while (lr > 1 && rp[0] == 0 && rp[1] < 0x80) { lr--; rp++; }
    while (ls > 1 && sp[0] == 0 && sp[1] < 0x80) { ls--; sp++; }
    if (*derlen < 6 + lr + ls) return 0;
    der[0] = 0x30; der[1] = (uint8_t)(4 + lr + ls); der[2] = 0x02; der[3] = (uint8_t)lr;
    memcpy(der + 4, rp, lr);
    der[4 + lr] = 0x02; der[5 + lr] = (uint8_t)ls;
    memcpy(der + 6 + lr, sp, ls);
    *derlen = 6 + lr + ls;
    return 1;
}

/* ---- synthetic signature set: key j = SHA256(seed||"key"||LE64(j)) mod n, msg i =
 *      SHA256d(seed||"msg"||LE64(i)), signer = RFC6979; pubkey compressed unless i % 10 == 9 */
typedef struct {
    const uint8_t *seed; size_t seedlen, nkeys; uint8_t *msg, *pub, *sig; uint8_t *keys; uint8_t *pubc, *pubu;
} gs_arg;
static void tagged(const uint8_t *seed, size_t seedlen, const char *tag, uint64_t i, int dbl, uint8_t out[32]) {
    uint8_t buf[256];
    size_t tl = strlen(tag);
    memcpy(buf, seed, seedlen); memcpy(buf + seedlen, tag, tl);
    for (int b = 0; b < 8; b++) buf[seedlen + tl + b] = (uint8_t)(i >> (8 * b));
    if (dbl) orc_sha256d(buf, seedlen + tl + 8, out); else orc_sha256(buf, seedlen + tl + 8, out);
}
static void gs_keys(size_t lo, size_t hi, void *p) {
    gs_arg *a = (gs_arg *)p;
    for (size_t j = lo; j < hi; j++) {
        uint8_t *k = a->keys + 32 * j;
        tagged(a->seed, a->seedlen, "key", j, 0, k);
        u256 d;
        sc_set_b32(&d, k);
        if (u256_is_zero(&d)) d.v[0] = 1;
        u256_to_be(k, &d);
        size_t l;
        orc_pubkey_create(k, 1, a->pubc + 33 * j, &l);
        orc_pubkey_create(k, 0, a->pubu + 65 * j, &l);
    }
}
static void gs_sigs(size_t lo, size_t hi, void *p) {
    gs_arg *a = (gs_arg *)p;
    for (size_t i = lo; i < hi; i++) {
        size_t j = i % a->nkeys;
        tagged(a->seed, a->seedlen, "msg", i, 1, a->msg + 32 * i);
        orc_sign_rfc6979(a->keys + 32 * j, a->msg + 32 * i, a->sig + 64 * i);
        memset(a->pub + 65 * i, 0, 65);
        if (i % 10 == 9) memcpy(a->pub + 65 * i, a->pubu + 65 * j, 65);
        else memcpy(a->pub + 65 * i, a->pubc + 33 * j, 33);
    }
}
void orc_gen_sigset(const uint8_t *seed, size_t seedlen, size_t n, size_t nkeys, int nthreads, uint8_t *msg32,
                    uint8_t *pub65, uint8_t *sig64) {
    if (seedlen > 200) seedlen = 200;
    gs_arg a = {seed, seedlen, nkeys, msg32, pub65, sig64, NULL, NULL, NULL};
    a.keys = (uint8_t *)malloc(32 * nkeys);
    a.pubc = (uint8_t *)malloc(33 * nkeys);
    a.pubu = (uint8_t *)malloc(65 * nkeys);
    pthread_once(&g_tab_once, build_g_tab);
    parallel_for(nkeys, nthreads, gs_keys, &a);
    parallel_for(n, nthreads, gs_sigs, &a);
    free(a.keys); free(a.pubc); free(a.pubu);
}

/* ============================================================== FORKID sighash ============== */
/* SignatureHash, FORKID branch (src/script/interpreter.cpp:1357-1404) with GetPrevoutHash /
 * GetSequenceHash / GetOutputsHash (:1321-1343) computed directly from the serialized tx. */
int orc_sighash_forkid(const uint8_t *script_code, size_t sc_len, const uint8_t *tx, size_t txlen, unsigned nin,
                       uint32_t hashtype, int64_t amount, uint8_t out32[32]) {
    size_t pos = 4;
    uint64_t nvin, nvout, sl;
    if (txlen < 10 || !read_compact(tx, txlen, &pos, &nvin)) return 0;
    if (nin >= nvin) return 0;
    size_t vin_start = pos;
    sha256_ctx cp, cs;
    sha256_init(&cp); sha256_init(&cs);
    const uint8_t *my_outpoint = NULL, *my_seq = NULL;
    for (uint64_t i = 0; i < nvin; i++) {
        if (pos + 36 > txlen) return 0;
        const uint8_t *op = tx + pos;
        pos += 36;
        if (!read_compact(tx, txlen, &pos, &sl) || pos + sl + 4 > txlen) return 0;
        pos += sl;
        sha256_update(&cp, op, 36);
        sha256_update(&cs, tx + pos, 4);
        if (i == nin) { my_outpoint = op; my_seq = tx + pos; }
        pos += 4;
    }
    (void)vin_start;
    if (!read_compact(tx, txlen, &pos, &nvout)) return 0;
    size_t outs_start = pos, single_start = 0, single_end = 0;
    for (uint64_t i = 0; i < nvout; i++) {
        size_t st = pos;
        if (pos + 8 > txlen) return 0;
        pos += 8;
        if (!read_compact(tx, txlen, &pos, &sl) || pos + sl > txlen) return 0;
        pos += sl;
        if (i == nin) { single_start = st; single_end = pos; }
    }
    size_t outs_end = pos;
    if (pos + 4 != txlen) return 0;
    uint8_t hp[32] = {0}, hs[32] = {0}, ho[32] = {0}, t[32];
    unsigned base = hashtype & 0x1f;
    int anyone = (hashtype & 0x80) != 0;
    if (!anyone) { sha256_final(&cp, t); orc_sha256(t, 32, hp); }
    if (!anyone && base != 2 && base != 3) { sha256_final(&cs, t); orc_sha256(t, 32, hs); }
    if (base != 2 && base != 3) orc_sha256d(tx + outs_start, outs_end - outs_start, ho);
    else if (base == 3 && nin < nvout) orc_sha256d(tx + single_start, single_end - single_start, ho);
    sha256_ctx c;
    sha256_init(&c);
    sha256_update(&c, tx, 4);
    sha256_update(&c, hp, 32);
    sha256_update(&c, hs, 32);
    sha256_update(&c, my_outpoint, 36);
    uint8_t cs_len[9];
    size_t cl;
    if (sc_len < 253) { cs_len[0] = (uint8_t)sc_len; cl = 1; }
    else if (sc_len <= 0xffff) { cs_len[0] = 253; cs_len[1] = (uint8_t)sc_len; cs_len[2] = (uint8_t)(sc_len >> 8); cl = 3; }
    else { cs_len[0] = 254; for (int i = 0; i < 4; i++) cs_len[1 + i] = (uint8_t)(sc_len >> (8 * i)); cl = 5; }
    sha256_update(&c, cs_len, cl);
    sha256_update(&c, script_code, sc_len);
    uint8_t le[8];
    for (int i = 0; i < 8; i++) le[i] = (uint8_t)((uint64_t)amount >> (8 * i));
    sha256_update(&c, le, 8);
    sha256_update(&c, my_seq, 4);
    sha256_update(&c, ho, 32);
    sha256_update(&c, tx + txlen - 4, 4);
    for (int i = 0; i < 4; i++) le[i] = (uint8_t)(hashtype >> (8 * i));
    sha256_update(&c, le, 4);
    sha256_final(&c, t);
    orc_sha256(t, 32, out32);
    return 1;
}

/* ============================================================== P2PKH input checks ========== */
/* IsValidSignatureEncoding (src/script/interpreter.cpp:107-174): strict DER + one hashtype byte */
int orc_is_valid_signature_encoding(const uint8_t *sig, size_t len) {
    if (len < 9 || len > 73) return 0;
    if (sig[0] != 0x30) return 0;
    if (sig[1] != len - 3) return 0;
    unsigned lenR = sig[3];
    if (5 + lenR >= len) return 0;
    unsigned lenS = sig[5 + lenR];
    if ((size_t)(lenR + lenS + 7) != len) return 0;
    if (sig[2] != 0x02) return 0;
    if (lenR == 0) return 0;
    if (sig[4] & 0x80) return 0;
    if (lenR > 1 && sig[4] == 0x00 && !(sig[5] & 0x80)) return 0;
    if (sig[lenR + 4] != 0x02) return 0;
    if (lenS == 0) return 0;
    if (sig[lenR + 6] & 0x80) return 0;
    if (lenS > 1 && sig[lenR + 6] == 0x00 && !(sig[lenR + 7] & 0x80)) return 0;
    return 1;
}
/* same scan as orc_block_index, also listing every non-coinbase input: byte offset of its outpoint, its
 * transaction and its ordinal within the transaction (CTransaction::IsCoinBase: one input with a null prevout) */
long orc_block_index_inputs(const uint8_t *blk, size_t len, int legacy, uint32_t *tx_off, size_t max_tx, uint32_t *in_off,
                            uint32_t *in_tx, uint32_t *in_n, size_t max_inputs, long *ntx_out) {
    long ntx = orc_block_index(blk, len, legacy, tx_off, max_tx);
    if (ntx < 0) return ntx;
    if (ntx_out) *ntx_out = ntx;
    size_t cnt = 0;
    for (long t = 0; t < ntx; t++) {
        size_t pos = tx_off[t] + 4;
        uint64_t nin, sl;
        read_compact(blk, len, &pos, &nin);
        int coinbase = 0;
        if (nin == 1) {
            coinbase = 1;
            for (int b = 0; b < 32; b++) if (blk[pos + b]) coinbase = 0;
            if (ld64(blk + pos + 28) >> 32 != 0xFFFFFFFFu) coinbase = 0;
        }
        for (uint64_t i = 0; i < nin; i++) {
            if (!coinbase) {
                if (cnt >= max_inputs) return -2;
                in_off[cnt] = (uint32_t)pos; in_tx[cnt] = (uint32_t)t; in_n[cnt] = (uint32_t)i;
                cnt++;
            }
            pos += 36;
            read_compact(blk, len, &pos, &sl);
            pos += sl + 4;
        }
    }
    return (long)cnt;
}
#define ORC_SCRIPT_FLAGS_POSTFORK ((1u << 1) | (1u << 2) | (1u << 3) | (1u << 14) | (1u << 16))
typedef struct {
    const uint8_t *blk; const uint32_t *tx_off, *in_off, *in_tx, *in_n; const uint8_t *spk; const uint32_t *spk_off;
    const int64_t *amounts; uint8_t *err;
} pk_arg;
static uint8_t p2pkh_one(const uint8_t *blk, const uint32_t *tx_off, uint32_t in_off, uint32_t t, uint32_t nin_idx,
                         const uint8_t *spk, size_t spk_len, int64_t amount) {
    if (spk_len != 25 || spk[0] != 0x76 || spk[1] != 0xa9 || spk[2] != 0x14 || spk[23] != 0x88 || spk[24] != 0xac) return 255;
    const uint8_t *tx = blk + tx_off[t];
    size_t txlen = tx_off[t + 1] - tx_off[t];
    size_t pos = in_off + 36 - tx_off[t];
    uint64_t sl;
    if (!read_compact(tx, txlen, &pos, &sl)) return 255;
    const uint8_t *ss = tx + pos;
    /* scriptSig must be exactly two direct pushes (opcodes 0x00..0x4b) */
    if (sl < 2) return 255;
    size_t siglen = ss[0];
    if (siglen > 75 || 1 + siglen >= sl) return 255;
    size_t publen = ss[1 + siglen];
    if (publen == 0 || publen > 75 || 1 + siglen + 1 + publen != sl) return 255;
    const uint8_t *sig = ss + 1, *pub = ss + 2 + siglen;
    uint8_t h160[20];
    orc_hash160(pub, publen, h160);
    if (memcmp(h160, spk + 3, 20) != 0) return 11;                       /* SCRIPT_ERR_EQUALVERIFY */
    uint8_t rs[64];
    memset(rs, 0, 64);
    if (siglen) { /* CheckSignatureEncoding (:221-254) */
        if (!orc_is_valid_signature_encoding(sig, siglen)) return 23;   /* SIG_DER */
        if (!orc_check_low_s(sig, siglen - 1)) return 26;                /* SIG_HIGH_S */
        uint8_t ht = sig[siglen - 1];
        uint8_t base = ht & ~(0x80 | 0x40);
        if (base < 1 || base > 3) return 22;                             /* SIG_HASHTYPE */
        if (!(ht & 0x40)) return 36;                                     /* MUST_USE_FORKID */
    }
    /* CheckPubKeyEncoding (:256-269, :62-82) */
    if (publen < 33) return 28;
    if (pub[0] == 0x04) { if (publen != 65) return 28; }
    else if (pub[0] == 0x02 || pub[0] == 0x03) { if (publen != 33) return 28; }
    else return 28;
    if (!siglen) return 2;                                               /* empty signature: FALSE on the stack -> EVAL_FALSE */
    uint8_t sh[32];
    if (!orc_sighash_forkid(spk, 25, tx, txlen, nin_idx, sig[siglen - 1], amount, sh)) return 1;
    if (!orc_pubkey_verify(sh, pub, publen, sig, siglen - 1)) return 31; /* SIG_NULLFAIL */
    return 0;
}
static void pk_run(size_t lo, size_t hi, void *p) {
    pk_arg *a = (pk_arg *)p;
    for (size_t i = lo; i < hi; i++)
        a->err[i] = p2pkh_one(a->blk, a->tx_off, a->in_off[i], a->in_tx[i], a->in_n[i], a->spk + a->spk_off[i],
                              a->spk_off[i + 1] - a->spk_off[i], a->amounts[i]);
}
long orc_block_verify_p2pkh(const uint8_t *blk, size_t len, int legacy, const uint8_t *spk, const uint32_t *spk_off,
                            const int64_t *amounts, uint32_t flags, int nthreads, uint8_t *err, size_t max_inputs) {
    if ((flags & ORC_SCRIPT_FLAGS_POSTFORK) != ORC_SCRIPT_FLAGS_POSTFORK || (flags & (1u << 17))) return -3;
    size_t max_tx = len / 60 + 2;
    uint32_t *tx_off = (uint32_t *)malloc(4 * (max_tx + 1));
    uint32_t *in_off = (uint32_t *)malloc(4 * max_inputs), *in_tx = (uint32_t *)malloc(4 * max_inputs),
             *in_n = (uint32_t *)malloc(4 * max_inputs);
    long ntx = 0;
    long n = orc_block_index_inputs(blk, len, legacy, tx_off, max_tx, in_off, in_tx, in_n, max_inputs, &ntx);
    if (n > 0) {
        pk_arg a = {blk, tx_off, in_off, in_tx, in_n, spk, spk_off, amounts, err};
        pthread_once(&g_tab_once, build_g_tab);
        parallel_for((size_t)n, nthreads, pk_run, &a);
    }
    free(tx_off); free(in_off); free(in_tx); free(in_n);
    return n;
}

/* ============================================================== synthetic chain generator ==== */
/* Transactions of one block of the synthetic chain: tx j spends output 0 of prev_txid[j]
 * (value prev_amt[j], locked to key prev_key[j]) and pays value-1000 to key (prev_key[j]+1) % nkeys; 1-in/1-out
 * P2PKH, SIGHASH_ALL|FORKID, RFC 6979 signatures — the same shape tests/synth.py builds in Python, at C speed.
 * On return prev_* describe the new outputs (for the next block) and out holds the serialized transactions back to
 * back; returns their total size. */
typedef struct {
    const uint8_t *seed; size_t seedlen, key_base, nkeys; uint8_t *keys, *pubs, *spks;
    uint8_t *prev_txid; uint32_t *prev_key; int64_t *prev_amt; uint8_t *slots; uint32_t *lens;
} gc_arg;
static void gc_keys(size_t lo, size_t hi, void *p) {
    gc_arg *a = (gc_arg *)p;
    for (size_t j = lo; j < hi; j++) {
        uint8_t *k = a->keys + 32 * j;
        tagged(a->seed, a->seedlen, "key", a->key_base + j, 0, k);
        u256 d;
        sc_set_b32(&d, k);
        if (u256_is_zero(&d)) d.v[0] = 1;
        u256_to_be(k, &d);
        size_t l;
        orc_pubkey_create(k, 1, a->pubs + 33 * j, &l);
        uint8_t *spk = a->spks + 25 * j;
        spk[0] = 0x76; spk[1] = 0xa9; spk[2] = 0x14; spk[23] = 0x88; spk[24] = 0xac;
        orc_hash160(a->pubs + 33 * j, 33, spk + 3);
    }
}
static void gc_txs(size_t lo, size_t hi, void *p) {
    gc_arg *a = (gc_arg *)p;
    for (size_t j = lo; j < hi; j++) {
        uint32_t kj = a->prev_key[j], nk = (uint32_t)((kj + 1) % a->nkeys);
        int64_t amt = a->prev_amt[j], outv = amt - 1000;
        uint8_t tx[256], un[128];
        /* unsigned form for the sighash: empty scriptSig */
        size_t q = 0;
        un[q++] = 2; un[q++] = 0; un[q++] = 0; un[q++] = 0; un[q++] = 1;
        memcpy(un + q, a->prev_txid + 32 * j, 32); q += 32;
        memset(un + q, 0, 4); q += 4;
        un[q++] = 0;
        memset(un + q, 0xff, 4); q += 4;
        un[q++] = 1;
        for (int b = 0; b < 8; b++) un[q++] = (uint8_t)((uint64_t)outv >> (8 * b));
        un[q++] = 25;
        memcpy(un + q, a->spks + 25 * nk, 25); q += 25;
        memset(un + q, 0, 4); q += 4;
        uint8_t sh[32], sig64[64], der[80];
        orc_sighash_forkid(a->spks + 25 * kj, 25, un, q, 0, 0x41, amt, sh);
        orc_sign_rfc6979(a->keys + 32 * kj, sh, sig64);
        size_t dl = sizeof der;
        orc_sig_to_der(sig64, der, &dl);
        size_t w = 0;
        memcpy(tx, un, 41); w = 41; /* version, count, outpoint */
        tx[w++] = (uint8_t)(1 + dl + 1 + 1 + 33);
        tx[w++] = (uint8_t)(dl + 1);
        memcpy(tx + w, der, dl); w += dl;
        tx[w++] = 0x41;
        tx[w++] = 33;
        memcpy(tx + w, a->pubs + 33 * kj, 33); w += 33;
        memcpy(tx + w, un + 42, q - 42); w += q - 42; /* sequence, outputs, locktime */
        memcpy(a->slots + 256 * j, tx, w);
        a->lens[j] = (uint32_t)w;
        orc_sha256d(tx, w, a->prev_txid + 32 * j);
        a->prev_key[j] = nk;
        a->prev_amt[j] = outv;
    }
}
size_t orc_gen_chain_txs(const uint8_t *seed, size_t seedlen, size_t key_base, size_t nkeys, size_t ntx, uint8_t *prev_txid,
                         uint32_t *prev_key, int64_t *prev_amt, int nthreads, uint8_t *out, size_t out_cap) {
    if (seedlen > 200) seedlen = 200;
    gc_arg a = {seed, seedlen, key_base, nkeys, NULL, NULL, NULL, prev_txid, prev_key, prev_amt, NULL, NULL};
    a.keys = (uint8_t *)malloc(32 * nkeys); a.pubs = (uint8_t *)malloc(33 * nkeys); a.spks = (uint8_t *)malloc(25 * nkeys);
    a.slots = (uint8_t *)malloc(256 * ntx); a.lens = (uint32_t *)malloc(4 * ntx);
    pthread_once(&g_tab_once, build_g_tab);
    parallel_for(nkeys, nthreads, gc_keys, &a);
    parallel_for(ntx, nthreads, gc_txs, &a);
    size_t tot = 0;
    for (size_t j = 0; j < ntx; j++) {
        if (tot + a.lens[j] > out_cap) { tot = 0; break; }
        memcpy(out + tot, a.slots + 256 * j, a.lens[j]);
        tot += a.lens[j];
    }
    free(a.keys); free(a.pubs); free(a.spks); free(a.slots); free(a.lens);
    return tot;
}
