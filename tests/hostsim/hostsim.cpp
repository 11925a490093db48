// tests/hostsim/hostsim.cpp — TEST-ONLY host build of the device arithmetic headers.
//
// The CI container has no GPU.  The device headers (bitcoincashplus_b200/csrc/*.cuh) compile as
// plain C++ with portable stand-ins for the PTX limb primitives, so the field / scalar / group law,
// the GLV split, the window recoding and the whole verdict logic can be checked against the oracle
// here before GPU time is spent.  Nothing in the product links this file: libbcp_b200.so is built
// by nvcc from the .cu files only and has no host execution path.
#include <cstring>
#include <thread>
#include <vector>
#include "../../bitcoincashplus_b200/csrc/ecdsa.cuh"
#include "../../bitcoincashplus_b200/csrc/sha256.cuh"
#include "../../bitcoincashplus_b200/csrc/blake2b.cuh"
#include "../../bitcoincashplus_b200/csrc/ripemd160.cuh"
#include "../../bitcoincashplus_b200/csrc/script.cuh"
#include "../../bitcoincashplus_b200/csrc/pow.cuh"

static void fe_from_be(fe &r, const uint8_t *b) { fe_set_b32(r, b); }
static void fe_to_be(uint8_t *b, fe a) {
    fe_normalize(a);
    for (int i = 0; i < 8; i++) {
        b[4 * (7 - i)] = a.v[i] >> 24; b[4 * (7 - i) + 1] = a.v[i] >> 16; b[4 * (7 - i) + 2] = a.v[i] >> 8; b[4 * (7 - i) + 3] = a.v[i];
    }
}
static void words_from_be(uint32_t *v, const uint8_t *b) {
    for (int i = 0; i < 8; i++) {
        const uint8_t *q = b + 4 * (7 - i);
        v[i] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
    }
}
static void words_to_be(uint8_t *b, const uint32_t *v) {
    for (int i = 0; i < 8; i++) {
        b[4 * (7 - i)] = v[i] >> 24; b[4 * (7 - i) + 1] = v[i] >> 16; b[4 * (7 - i) + 2] = v[i] >> 8; b[4 * (7 - i) + 3] = v[i];
    }
}

#include <atomic>
static std::atomic<long> hs_quick_redone{0}, hs_quick_mismatch{0};
extern "C" {
long hs_quick_counters(int which) { return which ? hs_quick_mismatch.load() : hs_quick_redone.load(); }
// the quick (branch-free, carry-counting) field forms: op 0 mul, 1 add, 2 sub, 3 sqr(a), 4 neg(a), 9 8*a; returns `bad`,
// out32 = the canonical value of whatever the form produced (only meaningful when bad == 0)
uint32_t hs_fe_op_q(int op, const uint8_t *a32, const uint8_t *b32, uint8_t *out32) {
    fe a, b, r;
    uint32_t bad = 0;
    words_from_be(a.v, a32);
    words_from_be(b.v, b32);
    switch (op) {
        case 0: fe_mul_q(r, a, b, bad); break;
        case 1: fe_add_q(r, a, b, bad); break;
        case 2: fe_sub_q(r, a, b, bad); break;
        case 3: fe_sqr_q(r, a, bad); break;
        case 4: fe_neg_q(r, a, bad); break;
        default: fe_mul8_q(r, a, bad); break;
    }
    fe_to_be(out32, r);
    return bad;
}
// fe_reduce_wide_q on a chosen 512-bit input (big-endian)
uint32_t hs_fe_reduce_q(const uint8_t *t64_be, uint8_t *out32) {
    uint32_t w[16], bad = 0;
    for (int i = 0; i < 16; i++) {
        const uint8_t *q = t64_be + 60 - 4 * i;
        w[i] = (uint32_t)q[0] << 24 | (uint32_t)q[1] << 16 | (uint32_t)q[2] << 8 | q[3];
    }
    fe r;
    fe_reduce_wide_q(r, w, bad);
    fe_to_be(out32, r);
    return bad;
}
// op: 0 mul, 1 add, 2 sub, 3 sqr(a), 4 neg(a), 5 inv(a), 6 sqrt candidate(a); inputs are ANY 256-bit values
// (weakly reduced representation), outputs canonical
void hs_fe_op(int op, const uint8_t *a32, const uint8_t *b32, uint8_t *out32) {
    fe a, b, r;
    words_from_be(a.v, a32);
    words_from_be(b.v, b32);
    switch (op) {
        case 0: fe_mul(r, a, b); break;
        case 1: fe_add(r, a, b); break;
        case 2: fe_sub(r, a, b); break;
        case 3: fe_sqr(r, a); break;
        case 4: fe_neg(r, a); break;
        case 5: fe_inv(r, a); break;
        case 7: fe_add_full(r, a, b); break;
        case 8: fe_sub_full(r, a, b); break;
        default: fe_sqrt_candidate(r, a); break;
    }
    fe_to_be(out32, r);
}
// fe_reduce_wide on a chosen 512-bit input (big-endian), with and without the short-cut second fold
void hs_fe_reduce(const uint8_t *t64_be, int short_fold, uint8_t *out32) {
    uint32_t w[16];
    for (int i = 0; i < 16; i++) {
        const uint8_t *q = t64_be + 60 - 4 * i;
        w[i] = (uint32_t)q[0] << 24 | (uint32_t)q[1] << 16 | (uint32_t)q[2] << 8 | q[3];
    }
    fe r;
    if (short_fold) fe_reduce_wide_t<true>(r, w);
    else fe_reduce_wide_t<false>(r, w);
    fe_to_be(out32, r);
}
// 512-bit product via the Karatsuba variant (an experiment kept in fe.cuh), big-endian in/out
void hs_mul_wide_karatsuba(const uint8_t *a32, const uint8_t *b32, uint8_t *out64) {
    uint32_t a[8], b[8], r[16], r2[16];
    words_from_be(a, a32);
    words_from_be(b, b32);
    mul_wide_karatsuba(r, a, b);
    mul_wide(r2, a, b);
    for (int i = 0; i < 16; i++) {
        if (r[i] != r2[i]) r[i] = ~r2[i]; /* make any disagreement with the schoolbook product visible */
        out64[4 * (15 - i)] = r[i] >> 24; out64[4 * (15 - i) + 1] = r[i] >> 16; out64[4 * (15 - i) + 2] = r[i] >> 8; out64[4 * (15 - i) + 3] = r[i];
    }
}
int hs_fe_is_zero(const uint8_t *a32) {
    fe a;
    words_from_be(a.v, a32);
    return fe_is_zero(a);
}
// op: 0 mul, 1 add, 2 sub, 3 inverse(a), 4 negate(a); inputs must be < n
void hs_sc_op(int op, const uint8_t *a32, const uint8_t *b32, uint8_t *out32) {
    scalar a, b, r;
    words_from_be(a.v, a32);
    words_from_be(b.v, b32);
    switch (op) {
        case 0: sc_mul(r, a, b); break;
        case 1: sc_add(r, a, b); break;
        case 2: sc_sub(r, a, b); break;
        case 3: sc_inverse(r, a); break;
        case 5: sc_inverse_divsteps(r, a); break;
        default: sc_negate(r, a); break;
    }
    words_to_be(out32, r.v);
}
void hs_sc_reduce512(const uint8_t *t64_be, uint8_t *out32) {
    uint32_t t[16];
    for (int i = 0; i < 16; i++) {
        const uint8_t *q = t64_be + 4 * (15 - i);
        t[i] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
    }
    scalar r;
    sc_reduce512(r, t);
    words_to_be(out32, r.v);
}
// split: out k1abs/k2abs as 5 little-endian words each, neg flags
void hs_split(const uint8_t *k32, uint32_t *k1abs, int *neg1, uint32_t *k2abs, int *neg2) {
    scalar k;
    words_from_be(k.v, k32);
    bool n1, n2;
    sc_split_lambda(k1abs, n1, k2abs, n2, k);
    *neg1 = n1; *neg2 = n2;
}
// digits of a recoded magnitude: out[34]
void hs_recode(const uint32_t *kabs5, int *digits34) {
    uint32_t k[5];
    memcpy(k, kabs5, 20);
    recode_w4(k);
    for (int w = 0; w < 34; w++) digits34[w] = digit_w4(k, w);
}
int hs_pubkey_parse(const uint8_t *pub65, uint8_t *xy64) {
    fe x, y;
    if (!pubkey_parse(x, y, pub65)) return 0;
    fe_to_be(xy64, x);
    fe_to_be(xy64 + 32, y);
    return 1;
}
void hs_gcomb_entry(int j, uint32_t d, uint32_t *out16) { gcomb_entry(out16, j, d); }

// Full table, built incrementally with the same device group law (entry d = entry d-1 + base_j)
void hs_gcomb_build(uint32_t *gcomb) {
    std::vector<std::thread> th;
    for (int j = 0; j < BCP_GCOMB_WINDOWS; j++) {
        th.emplace_back([=] {
            uint32_t b16[16];
            gcomb_entry(b16, j, 1);
            fe bx, by;
            for (int i = 0; i < 8; i++) { bx.v[i] = b16[i]; by.v[i] = b16[8 + i]; }
            gej acc;
            acc.inf = true;
            fe_set_zero(acc.x); fe_set_zero(acc.y); fe_set_zero(acc.z);
            uint32_t *tj = gcomb + (size_t)j * BCP_GCOMB_ENTRIES * 16;
            memset(tj, 0, 64);
            for (uint32_t d = 1; d < BCP_GCOMB_ENTRIES; d++) {
                gej_add_ge(acc, bx, by, nullptr);
                fe ax, ay;
                gej_to_affine(ax, ay, acc);
                for (int i = 0; i < 8; i++) { tj[d * 16 + i] = ax.v[i]; tj[d * 16 + 8 + i] = ay.v[i]; }
            }
        });
    }
    for (auto &t : th) t.join();
}

// Emulates k_ecdsa_prep (T logical threads x K signatures) followed by k_ecdsa_point.
void hs_ecdsa_verify_batch(const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64, size_t n, uint32_t flags,
                           const uint32_t *gcomb, int nthreads, uint8_t *ok) {
    const int K = 8;
    std::vector<uint32_t> rec(n * BCP_REC_WORDS);
    size_t T = (n + K - 1) / K;
    T = ((T + 127) / 128) * 128; // grid-rounded like the launch
    auto prep = [&](size_t lo, size_t hi) {
        for (size_t t = lo; t < hi; t++) {
            if (t >= n) continue;
            sig_scalars g[K];
            scalar pre[K], acc, inv;
            for (int i = 0; i < 8; i++) acc.v[i] = (i == 0);
            int cnt = 0;
            for (int j = 0; j < K; j++) {
                size_t i = t + (size_t)j * T;
                if (i >= n) break;
                sig_load(g[j], msg32 + 32 * i, sig64 + 64 * i, flags);
                pre[j] = acc;
                sc_mul(acc, acc, g[j].s);
                cnt++;
            }
            sc_inverse_divsteps(inv, acc); /* what k_ecdsa_prep calls */
            for (int j = cnt - 1; j >= 0; j--) {
                size_t i = t + (size_t)j * T;
                scalar sinv;
                sc_mul(sinv, inv, pre[j]);
                sc_mul(inv, inv, g[j].s);
                sig_finish(&rec[i * BCP_REC_WORDS], g[j], sinv);
            }
        }
    };
    auto point = [&](size_t lo, size_t hi) {
        std::vector<uint32_t> smem(BCP_QTAB_WORDS);
        for (size_t i = lo; i < hi; i++) {
            qtable tab;
            tab.base = smem.data();
            tab.stride = 1;
            // like k_ecdsa_point: the walk over the quick field forms, repeated with the exact ones when a fold was left
            // unfinished; both walks must give the same verdict whenever the quick one is trusted
            uint32_t bad = 0;
            bool v = ecdsa_point_phase<qtable, false, fe_quick>(&rec[i * BCP_REC_WORDS], pub65 + 65 * i, sig64 + 64 * i, tab, gcomb, 0, nullptr, &bad);
            const bool vx = ecdsa_point_phase<qtable, false, fe_exact>(&rec[i * BCP_REC_WORDS], pub65 + 65 * i, sig64 + 64 * i, tab, gcomb);
            if (bad) { hs_quick_redone++; v = vx; }
            else if (v != vx) hs_quick_mismatch++;
            // the other bracketing of the same walk (ecdsa_point_phase_v2), over both sets of field forms
            uint32_t bad2 = 0;
            const bool v2q = ecdsa_point_phase_v2<qtable, false, fe_quick>(&rec[i * BCP_REC_WORDS], pub65 + 65 * i, sig64 + 64 * i, tab, gcomb, 0, nullptr, &bad2);
            const bool v2x = ecdsa_point_phase_v2<qtable, false, fe_exact>(&rec[i * BCP_REC_WORDS], pub65 + 65 * i, sig64 + 64 * i, tab, gcomb);
            if (v2x != vx || (!bad2 && v2q != vx)) hs_quick_mismatch++;
            ok[i] = v ? 1 : 0;
        }
    };
    auto run = [&](size_t total, auto fn) {
        std::vector<std::thread> th;
        size_t chunk = (total + nthreads - 1) / nthreads;
        for (int t = 0; t < nthreads; t++) {
            size_t lo = std::min(total, t * chunk), hi = std::min(total, lo + chunk);
            if (lo < hi) th.emplace_back(fn, lo, hi);
        }
        for (auto &x : th) x.join();
    };
    run(T, prep);
    run(n, point);
}

// ---- hashing device functions
void hs_sha256d(const uint8_t *data, size_t len, uint8_t *out32) { dev_sha256d_bytes(out32, data, len); }
void hs_sha256(const uint8_t *data, size_t len, uint8_t *out32) { dev_sha256_bytes(out32, data, len); }
void hs_hash160(const uint8_t *data, size_t len, uint8_t *out20) {
    uint32_t s[8], r[5];
    dev_sha256_words(s, data, len);
    ripemd160_of_sha256(r, s);
    for (int i = 0; i < 5; i++)
        for (int b = 0; b < 4; b++) out20[4 * i + b] = (uint8_t)(r[i] >> (8 * b));
}
// Emulates k_tx_midstate + k_p2pkh_extract: status per input and the ECDSA triples
void hs_block_p2pkh_extract(const uint8_t *blk, const uint32_t *tx_off, size_t ntx, const uint32_t *in_off,
                            const uint32_t *in_tx, const uint32_t *in_n, size_t nin, const uint8_t *spk,
                            const uint32_t *spk_off, const int64_t *amounts, uint8_t *status, uint8_t *msg32, uint8_t *pub65,
                            uint8_t *sig64) {
    std::vector<uint8_t> mid(ntx * BCP_TXMID_BYTES);
    for (size_t t = 0; t < ntx; t++) tx_midstate(&mid[t * BCP_TXMID_BYTES], blk + tx_off[t], tx_off[t + 1] - tx_off[t]);
    for (size_t i = 0; i < nin; i++) {
        uint32_t t = in_tx[i];
        status[i] = (uint8_t)p2pkh_extract_one(blk + tx_off[t], tx_off[t + 1] - tx_off[t], in_off[i] - tx_off[t], in_n[i],
                                               spk + spk_off[i], spk_off[i + 1] - spk_off[i], amounts[i],
                                               &mid[t * BCP_TXMID_BYTES], msg32 + 32 * i, pub65 + 65 * i, sig64 + 64 * i);
    }
}
// pow.cuh: CheckProofOfWork on the device limbs; hash / limit as uint256 memory bytes
int hs_pow_check(const uint8_t *hash32, uint32_t nbits, const uint8_t *limit32, uint8_t *target32, int *neg, int *ovf) {
    uint32_t h[8], l[8], t[8];
    memcpy(h, hash32, 32);
    memcpy(l, limit32, 32);
    bool n, o;
    pow_set_compact(nbits, t, n, o);
    memcpy(target32, t, 32);
    *neg = n;
    *ovf = o;
    return pow_check(h, nbits, l) ? 1 : 0;
}
void hs_blake2b_final_block(const uint64_t *h_mid, const uint8_t *blk128, uint64_t t0, uint64_t *h_out) {
    uint64_t h[8], m[16];
    for (int i = 0; i < 8; i++) h[i] = h_mid[i];
    for (int i = 0; i < 16; i++) {
        uint64_t v = 0;
        for (int b = 0; b < 8; b++) v |= (uint64_t)blk128[8 * i + b] << (8 * b);
        m[i] = v;
    }
    blake2b_compress(h, m, t0, true);
    for (int i = 0; i < 8; i++) h_out[i] = h[i];
}
}
