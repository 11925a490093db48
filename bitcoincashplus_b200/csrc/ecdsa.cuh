// ecdsa.cuh — the two device phases of one ECDSA verification.
//
// Together they compute exactly the verdict of the reference chain
//   secp256k1_ec_pubkey_parse (src/secp256k1/src/secp256k1.c:150-163, eckey_impl.h:17-35)
//   secp256k1_ecdsa_signature_parse_compact (:228-247)  [+ _normalize (:274-291) when BCP_ECDSA_NORMALIZE_S]
//   secp256k1_ecdsa_verify (:293-308) -> secp256k1_ecdsa_sig_verify (ecdsa_impl.h:201-269)
// i.e. accept iff  pubkey parses,  r,s in [1,n-1],  s <= n/2,  R = (m/s)G + (r/s)Q != infinity  and
// x(R) == r or (r < p-n and x(R) == r+n)  (the comparison is done projectively, r*Z^2 == X, like
// gej_eq_x_var group_impl.h:246-252, so no field inversion is needed).
//
// How the multiplication is evaluated is NOT the reference's (wNAF Strauss with a 1 MiB G table,
// ecmult_impl.h:286-404): a warp must stay convergent, so
//   phase 1 (prep, one thread per K signatures): range checks, ONE scalar inversion per K signatures
//           (Montgomery's trick; the inversion itself by division steps, scalar.cuh), u1 = m/s, u2 = r/s, lattice
//           split u2 = k1 + k2*lambda, signed 4-bit window recoding.  Output: a 80-byte record per signature in scratch memory.
//   phase 2 (point, one thread per signature; one rolled loop, see ecdsa_point_phase_v2): decompress Q, build {1..8}Q on a common Z without any
//           inversion (z-ratio rescaling, the idea of ge_globalz_set_table_gej group_impl.h:173-194),
//           34 windows of 4 doublings + 2 mixed additions (Q and lambda*Q share the table: lambda*(x,y) =
//           (beta*x, y)), then u1*G from a fixed-position comb table (BCP_GCOMB_WINDOWS windows of BCP_GCOMB_BITS
//           bits, no doublings), then the projective x comparison.
#pragma once
#include "ge.cuh"
#include "scalar.cuh"

#ifndef BCP_STEP_SYNC
#define BCP_STEP_SYNC() ((void)0)
#endif
#ifndef BCP_STEP_SYNC_V2
#define BCP_STEP_SYNC_V2() ((void)0)
#endif

#define BCP_ECDSA_NORMALIZE_S 1u /* CPubKey::Verify semantics: s := min(s, n-s) instead of rejecting high s */

#define BCP_REC_WORDS 20 /* u1[8] k1[5] k2[5] flags pad  = 80 bytes */
#define BCP_REC_VALID 1u
#define BCP_REC_NEG1 2u
#define BCP_REC_NEG2 4u

/* window width of the fixed-base comb: 22 bits = 12 lookups per signature from a 3.2 GB table in HBM (12 windows x
 * 4 M entries x 64 B, generated on the device at start-up).  The table is far larger than the L2, and that does not
 * matter: a lookup is one 64-byte random read per mixed addition (~700 multiplier instructions), other warps cover its
 * latency, and at 48 M signatures/s the reads are 37 GB/s of an HBM that offers 6,500.  Measured on B200, 1 M
 * signatures: 16 bits (64 MiB, L2-resident) 47.09 M/s, 20 bits (832 MB) 47.79, 22 bits 48.07 — each window less is one
 * mixed addition less.  The test-only host build uses 12 to keep its table generation short. */
#ifndef BCP_GCOMB_BITS
#define BCP_GCOMB_BITS 22
#endif
#define BCP_GCOMB_WINDOWS ((256 + BCP_GCOMB_BITS - 1) / BCP_GCOMB_BITS)
#define BCP_GCOMB_ENTRIES (1 << BCP_GCOMB_BITS)
#define BCP_GCOMB_WORDS (BCP_GCOMB_WINDOWS * BCP_GCOMB_ENTRIES * 16) /* 16 words (x,y) per entry; entry 0 unused */

#define BCP_QTAB_ENTRIES 8
#define BCP_QTAB_EWORDS 24 /* per entry: x, y and beta*x (the x of lambda*P), 8 limbs each */
#define BCP_QTAB_WORDS (BCP_QTAB_ENTRIES * BCP_QTAB_EWORDS)

// ---------------------------------------------------------------- phase 1: scalar side
struct sig_scalars {
    scalar r, s, m;
    bool valid;
};
// parse + range rules for one signature (msg is reduced mod n silently, secp256k1.c:303)
BCP_DEV void sig_load(sig_scalars &o, const uint8_t *msg32, const uint8_t *sig64, uint32_t flags) {
    bool ov_r = sc_set_b32(o.r, sig64);
    bool ov_s = sc_set_b32(o.s, sig64 + 32);
    sc_set_b32(o.m, msg32);
    bool ok = !(ov_r || ov_s);                          /* parse_compact overflow => rejected            */
    ok = ok && !sc_is_zero(o.r) && !sc_is_zero(o.s);    /* ecdsa_impl.h:210                              */
    if (sc_is_high(o.s)) {
        if (flags & BCP_ECDSA_NORMALIZE_S) sc_negate(o.s, o.s);
        else ok = false;                                /* secp256k1.c:305                               */
    }
    o.valid = ok;
    if (!ok) { /* keep the batch product invertible */
        BCP_UNROLL
        for (int i = 0; i < 8; i++) o.s.v[i] = (i == 0);
    }
}
// signed 4-bit window recoding of a magnitude < 2^132 held in 5 limbs: after adding 8 to every
// nibble 0..32 (with carries), digit_w = nibble_w - 8 in [-8,7] for w <= 32 and digit_33 = nibble_33 in {0,1}
BCP_DEV void recode_w4(uint32_t *k /*5, in place*/) {
    uint32_t add[5] = {0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u, 0x00000008u};
    addn_cc<5>(k, k, add);
}
BCP_DEV int digit_w4(const uint32_t *k, int w) {
    int nib = (int)((k[w >> 3] >> ((w & 7) * 4)) & 15u);
    return w <= 32 ? nib - 8 : nib;
}
// finish one signature once 1/s is known: write the scratch record
BCP_DEV void sig_finish(uint32_t *rec /*BCP_REC_WORDS*/, const sig_scalars &g, const scalar &sinv) {
    scalar u1, u2;
    sc_mul(u1, g.m, sinv);
    sc_mul(u2, g.r, sinv);
    uint32_t k1[5], k2[5];
    bool n1, n2;
    sc_split_lambda(k1, n1, k2, n2, u2);
    recode_w4(k1);
    recode_w4(k2);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) rec[i] = u1.v[i];
    BCP_UNROLL
    for (int i = 0; i < 5; i++) { rec[8 + i] = k1[i]; rec[13 + i] = k2[i]; }
    rec[18] = (g.valid ? BCP_REC_VALID : 0u) | (n1 ? BCP_REC_NEG1 : 0u) | (n2 ? BCP_REC_NEG2 : 0u);
    rec[19] = 0;
}

// ---------------------------------------------------------------- phase 2: point side
// Per-thread table of 8 affine-like points in a strided word array (shared memory on the GPU:
// word (e*16 + i) of thread t lives at base[(e*16 + i) * stride + t], so a warp reading "entry
// e_lane, limb i" touches 32 different banks whatever the per-lane entries are).
struct qtable {
    uint32_t *base;
    int stride;
    BCP_DEV_M void store(int e, const fe &x, const fe &y) {
        BCP_UNROLL
        for (int i = 0; i < 8; i++) {
            base[(e * BCP_QTAB_EWORDS + i) * stride] = x.v[i];
            base[(e * BCP_QTAB_EWORDS + 8 + i) * stride] = y.v[i];
        }
    }
    BCP_DEV_M void store_xb(int e, const fe &xb) {
        BCP_UNROLL
        for (int i = 0; i < 8; i++) base[(e * BCP_QTAB_EWORDS + 16 + i) * stride] = xb.v[i];
    }
    BCP_DEV_M void load(int e, fe &x, fe &y) const {
        BCP_UNROLL
        for (int i = 0; i < 8; i++) {
            x.v[i] = base[(e * BCP_QTAB_EWORDS + i) * stride];
            y.v[i] = base[(e * BCP_QTAB_EWORDS + 8 + i) * stride];
        }
    }
    BCP_DEV_M void load_b(int e, fe &xb, fe &y) const { /* lambda * entry = (beta * x, y) */
        BCP_UNROLL
        for (int i = 0; i < 8; i++) {
            xb.v[i] = base[(e * BCP_QTAB_EWORDS + 16 + i) * stride];
            y.v[i] = base[(e * BCP_QTAB_EWORDS + 8 + i) * stride];
        }
    }
};

// Same table, one contiguous 768-byte slab per signature in global memory (entry e = 96 bytes at word e*24:
// x, y, beta*x): every lookup is four 16-byte loads of one lane's own entry, i.e. fully used sectors whatever
// entry each lane picks.  Frees the shared memory that capped the point kernel at 12 warps per SM.
struct qtable_packed {
    uint32_t *base; /* this signature's slab, 16-byte aligned */
    BCP_DEV_M void store(int e, const fe &x, const fe &y) {
#if BCP_PTX
        uint4 *p = reinterpret_cast<uint4 *>(base + e * BCP_QTAB_EWORDS);
        p[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]);
        p[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
        p[2] = make_uint4(y.v[0], y.v[1], y.v[2], y.v[3]);
        p[3] = make_uint4(y.v[4], y.v[5], y.v[6], y.v[7]);
#else
        for (int i = 0; i < 8; i++) { base[e * BCP_QTAB_EWORDS + i] = x.v[i]; base[e * BCP_QTAB_EWORDS + 8 + i] = y.v[i]; }
#endif
    }
    BCP_DEV_M void store_xb(int e, const fe &xb) {
#if BCP_PTX
        uint4 *p = reinterpret_cast<uint4 *>(base + e * BCP_QTAB_EWORDS + 16);
        p[0] = make_uint4(xb.v[0], xb.v[1], xb.v[2], xb.v[3]);
        p[1] = make_uint4(xb.v[4], xb.v[5], xb.v[6], xb.v[7]);
#else
        for (int i = 0; i < 8; i++) base[e * BCP_QTAB_EWORDS + 16 + i] = xb.v[i];
#endif
    }
    BCP_DEV_M void load(int e, fe &x, fe &y) const { load_at(e, 0, x, y); }
    BCP_DEV_M void load_b(int e, fe &xb, fe &y) const { load_at(e, 16, xb, y); } /* lambda * entry = (beta * x, y) */
    BCP_DEV_M void load_at(int e, int xo, fe &x, fe &y) const {
#if BCP_PTX
        const uint4 *px = reinterpret_cast<const uint4 *>(base + e * BCP_QTAB_EWORDS + xo);
        const uint4 *py = reinterpret_cast<const uint4 *>(base + e * BCP_QTAB_EWORDS + 8);
        uint4 a = px[0], b = px[1], c = py[0], d = py[1];
        x.v[0] = a.x; x.v[1] = a.y; x.v[2] = a.z; x.v[3] = a.w; x.v[4] = b.x; x.v[5] = b.y; x.v[6] = b.z; x.v[7] = b.w;
        y.v[0] = c.x; y.v[1] = c.y; y.v[2] = c.z; y.v[3] = c.w; y.v[4] = d.x; y.v[5] = d.y; y.v[6] = d.z; y.v[7] = d.w;
#else
        for (int i = 0; i < 8; i++) { x.v[i] = base[e * BCP_QTAB_EWORDS + xo + i]; y.v[i] = base[e * BCP_QTAB_EWORDS + 8 + i]; }
#endif
    }
};

// The whole point phase for one signature, written as ONE loop over "steps" so that the kernel contains a
// single instance of the doubling and of the mixed addition (the fully unrolled first version was 800 KB of
// SASS and stalled on instruction fetch; profiles/r01_*).  rec = phase-1 record (global memory), gcomb =
// fixed-base comb table.  Step schedule:
//   t in [0, 7)            table build: acc = Q; double, then add Q six times; after each step store acc.(x,y)
//                          as table entry t+1 and remember the z-ratio (Z_{k+1} / Z_k)
//   t == 7                 (before the step) rescale entries 0..6 to the common Z = Z_8 with the products of
//                          the later ratios — no inversion (idea of ge_globalz_set_table_gej, group_impl.h:173-194;
//                          kQ, k <= 8, is never infinity nor +-Q, so every table addition is generic) — then acc = infinity
//   t in [7, 7 + 34*6)     window w = 33..0: four doublings, add digit(k1,w)*Q, add digit(k2,w)*lambda*Q
//   next step              leave the isomorphic curve: Z *= Z_8
//   last 16 steps          u1*G: one comb lookup + mixed addition per 16-bit window, no doublings
// sync_n = how many threads of the block walk the steps together (what BCP_STEP_SYNC waits for); unused on the host build.
// WANT_POINT: hand back R = u1*G + u2*Q itself (Jacobian) instead of comparing its x with r — the recovery and
// tweak-add entry points are the same walk with other scalars; returns whether R is a finite point of a valid input.
// F: the field operations of the walk (ge.cuh): fe_exact, or fe_quick with the count of unfinished folds in *bad_out — the
// caller then repeats the call with fe_exact (and sync_n = 0: no barrier, the other threads are not coming).
template <class Table, bool WANT_POINT = false, class F = fe_exact>
BCP_DEV bool ecdsa_point_phase(const uint32_t *rec, const uint8_t *pub65, const uint8_t *sig64, Table tab,
                               const uint32_t *gcomb, const int sync_n = 0, gej *out = nullptr, uint32_t *bad_out = nullptr) {
    (void)sync_n;
    uint32_t bad = 0;
    const uint32_t fl = rec[18];
    gej acc;
    // A signature that is already known to be invalid keeps walking (on the generator instead of its own
    // key) so that every thread of the block reaches every BCP_STEP_SYNC; its verdict is forced to false.
    bool valid = (fl & BCP_REC_VALID) != 0;
    valid = pubkey_parse(acc.x, acc.y, pub65) && valid;
    if (!valid) {
        BCP_UNROLL
        for (int i = 0; i < 8; i++) { acc.x.v[i] = FE_GX[i]; acc.y.v[i] = FE_GY[i]; }
    }
    fe_set_u32(acc.z, 1);
    acc.inf = false;
    tab.store(0, acc.x, acc.y);
    const bool n1 = fl & BCP_REC_NEG1, n2 = fl & BCP_REC_NEG2;
    fe zr[BCP_QTAB_ENTRIES]; /* zr[k] = Z_{k+1} / Z_k  (1-based multiples); run-time index -> local memory */
    fe zglob;
    fe_set_u32(zglob, 1);
    const int T_TAB = BCP_QTAB_ENTRIES - 1, T_MAIN = 34 * 6, T_ALL = T_TAB + T_MAIN + 1 + BCP_GCOMB_WINDOWS;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int t = 0; t < T_ALL; t++) {
        BCP_STEP_SYNC(); /* keeps the warps of a block in the same code region (instruction-cache locality) */
        bool dbl = false, have = false;
        fe x, y, h;
        if (t < T_TAB) {
            if (t == 0) dbl = true;
            else { tab.load(0, x, y); have = true; }
        } else if (t < T_TAB + T_MAIN) {
            const int u = t - T_TAB;
            if (u == 0) {
                zglob = acc.z;
                fe zs = zr[BCP_QTAB_ENTRIES - 1], beta;
                BCP_UNROLL
                for (int i = 0; i < 8; i++) beta.v[i] = FE_BETA[i];
                /* k = ENTRIES-1 is already on the common Z and only gets its beta*x; the others are rescaled first */
#if BCP_PTX
#pragma unroll 1
#endif
                for (int k = BCP_QTAB_ENTRIES - 1; k >= 0; k--) {
                    fe ex, ey, zs2, zs3;
                    tab.load(k, ex, ey);
                    if (k < BCP_QTAB_ENTRIES - 1) {
                        F::sqr(zs2, zs, bad);
                        F::mul(zs3, zs2, zs, bad);
                        F::mul(ex, ex, zs2, bad);
                        F::mul(ey, ey, zs3, bad);
                        tab.store(k, ex, ey);
                        if (k > 0) F::mul(zs, zs, zr[k], bad);
                    }
                    F::mul(ex, ex, beta, bad); /* x of lambda * (k+1)Q: one product per entry instead of one per use */
                    tab.store_xb(k, ex);
                }
                acc.inf = true;
            }
            const int w = 33 - u / 6, ph = u % 6;
            if (ph < 4) {
                dbl = true;
            } else {
                const bool second = ph == 5;
                const uint32_t word = rec[(second ? 13 : 8) + (w >> 3)];
                int d = (int)((word >> ((w & 7) * 4)) & 15u);
                if (w <= 32) d -= 8;
                if (d != 0) {
                    const bool neg = (d < 0) != (second ? n2 : n1);
                    if (second) tab.load_b((d < 0 ? -d : d) - 1, x, y); /* lambda * (x, y) = (beta * x, y) */
                    else tab.load((d < 0 ? -d : d) - 1, x, y);
                    if (neg) F::neg(y, y, bad);
                    have = true;
                }
            }
        } else if (t == T_TAB + T_MAIN) {
            if (!acc.inf) F::mul(acc.z, acc.z, zglob, bad);
            continue;
        } else {
            const int j = t - (T_TAB + T_MAIN + 1);
            const int bit = j * BCP_GCOMB_BITS;
            const int limb = bit >> 5, sh = bit & 31;
            const uint32_t lo = rec[limb];
            const uint32_t hi = (limb + 1 < 8) ? rec[limb + 1] : 0u;
            const uint32_t d = (sh == 0 ? lo : ((lo >> sh) | (hi << (32 - sh)))) & (BCP_GCOMB_ENTRIES - 1);
            if (d != 0) {
                const uint32_t *e = gcomb + ((size_t)j * BCP_GCOMB_ENTRIES + d) * 16;
#if BCP_PTX
                const uint4 *e4 = reinterpret_cast<const uint4 *>(e);
                uint4 qa = __ldg(e4), qb = __ldg(e4 + 1), qc = __ldg(e4 + 2), qd = __ldg(e4 + 3);
                x.v[0] = qa.x; x.v[1] = qa.y; x.v[2] = qa.z; x.v[3] = qa.w; x.v[4] = qb.x; x.v[5] = qb.y; x.v[6] = qb.z; x.v[7] = qb.w;
                y.v[0] = qc.x; y.v[1] = qc.y; y.v[2] = qc.z; y.v[3] = qc.w; y.v[4] = qd.x; y.v[5] = qd.y; y.v[6] = qd.z; y.v[7] = qd.w;
#else
                for (int i = 0; i < 8; i++) { x.v[i] = e[i]; y.v[i] = e[8 + i]; }
#endif
                have = true;
            }
        }
        if (!dbl) {
            if (!have) continue;
            if (acc.inf) { /* first addition into infinity copies the operand (group_impl.h:417-421) */
                acc.x = x; acc.y = y;
                fe_set_u32(acc.z, 1);
                acc.inf = false;
                continue;
            }
            // mixed addition acc += (x, y): 8M + 3S, with the degenerate branches of gej_add_ge_var (:438-448)
            fe z2, u2, s2, rr, h2, h3, v, tt;
            F::sqr(z2, acc.z, bad);
            F::mul(u2, x, z2, bad);
            F::mul(tt, acc.z, z2, bad);
            F::mul(s2, y, tt, bad);
            F::sub(h, u2, acc.x, bad);
            F::sub(rr, s2, acc.y, bad);
            if (fe_is_zero(h)) {
                if (fe_is_zero(rr)) dbl = true;           /* P + P: fall through to the shared doubling  */
                else { acc.inf = true; continue; }        /* P + (-P)                                       */
            } else {
                F::sqr(h2, h, bad);
                F::mul(h3, h2, h, bad);
                F::mul(v, acc.x, h2, bad);
                F::mul(acc.z, acc.z, h, bad);
                F::sqr(tt, rr, bad);
                F::sub(tt, tt, h3, bad); F::sub(tt, tt, v, bad); F::sub(acc.x, tt, v, bad);
                F::sub(tt, v, acc.x, bad);
                F::mul(tt, tt, rr, bad);
                F::mul(h3, h3, acc.y, bad);
                F::sub(acc.y, tt, h3, bad);
            }
        }
        if (dbl && !acc.inf) gej_double_t<F>(acc, acc, bad);
        if (t < T_TAB) {
            tab.store(t + 1, acc.x, acc.y);
            zr[t + 1] = (t == 0) ? acc.z : h;
        }
    }
    if (bad_out) *bad_out = bad;
    if (WANT_POINT) {
        *out = acc;
        return valid && !acc.inf;
    }
    if (acc.inf || !valid) return false; /* ecdsa_impl.h:220 */

    // x(R) == r (mod n), projectively: r * Z^2 == X, then the r + n < p second chance (:246-266)
    scalar r;
    sc_set_b32(r, sig64); /* already known to be in range */
    fe z2, t1, fr;
    fe_sqr(z2, acc.z);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) fr.v[i] = r.v[i];
    uint32_t pmn[8], n[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { pmn[i] = SC_P_MINUS_N[i]; n[i] = SC_N[i]; }
    const bool second_chance = !u256_ge(r.v, pmn);
    bool ok = false;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int pass = 0; pass < 2; pass++) {
        fe_mul(t1, fr, z2);
        if (fe_equal(t1, acc.x)) ok = true;
        if (ok || !second_chance) break;
        addn_cc<8>(fr.v, r.v, n); /* r + n < p: no overflow */
    }
    return ok;
}

// The same walk with the steps bracketed differently: one iteration = at most one mixed addition FOLLOWED by `nd`
// doublings (an inner loop around the single doubling instance).  The order of the group operations is the same as in
// ecdsa_point_phase (window w: four doublings, add digit(k1)*Q, add digit(k2)*lambda*Q); what changes is the code shape:
// no "is this step a doubling" flag merges with the addition's values, the doublings run with only the accumulator
// live, and the P + P case of an addition becomes one more doubling of the same iteration.
//   iteration 0                 table build: no addition, one doubling (2Q)
//   iterations 1..6             table build: add Q
//   then, for w = 33..0         A: add digit(k1,w)*Q;  B: add digit(k2,w)*lambda*Q, then (w > 0) the four doublings of window w-1
//   one iteration               Z *= Z_8 (leave the isomorphic curve)
//   last BCP_GCOMB_WINDOWS      add the comb entry of u1
template <class Table, bool WANT_POINT = false, class F = fe_exact>
BCP_DEV bool ecdsa_point_phase_v2(const uint32_t *rec, const uint8_t *pub65, const uint8_t *sig64, Table tab,
                                  const uint32_t *gcomb, const int sync_n = 0, gej *out = nullptr, uint32_t *bad_out = nullptr) {
    (void)sync_n;
    uint32_t bad = 0;
    const uint32_t fl = rec[18];
    gej acc;
    bool valid = (fl & BCP_REC_VALID) != 0;
    valid = pubkey_parse(acc.x, acc.y, pub65) && valid;
    if (!valid) {
        BCP_UNROLL
        for (int i = 0; i < 8; i++) { acc.x.v[i] = FE_GX[i]; acc.y.v[i] = FE_GY[i]; }
    }
    fe_set_u32(acc.z, 1);
    acc.inf = false;
    tab.store(0, acc.x, acc.y);
    const bool n1 = fl & BCP_REC_NEG1, n2 = fl & BCP_REC_NEG2;
    fe zr[BCP_QTAB_ENTRIES]; /* zr[k] = Z_{k+1} / Z_k */
    fe zglob;
    fe_set_u32(zglob, 1);
    const int T_TAB = BCP_QTAB_ENTRIES - 1, T_MAIN = 34 * 2, T_ALL = T_TAB + T_MAIN + 1 + BCP_GCOMB_WINDOWS;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int t = 0; t < T_ALL; t++) {
        BCP_STEP_SYNC_V2();
        bool have = false;
        int nd = 0;
        fe x, y, h;
        if (t < T_TAB) {
            if (t == 0) nd = 1;
            else { tab.load(0, x, y); have = true; }
        } else if (t < T_TAB + T_MAIN) {
            const int u = t - T_TAB;
            if (u == 0) {
                zglob = acc.z;
                fe zs = zr[BCP_QTAB_ENTRIES - 1], beta;
                BCP_UNROLL
                for (int i = 0; i < 8; i++) beta.v[i] = FE_BETA[i];
#if BCP_PTX
#pragma unroll 1
#endif
                for (int k = BCP_QTAB_ENTRIES - 1; k >= 0; k--) {
                    fe ex, ey, zs2, zs3;
                    tab.load(k, ex, ey);
                    if (k < BCP_QTAB_ENTRIES - 1) {
                        F::sqr(zs2, zs, bad);
                        F::mul(zs3, zs2, zs, bad);
                        F::mul(ex, ex, zs2, bad);
                        F::mul(ey, ey, zs3, bad);
                        tab.store(k, ex, ey);
                        if (k > 0) F::mul(zs, zs, zr[k], bad);
                    }
                    F::mul(ex, ex, beta, bad);
                    tab.store_xb(k, ex);
                }
                acc.inf = true;
            }
            const int w = 33 - (u >> 1);
            const bool second = (u & 1) != 0;
            if (second && w > 0) nd = 4;
            const uint32_t word = rec[(second ? 13 : 8) + (w >> 3)];
            int d = (int)((word >> ((w & 7) * 4)) & 15u);
            if (w <= 32) d -= 8;
            if (d != 0) {
                const bool neg = (d < 0) != (second ? n2 : n1);
                if (second) tab.load_b((d < 0 ? -d : d) - 1, x, y);
                else tab.load((d < 0 ? -d : d) - 1, x, y);
                if (neg) F::neg(y, y, bad);
                have = true;
            }
        } else if (t == T_TAB + T_MAIN) {
            if (!acc.inf) F::mul(acc.z, acc.z, zglob, bad);
            continue;
        } else {
            const int j = t - (T_TAB + T_MAIN + 1);
            const int bit = j * BCP_GCOMB_BITS;
            const int limb = bit >> 5, sh = bit & 31;
            const uint32_t lo = rec[limb];
            const uint32_t hi = (limb + 1 < 8) ? rec[limb + 1] : 0u;
            const uint32_t d = (sh == 0 ? lo : ((lo >> sh) | (hi << (32 - sh)))) & (BCP_GCOMB_ENTRIES - 1);
            if (d != 0) {
                const uint32_t *e = gcomb + ((size_t)j * BCP_GCOMB_ENTRIES + d) * 16;
#if BCP_PTX
                const uint4 *e4 = reinterpret_cast<const uint4 *>(e);
                uint4 qa = __ldg(e4), qb = __ldg(e4 + 1), qc = __ldg(e4 + 2), qd = __ldg(e4 + 3);
                x.v[0] = qa.x; x.v[1] = qa.y; x.v[2] = qa.z; x.v[3] = qa.w; x.v[4] = qb.x; x.v[5] = qb.y; x.v[6] = qb.z; x.v[7] = qb.w;
                y.v[0] = qc.x; y.v[1] = qc.y; y.v[2] = qc.z; y.v[3] = qc.w; y.v[4] = qd.x; y.v[5] = qd.y; y.v[6] = qd.z; y.v[7] = qd.w;
#else
                for (int i = 0; i < 8; i++) { x.v[i] = e[i]; y.v[i] = e[8 + i]; }
#endif
                have = true;
            }
        }
        if (have) {
            if (acc.inf) { /* first addition into infinity copies the operand (group_impl.h:417-421) */
                acc.x = x; acc.y = y;
                fe_set_u32(acc.z, 1);
                acc.inf = false;
            } else {
                // mixed addition acc += (x, y): 8M + 3S, with the degenerate branches of gej_add_ge_var (:438-448)
                fe z2, u2, s2, rr, h2, h3, v, tt;
                F::sqr(z2, acc.z, bad);
                F::mul(u2, x, z2, bad);
                F::mul(tt, acc.z, z2, bad);
                F::mul(s2, y, tt, bad);
                F::sub(h, u2, acc.x, bad);
                F::sub(rr, s2, acc.y, bad);
                if (fe_is_zero(h)) {
                    if (fe_is_zero(rr)) nd += 1;              /* P + P: one more doubling of this iteration      */
                    else acc.inf = true;                      /* P + (-P)                                       */
                } else {
                    F::sqr(h2, h, bad);
                    F::mul(h3, h2, h, bad);
                    F::mul(v, acc.x, h2, bad);
                    F::mul(acc.z, acc.z, h, bad);
                    F::sqr(tt, rr, bad);
                    F::sub(tt, tt, h3, bad); F::sub(tt, tt, v, bad); F::sub(acc.x, tt, v, bad);
                    F::sub(tt, v, acc.x, bad);
                    F::mul(tt, tt, rr, bad);
                    F::mul(h3, h3, acc.y, bad);
                    F::sub(acc.y, tt, h3, bad);
                }
            }
        }
#if BCP_PTX
#pragma unroll 1
#endif
        for (int k = 0; k < nd; k++)
            if (!acc.inf) gej_double_t<F>(acc, acc, bad);
        if (t < T_TAB) {
            tab.store(t + 1, acc.x, acc.y);
            zr[t + 1] = (t == 0) ? acc.z : h;
        }
    }
    if (bad_out) *bad_out = bad;
    if (WANT_POINT) {
        *out = acc;
        return valid && !acc.inf;
    }
    if (acc.inf || !valid) return false; /* ecdsa_impl.h:220 */
    scalar r;
    sc_set_b32(r, sig64);
    fe z2, t1, fr;
    fe_sqr(z2, acc.z);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) fr.v[i] = r.v[i];
    uint32_t pmn[8], n[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { pmn[i] = SC_P_MINUS_N[i]; n[i] = SC_N[i]; }
    const bool second_chance = !u256_ge(r.v, pmn);
    bool ok = false;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int pass = 0; pass < 2; pass++) {
        fe_mul(t1, fr, z2);
        if (fe_equal(t1, acc.x)) ok = true;
        if (ok || !second_chance) break;
        addn_cc<8>(fr.v, r.v, n);
    }
    return ok;
}

// ---------------------------------------------------------------- one-time: fixed-base comb entry
// entry (j, d) = d * 2^(12 j) * G as a normalized affine point (d >= 1).  Run once per device by the
// table-generation kernel (plain double-and-add with two true field inversions: not a hot path).
BCP_DEV void gej_to_affine(fe &ax, fe &ay, const gej &p) {
    fe zi, zi2, zi3;
    fe_inv(zi, p.z);
    fe_sqr(zi2, zi);
    fe_mul(zi3, zi2, zi);
    fe_mul(ax, p.x, zi2);
    fe_mul(ay, p.y, zi3);
    fe_normalize(ax);
    fe_normalize(ay);
}
BCP_DEV void gcomb_entry(uint32_t *out16, int j, uint32_t d) {
    gej base;
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { base.x.v[i] = FE_GX[i]; base.y.v[i] = FE_GY[i]; }
    fe_set_u32(base.z, 1);
    base.inf = false;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int i = 0; i < j * BCP_GCOMB_BITS; i++) gej_double(base, base);
    fe bx, by;
    gej_to_affine(bx, by, base);
    gej acc;
    fe_set_zero(acc.x); fe_set_zero(acc.y); fe_set_zero(acc.z);
    acc.inf = true;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int b = BCP_GCOMB_BITS - 1; b >= 0; b--) {
        if (!acc.inf) gej_double(acc, acc);
        if ((d >> b) & 1u) gej_add_ge(acc, bx, by, nullptr);
    }
    fe ax, ay;
    gej_to_affine(ax, ay, acc);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { out16[i] = ax.v[i]; out16[8 + i] = ay.v[i]; }
}
