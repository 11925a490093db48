// scalar.cuh — arithmetic modulo the group order n, and the scalar-side preparation of a verify.
//
// Covers what the reference does in src/secp256k1/src/scalar_8x32_impl.h (set_b32 :165-179,
// is_high :220-236, negate :196, mul + reduce_512 :349-651) and scalar_impl.h (inverse :55-244), plus a
// lattice (GLV) split of the point multiplier that the reference build does NOT use (its
// USE_ENDOMORPHISM is off); the split changes only how u2*Q is evaluated, never the value.
// All constants come from tools/secp_consts.py (derived from the SEC2 curve parameters).
//
// This layer runs in the small "prep" kernel (about 5% of the work), so it is written as plain
// loops over 32-bit limbs with 64-bit accumulators that nvcc lowers to IMAD.WIDE; the hot field
// layer (fe.cuh) is the hand-scheduled one.
#pragma once
#include "fe.cuh"

struct scalar {
    uint32_t v[8];
};

BCP_CONST uint32_t SC_N[8] = {0xd0364141u, 0xbfd25e8cu, 0xaf48a03bu, 0xbaaedce6u,
                              0xfffffffeu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
BCP_CONST uint32_t SC_NC[5] = {0x2fc9bebfu, 0x402da173u, 0x50b75fc4u, 0x45512319u, 0x00000001u}; /* 2^256 - n */
BCP_CONST uint32_t SC_NHALF[8] = {0x681b20a0u, 0xdfe92f46u, 0x57a4501du, 0x5d576e73u,
                                  0xffffffffu, 0xffffffffu, 0xffffffffu, 0x7fffffffu};
BCP_CONST uint32_t SC_LAMBDA[8] = {0x1b23bd72u, 0xdf02967cu, 0x20816678u, 0x122e22eau,
                                   0x8812645au, 0xa5261c02u, 0xc05c30e0u, 0x5363ad4cu};
BCP_CONST uint32_t SC_G1[8] = {0x45dbb031u, 0xe893209au, 0x71e8ca7fu, 0x3daa8a14u,
                               0x9284eb15u, 0xe86c90e4u, 0xa7d46bcdu, 0x3086d221u};
BCP_CONST uint32_t SC_G2[8] = {0x8ac47f71u, 0x1571b4aeu, 0x9df506c6u, 0x221208acu,
                               0x0abfe4c4u, 0x6f547fa9u, 0x010e8828u, 0xe4437ed6u};
BCP_CONST uint32_t SC_MINUS_B1[8] = {0x0abfe4c3u, 0x6f547fa9u, 0x010e8828u, 0xe4437ed6u, 0, 0, 0, 0};
/* -b2 mod n = n - b2 */
BCP_CONST uint32_t SC_MINUS_B2[8] = {0x3db1562cu, 0xd765cda8u, 0x0774346du, 0x8a280ac5u,
                                     0xfffffffeu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
/* p - n: bound of the "r + n" second chance in the x comparison (ecdsa_impl.h:45-47,258-266) */
BCP_CONST uint32_t SC_P_MINUS_N[8] = {0x2fc9baeeu, 0x402da172u, 0x50b75fc4u, 0x45512319u, 0x00000001u, 0, 0, 0};

BCP_DEV bool sc_is_zero(const scalar &a) {
    return (a.v[0] | a.v[1] | a.v[2] | a.v[3] | a.v[4] | a.v[5] | a.v[6] | a.v[7]) == 0;
}
// a >= b over 8 limbs
BCP_DEV bool u256_ge(const uint32_t *a, const uint32_t *b) {
    uint32_t t[8];
    return subn_cc<8>(t, a, b) == 0;
}
// big-endian bytes -> scalar, reduced by at most one subtraction of n; returns the overflow flag
BCP_DEV bool sc_set_b32(scalar &r, const uint8_t *b) {
    BCP_UNROLL
    for (int i = 0; i < 8; i++) {
        const uint8_t *q = b + 4 * (7 - i);
        r.v[i] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
    }
    uint32_t n[8], t[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) n[i] = SC_N[i];
    uint32_t bw = subn_cc<8>(t, r.v, n);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) r.v[i] = bw ? r.v[i] : t[i];
    return bw == 0;
}
BCP_DEV bool sc_is_high(const scalar &a) { /* a > n/2 */
    uint32_t h[8], t[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) h[i] = SC_NHALF[i];
    return subn_cc<8>(t, h, a.v) != 0;
}
BCP_DEV void sc_negate(scalar &r, const scalar &a) { /* n - a, 0 -> 0 */
    uint32_t n[8], t[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) n[i] = SC_N[i];
    subn_cc<8>(t, n, a.v);
    bool z = sc_is_zero(a);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) r.v[i] = z ? 0u : t[i];
}
// r (LEN limbs) += a (AL limbs) * b (BL limbs) at limb offset 0; r must be large enough to absorb the sum
template <int AL, int BL, int RL>
BCP_DEV void mac_limbs(uint32_t *r, const uint32_t *a, const uint32_t *b) {
    BCP_UNROLL
    for (int i = 0; i < AL; i++) {
        uint64_t c = 0;
        BCP_UNROLL
        for (int j = 0; j < BL; j++) {
            c += (uint64_t)a[i] * b[j] + r[i + j];
            r[i + j] = (uint32_t)c;
            c >>= 32;
        }
        BCP_UNROLL
        for (int k = i + BL; k < RL; k++) {
            c += r[k];
            r[k] = (uint32_t)c;
            c >>= 32;
        }
    }
}
// 512-bit t -> t mod n
BCP_DEV void sc_reduce512(scalar &r, const uint32_t *t) {
    uint32_t nc[5], n[8];
    BCP_UNROLL
    for (int i = 0; i < 5; i++) nc[i] = SC_NC[i];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) n[i] = SC_N[i];
    // fold 1: lo + hi * NC  (< 2^386) -> 13 limbs
    uint32_t a[13];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) a[i] = t[i];
    BCP_UNROLL
    for (int i = 8; i < 13; i++) a[i] = 0;
    mac_limbs<8, 5, 13>(a, t + 8, nc);
    // fold 2: a[0..7] + a[8..12] * NC (< 2^260) -> 9 limbs (mac writes up to limb 9; limb 9 stays 0)
    uint32_t b[10];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) b[i] = a[i];
    b[8] = 0; b[9] = 0;
    mac_limbs<5, 5, 10>(b, a + 8, nc);
    // fold 3: b[0..7] + b[8] * NC (< 2^256 + 2^134) -> 9 limbs
    uint32_t c[9];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) c[i] = b[i];
    c[8] = 0;
    mac_limbs<1, 5, 9>(c, b + 8, nc);
    // fold 4: a set limb 8 means the value wrapped past 2^256 == NC; the remainder is < 2^134
    uint32_t d[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    BCP_UNROLL
    for (int i = 0; i < 5; i++) d[i] = c[8] ? nc[i] : 0u;
    addn_cc<8>(c, c, d);
    // final conditional subtraction
    uint32_t s[8];
    uint32_t bw = subn_cc<8>(s, c, n);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) r.v[i] = bw ? c[i] : s[i];
}
BCP_DEV void sc_mul(scalar &r, const scalar &a, const scalar &b) {
    uint32_t t[16];
    mul_wide(t, a.v, b.v);
    sc_reduce512(r, t);
}
BCP_DEV void sc_sub(scalar &r, const scalar &a, const scalar &b) {
    uint32_t n[8], t[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) n[i] = SC_N[i];
    uint32_t bw = subn_cc<8>(r.v, a.v, b.v);
    addn_cc<8>(t, r.v, n);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) r.v[i] = bw ? t[i] : r.v[i];
}
BCP_DEV void sc_add(scalar &r, const scalar &a, const scalar &b) {
    uint32_t n[8], t[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) n[i] = SC_N[i];
    uint32_t c = addn_cc<8>(r.v, a.v, b.v);
    uint32_t bw = subn_cc<8>(t, r.v, n);
    bool take = c || !bw;
    BCP_UNROLL
    for (int i = 0; i < 8; i++) r.v[i] = take ? t[i] : r.v[i];
}
// a^(n-2) by fixed 4-bit windows (same value as secp256k1_scalar_inverse, scalar_impl.h:55-218)
BCP_DEV_NOINLINE void sc_inverse(scalar &r, const scalar &a) {
    scalar tab[16];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) tab[0].v[i] = (i == 0);
    tab[1] = a;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int i = 2; i < 16; i++) sc_mul(tab[i], tab[i - 1], a);
    uint32_t e[8];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) e[i] = SC_N[i];
    e[0] -= 2; /* n - 2, no borrow: low limb of n is 0xd0364141 */
    scalar acc = tab[(e[7] >> 28) & 15];
#if BCP_PTX
#pragma unroll 1
#endif
    for (int w = 62; w >= 0; w--) {
        sc_mul(acc, acc, acc); sc_mul(acc, acc, acc); sc_mul(acc, acc, acc); sc_mul(acc, acc, acc);
        uint32_t d = (e[w >> 3] >> ((w & 7) * 4)) & 15;
        sc_mul(acc, acc, tab[d]); /* d == 0 multiplies by 1: keeps the loop uniform */
    }
    r = acc;
}
// ---------------------------------------------------------------- inversion by division steps
// 1/a mod n with the Bernstein-Yang "safegcd" iteration ("Fast constant-time gcd computation and modular inversion", 2019)
// instead of the 329 modular multiplications of a^(n-2): 600 division steps on (f, g) = (n, a), taken 30 at a time on
// the low 30 bits of f and g alone (each batch yields a 2x2 integer matrix), and per batch ONE application of that matrix
// to the full-width pair (f, g) and to the pair (d, e) that tracks a's multiplier modulo n.  Every step is branch-free and
// the same for all inputs, so a warp stays convergent; 590 steps are enough for any 256-bit input.  About a tenth of the
// instructions of sc_inverse — and the inversion is the serial chain that sets the prep kernel's time (one per K
// signatures) and a quarter of one block's signature-check latency.  Values are held as nine signed limbs of 30 bits.
// The reference inverts with an addition chain (scalar_impl.h:55-218); both give the same number, and
// tests/test_hostsim.py and the device self-test compare the two.
struct sc30 {
    int32_t v[9];
};
BCP_CONST int32_t SC_N30[9] = {0x10364141, 0x3f497a33, 0x348a03bb, 0x2bb739ab, 0x3ffffeba, 0x3fffffff, 0x3fffffff, 0x3fffffff, 0xffff};
#define SC_N_INV30 0x2a774ec1u /* n^-1 mod 2^30 */
#define SC_M30 0x3fffffff

// 30 division steps on the low bits; zeta = -(delta + 1/2).  Returns the new zeta, m = {u, v, q, r} with
// 2^30 * (f', g') = [u v; q r] (f, g)
BCP_DEV int32_t sc_divsteps30(int32_t zeta, uint32_t f, uint32_t g, int32_t *m) {
    uint32_t u = 1, v = 0, q = 0, r = 1;
#if BCP_PTX
#pragma unroll 5
#endif
    for (int i = 0; i < 30; i++) {
        const uint32_t neg = (uint32_t)(zeta >> 31); /* all ones when delta > 0 */
        const uint32_t odd = 0u - (g & 1u);          /* all ones when g is odd  */
        // g is odd: g += f or g -= f (by the sign of delta), and the same on the matrix row
        g += ((f ^ neg) - neg) & odd;
        q += ((u ^ neg) - neg) & odd;
        r += ((v ^ neg) - neg) & odd;
        // delta > 0 and g was odd: the old g becomes f (the new g already holds g - f; adding it back restores it)
        const uint32_t swap = neg & odd;
        zeta = (int32_t)((uint32_t)zeta ^ swap) - 1;
        f += g & swap;
        u += q & swap;
        v += r & swap;
        g >>= 1;
        u <<= 1;
        v <<= 1;
    }
    m[0] = (int32_t)u; m[1] = (int32_t)v; m[2] = (int32_t)q; m[3] = (int32_t)r;
    return zeta;
}
// (f, g) <- M (f, g) / 2^30, exact
BCP_DEV void sc_apply_fg(sc30 &f, sc30 &g, const int32_t *m) {
    const int64_t u = m[0], v = m[1], q = m[2], r = m[3];
    int64_t cf = u * f.v[0] + v * g.v[0];
    int64_t cg = q * f.v[0] + r * g.v[0];
    cf >>= 30; /* the low 30 bits are zero by construction */
    cg >>= 30;
    BCP_UNROLL
    for (int i = 1; i < 9; i++) {
        cf += u * f.v[i] + v * g.v[i];
        cg += q * f.v[i] + r * g.v[i];
        f.v[i - 1] = (int32_t)cf & SC_M30; cf >>= 30;
        g.v[i - 1] = (int32_t)cg & SC_M30; cg >>= 30;
    }
    f.v[8] = (int32_t)cf;
    g.v[8] = (int32_t)cg;
}
// (d, e) <- M (d, e) / 2^30 modulo n: a multiple of n is added that makes the low 30 bits vanish; d, e stay in (-2n, n)
BCP_DEV void sc_apply_de(sc30 &d, sc30 &e, const int32_t *m) {
    const int32_t u = m[0], v = m[1], q = m[2], r = m[3];
    const int32_t sd = d.v[8] >> 31, se = e.v[8] >> 31;
    int32_t md = (u & sd) + (v & se); /* + n for a negative d, e: keeps the results in range */
    int32_t me = (q & sd) + (r & se);
    int64_t cd = (int64_t)u * d.v[0] + (int64_t)v * e.v[0];
    int64_t ce = (int64_t)q * d.v[0] + (int64_t)r * e.v[0];
    md -= (int32_t)((SC_N_INV30 * (uint32_t)cd + (uint32_t)md) & SC_M30);
    me -= (int32_t)((SC_N_INV30 * (uint32_t)ce + (uint32_t)me) & SC_M30);
    cd += (int64_t)SC_N30[0] * md;
    ce += (int64_t)SC_N30[0] * me;
    cd >>= 30;
    ce >>= 30;
    BCP_UNROLL
    for (int i = 1; i < 9; i++) {
        cd += (int64_t)u * d.v[i] + (int64_t)v * e.v[i] + (int64_t)SC_N30[i] * md;
        ce += (int64_t)q * d.v[i] + (int64_t)r * e.v[i] + (int64_t)SC_N30[i] * me;
        d.v[i - 1] = (int32_t)cd & SC_M30; cd >>= 30;
        e.v[i - 1] = (int32_t)ce & SC_M30; ce >>= 30;
    }
    d.v[8] = (int32_t)cd;
    e.v[8] = (int32_t)ce;
}
// a^-1 mod n (0 for a = 0, like the reference's scalar_inverse); a in [0, n)
BCP_DEV_NOINLINE void sc_inverse_divsteps(scalar &out, const scalar &a) {
    sc30 d, e, f, g;
    BCP_UNROLL
    for (int i = 0; i < 9; i++) { d.v[i] = 0; e.v[i] = (i == 0); f.v[i] = SC_N30[i]; }
    BCP_UNROLL
    for (int i = 0; i < 9; i++) { /* 8 x 32 bits -> 9 x 30 bits */
        const int bit = 30 * i, w = bit >> 5, sh = bit & 31;
        uint32_t x = a.v[w] >> sh;
        if (sh > 2 && w + 1 < 8) x |= a.v[w + 1] << (32 - sh);
        g.v[i] = (int32_t)(x & SC_M30);
    }
    int32_t zeta = -1;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int it = 0; it < 20; it++) {
        int32_t m[4];
        zeta = sc_divsteps30(zeta, (uint32_t)f.v[0], (uint32_t)g.v[0], m);
        sc_apply_de(d, e, m);
        sc_apply_fg(f, g, m);
    }
    // now g = 0 and f = +-1 (f = +-n for a = 0): the inverse is d times the sign of f.  Bring d from (-2n, n) to [0, n):
    // add n when negative, negate when f is negative, carry, add n once more when still negative, carry.
    int32_t t[9];
    const int32_t addn = d.v[8] >> 31, neg = f.v[8] >> 31;
    BCP_UNROLL
    for (int i = 0; i < 9; i++) t[i] = ((d.v[i] + (SC_N30[i] & addn)) ^ neg) - neg;
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { t[i + 1] += t[i] >> 30; t[i] &= SC_M30; }
    const int32_t addn2 = t[8] >> 31;
    BCP_UNROLL
    for (int i = 0; i < 9; i++) t[i] += SC_N30[i] & addn2;
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { t[i + 1] += t[i] >> 30; t[i] &= SC_M30; }
    BCP_UNROLL
    for (int i = 0; i < 8; i++) { /* 9 x 30 bits -> 8 x 32 bits */
        const int bit = 32 * i, w = bit / 30, sh = bit % 30;
        uint32_t x = (uint32_t)t[w] >> sh;
        x |= (uint32_t)t[w + 1] << (30 - sh);
        if (30 - sh + 30 < 32 && w + 2 < 9) x |= (uint32_t)t[w + 2] << (60 - sh);
        out.v[i] = x;
    }
}

// round(k * g / 2^384) for a 256-bit g: top 128 bits of the 512-bit product plus the rounding bit
BCP_DEV void sc_mul_shift384(scalar &r, const scalar &k, const uint32_t *g_const) {
    uint32_t g[8], t[16];
    BCP_UNROLL
    for (int i = 0; i < 8; i++) g[i] = g_const[i];
    mul_wide(t, k.v, g);
    uint32_t rnd[8] = {t[11] >> 31, 0, 0, 0, 0, 0, 0, 0};
    uint32_t hi[8] = {t[12], t[13], t[14], t[15], 0, 0, 0, 0};
    addn_cc<8>(r.v, hi, rnd);
}
// k = k1 + k2*lambda (mod n) with |k1|, |k2| < 2^129; returns magnitudes (5 limbs) and signs
BCP_DEV void sc_split_lambda(uint32_t *k1abs /*5*/, bool &neg1, uint32_t *k2abs /*5*/, bool &neg2, const scalar &k) {
    scalar c1, c2, t1, t2, k1, k2, cst;
    sc_mul_shift384(c1, k, SC_G1);
    sc_mul_shift384(c2, k, SC_G2);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) cst.v[i] = SC_MINUS_B1[i];
    sc_mul(t1, c1, cst);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) cst.v[i] = SC_MINUS_B2[i];
    sc_mul(t2, c2, cst);
    sc_add(k2, t1, t2);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) cst.v[i] = SC_LAMBDA[i];
    sc_mul(t1, k2, cst);
    sc_sub(k1, k, t1);
    neg1 = sc_is_high(k1);
    neg2 = sc_is_high(k2);
    scalar m1, m2;
    sc_negate(m1, k1);
    sc_negate(m2, k2);
    BCP_UNROLL
    for (int i = 0; i < 5; i++) {
        k1abs[i] = neg1 ? m1.v[i] : k1.v[i];
        k2abs[i] = neg2 ? m2.v[i] : k2.v[i];
    }
}
