// ge.cuh — secp256k1 group law in Jacobian coordinates for sm_100a (y^2 = x^3 + 7, a = 0).
//
// Same maps as the reference's src/secp256k1/src/group_impl.h — gej_double_var (:301-354),
// gej_add_ge_var (:414-461) with its degenerate branches (P==Q -> double :438-441, P==-Q -> infinity
// :442-447, infinity operands :417-421), ge_set_xo_var (:216-237), ge_is_valid_var (:287-299) — on the
// 8x32 weakly-reduced field of fe.cuh.  Results (as group elements) are identical; limb layout
// and evaluation order are this project's own.
#pragma once
#include "fe.cuh"

struct gej {
    fe x, y, z;
    bool inf;
};

BCP_CONST uint32_t FE_BETA[8] = {0x719501eeu, 0xc1396c28u, 0x12f58995u, 0x9cf04975u,
                                 0xac3434e9u, 0x6e64479eu, 0x657c0710u, 0x7ae96a2bu};
BCP_CONST uint32_t FE_GX[8] = {0x16f81798u, 0x59f2815bu, 0x2dce28d9u, 0x029bfcdbu,
                               0xce870b07u, 0x55a06295u, 0xf9dcbbacu, 0x79be667eu};
BCP_CONST uint32_t FE_GY[8] = {0xfb10d4b8u, 0x9c47d08fu, 0xa6855419u, 0xfd17b448u,
                               0x0e1108a8u, 0x5da4fbfcu, 0x26a3c465u, 0x483ada77u};

// The field operations a group formula is written over: `fe_exact` finishes every fold in place (out-of-line branch for
// the practically-never carries), `fe_quick` counts those carries in `bad` instead and leaves the result unreduced then —
// the caller redoes the whole computation with fe_exact when bad != 0 (fe.cuh, "quick forms").
struct fe_exact {
    static BCP_DEV_M void mul(fe &r, const fe &a, const fe &b, uint32_t &) { fe_mul(r, a, b); }
    static BCP_DEV_M void sqr(fe &r, const fe &a, uint32_t &) { fe_sqr(r, a); }
    static BCP_DEV_M void add(fe &r, const fe &a, const fe &b, uint32_t &) { fe_add(r, a, b); }
    static BCP_DEV_M void sub(fe &r, const fe &a, const fe &b, uint32_t &) { fe_sub(r, a, b); }
    static BCP_DEV_M void dbl(fe &r, const fe &a, uint32_t &) { fe_dbl(r, a); }
    static BCP_DEV_M void mul8(fe &r, const fe &a, uint32_t &) { fe_dbl(r, a); fe_dbl(r, r); fe_dbl(r, r); }
    static BCP_DEV_M void neg(fe &r, const fe &a, uint32_t &) { fe_neg(r, a); }
};
struct fe_quick {
    static BCP_DEV_M void mul(fe &r, const fe &a, const fe &b, uint32_t &bad) { fe_mul_q(r, a, b, bad); }
    static BCP_DEV_M void sqr(fe &r, const fe &a, uint32_t &bad) { fe_sqr_q(r, a, bad); }
    static BCP_DEV_M void add(fe &r, const fe &a, const fe &b, uint32_t &bad) { fe_add_q(r, a, b, bad); }
    static BCP_DEV_M void sub(fe &r, const fe &a, const fe &b, uint32_t &bad) { fe_sub_q(r, a, b, bad); }
    static BCP_DEV_M void dbl(fe &r, const fe &a, uint32_t &bad) { fe_dbl_q(r, a, bad); }
    static BCP_DEV_M void mul8(fe &r, const fe &a, uint32_t &bad) { fe_mul8_q(r, a, bad); }
    static BCP_DEV_M void neg(fe &r, const fe &a, uint32_t &bad) { fe_neg_q(r, a, bad); }
};

// r = 2a.  2M + 5S (the a = 0 doubling with X*Y^2 taken from (X + Y^2)^2: one multiplication becomes a squaring,
// which is 28 IMAD.WIDE cheaper here, for three more additions).  No point of order two exists on this curve, so
// y = 0 never occurs for a finite point and the only special case is infinity (group_impl.h:320-326).
template <class F>
BCP_DEV void gej_double_t(gej &r, const gej &a, uint32_t &bad) {
    if (a.inf) { r = a; return; }
    fe t, A, B, C, D, E;
    F::mul(t, a.y, a.z, bad);
    F::sqr(A, a.x, bad);
    F::sqr(B, a.y, bad);
    F::dbl(r.z, t, bad);                 /* Z3 = 2YZ */
    F::add(t, a.x, B, bad);
    F::sqr(C, B, bad);                   /* C = Y^4 */
    F::sqr(t, t, bad);                   /* (X + Y^2)^2 */
    F::sub(t, t, A, bad); F::sub(t, t, C, bad);
    F::dbl(D, t, bad);                   /* D = 4XY^2 */
    F::dbl(E, A, bad); F::add(E, E, A, bad);  /* E = 3X^2 */
    F::sqr(t, E, bad);
    F::sub(t, t, D, bad); F::sub(r.x, t, D, bad); /* X3 = E^2 - 2D */
    F::mul8(C, C, bad);                  /* 8Y^4 */
    F::sub(t, D, r.x, bad);
    F::mul(t, E, t, bad);
    F::sub(r.y, t, C, bad);              /* Y3 = E(D - X3) - 8Y^4 */
    r.inf = false;
}
BCP_DEV void gej_double(gej &r, const gej &a) {
    uint32_t unused = 0;
    gej_double_t<fe_exact>(r, a, unused);
}

// a += (bx, by) with (bx, by) affine and finite.  8M + 3S.  If rzr != nullptr it receives the
// ratio Z3/Z1 (only meaningful on the generic path; table construction never hits the others).
BCP_DEV void gej_add_ge(gej &a, const fe &bx, const fe &by, fe *rzr) {
    if (a.inf) {
        a.x = bx; a.y = by;
        fe_set_u32(a.z, 1);
        a.inf = false;
        return;
    }
    fe z2, u2, s2, h, rr, h2, h3, v, t;
    fe_sqr(z2, a.z);
    fe_mul(u2, bx, z2);
    fe_mul(t, a.z, z2);
    fe_mul(s2, by, t);
    fe_sub(h, u2, a.x);
    fe_sub(rr, s2, a.y);
    if (fe_is_zero(h)) {
        if (fe_is_zero(rr)) {
            gej d = a;
            gej_double(a, d);
        } else {
            a.inf = true;
        }
        return;
    }
    if (rzr) *rzr = h;
    fe_sqr(h2, h);
    fe_mul(h3, h2, h);
    fe_mul(v, a.x, h2);
    fe_mul(a.z, a.z, h);
    fe_sqr(t, rr);
    fe_sub(t, t, h3); fe_sub(t, t, v); fe_sub(a.x, t, v); /* X3 = R^2 - H^3 - 2V */
    fe_sub(t, v, a.x);
    fe_mul(t, t, rr);
    fe_mul(h3, h3, a.y);
    fe_sub(a.y, t, h3);                                    /* Y3 = R(V - X3) - Y1 H^3 */
}

// y^2 == x^3 + 7 ?
BCP_DEV bool ge_on_curve(const fe &x, const fe &y) {
    fe l, r, seven;
    fe_sqr(l, y);
    fe_sqr(r, x);
    fe_mul(r, r, x);
    fe_set_u32(seven, 7);
    fe_add(r, r, seven);
    return fe_equal(l, r);
}

// secp256k1_eckey_pubkey_parse (eckey_impl.h:17-35): 33-byte 02/03 -> decompress (x < p, residue test,
// parity), 65-byte 04/06/07 -> range + (hybrid) parity + curve check.  `len` is decided by the caller
// from the prefix byte as CPubKey does (src/pubkey.h:47-51).  Outputs a normalized affine point.
BCP_DEV bool pubkey_parse(fe &qx, fe &qy, const uint8_t *pub) {
    uint8_t pre = pub[0];
    if (pre == 2 || pre == 3) {
        if (!fe_set_b32(qx, pub + 1)) return false;
        fe c, seven, y, chk;
        fe_sqr(c, qx);
        fe_mul(c, c, qx);
        fe_set_u32(seven, 7);
        fe_add(c, c, seven);
        fe_sqrt_candidate(y, c);
        fe_sqr(chk, y);
        if (!fe_equal(chk, c)) return false;
        fe_normalize(y);
        fe ny;
        fe_neg(ny, y);
        fe_normalize(ny);
        bool flip = (bool)(y.v[0] & 1u) != (pre == 3);
        BCP_UNROLL
        for (int i = 0; i < 8; i++) qy.v[i] = flip ? ny.v[i] : y.v[i];
        return true;
    }
    if (pre == 4 || pre == 6 || pre == 7) {
        if (!fe_set_b32(qx, pub + 1)) return false;
        if (!fe_set_b32(qy, pub + 33)) return false;
        if (pre != 4 && ((bool)(qy.v[0] & 1u) != (pre == 7))) return false;
        return ge_on_curve(qx, qy);
    }
    return false;
}
