// fe.cuh — arithmetic in F_p, p = 2^256 - 2^32 - 977, for sm_100a.
//
// Replaces (functionally) the reference's CPU field layer src/secp256k1/src/field_10x26_impl.h /
// field_5x52_impl.h (fe_mul :1059, fe_sqr :1074, fe_normalize :43-264, fe_set_b32 :323, fe_negate/
// fe_add :386-453) and field_impl.h (fe_sqrt :38-134).  It is NOT a port: the reference keeps 10x26 or
// 5x52 limbs with lazy "magnitudes"; here an element is 8 x 32-bit limbs (little-endian) holding
// a value in [0, 2^256) ("weakly reduced": it may exceed p by less than 2^32+977), because B200
// has no 64-bit multiplier and the 32x32->64 IMAD.WIDE with carry (IMAD.WIDE.U32.X) is the
// widest native multiply-accumulate.  Products are formed as two interleaved accumulators (even and
// odd limb positions) so every multiply-accumulate is a 64-bit-aligned IMAD.WIDE on a carry chain,
// then folded with 2^256 == 2^32 + 977 (mod p).
//
// Algorithmic cost convention used for the roofline: one field multiplication =
// 64 mul32, one squaring = 36 mul32 (reduction not counted).
#pragma once
#include "bcp_common.cuh"

struct fe {
    uint32_t v[8];
};

#define BCP_P0 0xFFFFFC2Fu
#define BCP_P1 0xFFFFFFFEu
#define BCP_FOLD 977u /* 2^256 = 2^32 + 977 (mod p) */
#ifndef BCP_FE_SHORT_FOLD
#define BCP_FE_SHORT_FOLD 1 /* finish the folds that practically never propagate out of line (see fe_add) */
#endif

// ------------------------------------------------------------------ limb primitives
// X[0..7] = { a0*b, a2*b, a4*b, a6*b } as four independent 64-bit products
BCP_DEV void row_set(uint32_t *X, uint32_t a0, uint32_t a2, uint32_t a4, uint32_t a6, uint32_t b) {
#if BCP_PTX
    /* mul.wide, not a mul.lo / mul.hi pair: ptxas splits a pair into IMAD + IMAD.HI (2 + 4 issue cycles) when it likes */
    asm("{\n\t.reg .u64 p0, p1, p2, p3;\n\t"
        "mul.wide.u32 p0, %8, %12;\n\tmul.wide.u32 p1, %9, %12;\n\tmul.wide.u32 p2, %10, %12;\n\tmul.wide.u32 p3, %11, %12;\n\t"
        "mov.b64 {%0, %1}, p0;\n\tmov.b64 {%2, %3}, p1;\n\tmov.b64 {%4, %5}, p2;\n\tmov.b64 {%6, %7}, p3;\n\t}"
        : "=r"(X[0]), "=r"(X[1]), "=r"(X[2]), "=r"(X[3]), "=r"(X[4]), "=r"(X[5]), "=r"(X[6]), "=r"(X[7])
        : "r"(a0), "r"(a2), "r"(a4), "r"(a6), "r"(b));
#else
    uint64_t p;
    p = (uint64_t)a0 * b; X[0] = (uint32_t)p; X[1] = (uint32_t)(p >> 32);
    p = (uint64_t)a2 * b; X[2] = (uint32_t)p; X[3] = (uint32_t)(p >> 32);
    p = (uint64_t)a4 * b; X[4] = (uint32_t)p; X[5] = (uint32_t)(p >> 32);
    p = (uint64_t)a6 * b; X[6] = (uint32_t)p; X[7] = (uint32_t)(p >> 32);
#endif
}
// X[0..8] += { a0*b, a2*b, a4*b, a6*b } placed at limbs (0,1),(2,3),(4,5),(6,7); X[8] absorbs the carry
BCP_DEV void row_acc9(uint32_t *X, uint32_t a0, uint32_t a2, uint32_t a4, uint32_t a6, uint32_t b) {
#if BCP_PTX
    asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\tmadc.hi.cc.u32 %1, %9, %13, %1;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %2;\n\tmadc.hi.cc.u32 %3, %10, %13, %3;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %4;\n\tmadc.hi.cc.u32 %5, %11, %13, %5;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, %6;\n\tmadc.hi.cc.u32 %7, %12, %13, %7;\n\t"
        "addc.u32 %8, %8, 0;"
        : "+r"(X[0]), "+r"(X[1]), "+r"(X[2]), "+r"(X[3]), "+r"(X[4]), "+r"(X[5]), "+r"(X[6]), "+r"(X[7]), "+r"(X[8])
        : "r"(a0), "r"(a2), "r"(a4), "r"(a6), "r"(b));
#else
    uint64_t c = 0;
    const uint32_t a[4] = {a0, a2, a4, a6};
    for (int j = 0; j < 4; j++) {
        uint64_t p = (uint64_t)a[j] * b;
        uint64_t lo = (uint64_t)X[2 * j] + (uint32_t)p + c;
        X[2 * j] = (uint32_t)lo;
        uint64_t hi = (uint64_t)X[2 * j + 1] + (uint32_t)(p >> 32) + (lo >> 32);
        X[2 * j + 1] = (uint32_t)hi;
        c = hi >> 32;
    }
    X[8] += (uint32_t)c;
#endif
}
// same, carry-out known to be zero (top of the product) and dropped
BCP_DEV void row_acc8(uint32_t *X, uint32_t a0, uint32_t a2, uint32_t a4, uint32_t a6, uint32_t b) {
#if BCP_PTX
    asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\tmadc.hi.cc.u32 %1, %8, %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9, %12, %2;\n\tmadc.hi.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\tmadc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, %6;\n\tmadc.hi.u32 %7, %11, %12, %7;"
        : "+r"(X[0]), "+r"(X[1]), "+r"(X[2]), "+r"(X[3]), "+r"(X[4]), "+r"(X[5]), "+r"(X[6]), "+r"(X[7])
        : "r"(a0), "r"(a2), "r"(a4), "r"(a6), "r"(b));
#else
    uint32_t t[9];
    for (int j = 0; j < 8; j++) t[j] = X[j];
    t[8] = 0;
    row_acc9(t, a0, a2, a4, a6, b);
    for (int j = 0; j < 8; j++) X[j] = t[j];
#endif
}
// X[0..2K-1] += { a[0]*b, ..., a[K-1]*b } at limb pairs (0,1),(2,3),...; X[2K] absorbs the carry
template <int K>
BCP_DEV void row_acc_k(uint32_t *X, const uint32_t *a, uint32_t b) {
#if BCP_PTX
    static_assert(K >= 1 && K <= 3, "row width");
    if (K == 1) {
        asm("mad.lo.cc.u32 %0, %3, %4, %0;\n\t"
            "madc.hi.cc.u32 %1, %3, %4, %1;\n\t"
            "addc.u32 %2, %2, 0;"
            : "+r"(X[0]), "+r"(X[1]), "+r"(X[2])
            : "r"(a[0]), "r"(b));
    } else if (K == 2) {
        asm("mad.lo.cc.u32 %0, %5, %7, %0;\n\t"
            "madc.hi.cc.u32 %1, %5, %7, %1;\n\t"
            "madc.lo.cc.u32 %2, %6, %7, %2;\n\t"
            "madc.hi.cc.u32 %3, %6, %7, %3;\n\t"
            "addc.u32 %4, %4, 0;"
            : "+r"(X[0]), "+r"(X[1]), "+r"(X[2]), "+r"(X[3]), "+r"(X[4])
            : "r"(a[0]), "r"(a[1]), "r"(b));
    } else {
        asm("mad.lo.cc.u32 %0, %7, %10, %0;\n\t"
            "madc.hi.cc.u32 %1, %7, %10, %1;\n\t"
            "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
            "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
            "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
            "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
            "addc.u32 %6, %6, 0;"
            : "+r"(X[0]), "+r"(X[1]), "+r"(X[2]), "+r"(X[3]), "+r"(X[4]), "+r"(X[5]), "+r"(X[6])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(b));
    }
#else
    uint64_t c = 0;
    for (int j = 0; j < K; j++) {
        uint64_t p = (uint64_t)a[j] * b;
        uint64_t lo = (uint64_t)X[2 * j] + (uint32_t)p + c;
        X[2 * j] = (uint32_t)lo;
        uint64_t hi = (uint64_t)X[2 * j + 1] + (uint32_t)(p >> 32) + (lo >> 32);
        X[2 * j + 1] = (uint32_t)hi;
        c = hi >> 32;
    }
    X[2 * K] += (uint32_t)c;
#endif
}
BCP_DEV void wide_mul(uint32_t &lo, uint32_t &hi, uint32_t a, uint32_t b) {
#if BCP_PTX
    asm("{\n\t.reg .u64 p;\n\tmul.wide.u32 p, %2, %3;\n\tmov.b64 {%0, %1}, p;\n\t}" : "=r"(lo), "=r"(hi) : "r"(a), "r"(b));
#else
    uint64_t p = (uint64_t)a * b;
    lo = (uint32_t)p;
    hi = (uint32_t)(p >> 32);
#endif
}
// r[0..n-1] = a + b, returns carry (0/1).  One asm statement per chain: the carry flag never has to
// survive between statements.  (Outputs go through temporaries so that r may alias a or b.)
template <int N>
BCP_DEV uint32_t addn_cc(uint32_t *r, const uint32_t *a, const uint32_t *b) {
    uint32_t c;
#if BCP_PTX
    static_assert(N == 8 || N == 5 || N == 4 || N == 3 || N == 2, "unsupported chain length");
    if (N == 4) {
        uint32_t r0, r1, r2, r3;
        asm("add.cc.u32 %0, %5, %9;\n\t"
            "addc.cc.u32 %1, %6, %10;\n\t"
            "addc.cc.u32 %2, %7, %11;\n\t"
            "addc.cc.u32 %3, %8, %12;\n\t"
            "addc.u32 %4, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
    } else if (N == 8) {
        uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
        asm("add.cc.u32 %0, %9, %17;\n\t"
            "addc.cc.u32 %1, %10, %18;\n\t"
            "addc.cc.u32 %2, %11, %19;\n\t"
            "addc.cc.u32 %3, %12, %20;\n\t"
            "addc.cc.u32 %4, %13, %21;\n\t"
            "addc.cc.u32 %5, %14, %22;\n\t"
            "addc.cc.u32 %6, %15, %23;\n\t"
            "addc.cc.u32 %7, %16, %24;\n\t"
            "addc.u32 %8, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
        r[4] = r4;
        r[5] = r5;
        r[6] = r6;
        r[7] = r7;
    } else if (N == 5) {
        uint32_t r0, r1, r2, r3, r4;
        asm("add.cc.u32 %0, %6, %11;\n\t"
            "addc.cc.u32 %1, %7, %12;\n\t"
            "addc.cc.u32 %2, %8, %13;\n\t"
            "addc.cc.u32 %3, %9, %14;\n\t"
            "addc.cc.u32 %4, %10, %15;\n\t"
            "addc.u32 %5, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
        r[4] = r4;
    } else if (N == 3) {
        uint32_t r0, r1, r2;
        asm("add.cc.u32 %0, %4, %7;\n\t"
            "addc.cc.u32 %1, %5, %8;\n\t"
            "addc.cc.u32 %2, %6, %9;\n\t"
            "addc.u32 %3, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(b[0]), "r"(b[1]), "r"(b[2]));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
    } else {
        uint32_t r0, r1;
        asm("add.cc.u32 %0, %3, %5;\n\t"
            "addc.cc.u32 %1, %4, %6;\n\t"
            "addc.u32 %2, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(b[0]), "r"(b[1]));
        r[0] = r0;
        r[1] = r1;
    }
#else
    uint64_t t = 0;
    for (int i = 0; i < N; i++) {
        t += (uint64_t)a[i] + b[i];
        r[i] = (uint32_t)t;
        t >>= 32;
    }
    c = (uint32_t)t;
#endif
    return c;
}
// r = a - b, returns borrow (0/1)
template <int N>
BCP_DEV uint32_t subn_cc(uint32_t *r, const uint32_t *a, const uint32_t *b) {
    uint32_t c;
#if BCP_PTX
    static_assert(N == 8 || N == 4 || N == 2, "unsupported chain length");
    if (N == 4) {
        uint32_t r0, r1, r2, r3;
        asm("sub.cc.u32 %0, %5, %9;\n\t"
            "subc.cc.u32 %1, %6, %10;\n\t"
            "subc.cc.u32 %2, %7, %11;\n\t"
            "subc.cc.u32 %3, %8, %12;\n\t"
            "subc.u32 %4, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
    } else if (N == 8) {
        uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
        asm("sub.cc.u32 %0, %9, %17;\n\t"
            "subc.cc.u32 %1, %10, %18;\n\t"
            "subc.cc.u32 %2, %11, %19;\n\t"
            "subc.cc.u32 %3, %12, %20;\n\t"
            "subc.cc.u32 %4, %13, %21;\n\t"
            "subc.cc.u32 %5, %14, %22;\n\t"
            "subc.cc.u32 %6, %15, %23;\n\t"
            "subc.cc.u32 %7, %16, %24;\n\t"
            "subc.u32 %8, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
        r[4] = r4;
        r[5] = r5;
        r[6] = r6;
        r[7] = r7;
    } else {
        uint32_t r0, r1;
        asm("sub.cc.u32 %0, %3, %5;\n\t"
            "subc.cc.u32 %1, %4, %6;\n\t"
            "subc.u32 %2, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(c)
            : "r"(a[0]), "r"(a[1]), "r"(b[0]), "r"(b[1]));
        r[0] = r0;
        r[1] = r1;
    }
    c &= 1u; /* subc 0 - 0 - borrow gives 0 or 0xFFFFFFFF */
#else
    uint64_t br = 0;
    for (int i = 0; i < N; i++) {
        uint64_t t = (uint64_t)a[i] - b[i] - br;
        r[i] = (uint32_t)t;
        br = (t >> 32) & 1;
    }
    c = (uint32_t)br;
#endif
    return c;
}

// r = a + b + cin (cin is 0/1), returns the carry out.  The carry flag is primed with cin by adding 0xFFFFFFFF to it.
template <int N>
BCP_DEV uint32_t addn_cin(uint32_t *r, const uint32_t *a, const uint32_t *b, uint32_t cin) {
    uint32_t c;
#if BCP_PTX
    static_assert(N == 8 || N == 4, "unsupported chain length");
    if (N == 8) {
        uint32_t r0, r1, r2, r3, r4, r5, r6, r7, scratch;
        asm("add.cc.u32 %9, %26, 0xffffffff;\n\t"
            "addc.cc.u32 %0, %10, %18;\n\t"
            "addc.cc.u32 %1, %11, %19;\n\t"
            "addc.cc.u32 %2, %12, %20;\n\t"
            "addc.cc.u32 %3, %13, %21;\n\t"
            "addc.cc.u32 %4, %14, %22;\n\t"
            "addc.cc.u32 %5, %15, %23;\n\t"
            "addc.cc.u32 %6, %16, %24;\n\t"
            "addc.cc.u32 %7, %17, %25;\n\t"
            "addc.u32 %8, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7), "=r"(c), "=r"(scratch)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]), "r"(cin));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
        r[4] = r4;
        r[5] = r5;
        r[6] = r6;
        r[7] = r7;
        (void)scratch;
    } else {
        uint32_t r0, r1, r2, r3, scratch;
        asm("add.cc.u32 %5, %14, 0xffffffff;\n\t"
            "addc.cc.u32 %0, %6, %10;\n\t"
            "addc.cc.u32 %1, %7, %11;\n\t"
            "addc.cc.u32 %2, %8, %12;\n\t"
            "addc.cc.u32 %3, %9, %13;\n\t"
            "addc.u32 %4, 0, 0;"
            : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(c), "=r"(scratch)
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(cin));
        r[0] = r0;
        r[1] = r1;
        r[2] = r2;
        r[3] = r3;
        (void)scratch;
    }
#else
    uint64_t t = cin;
    for (int i = 0; i < N; i++) {
        t += (uint64_t)a[i] + b[i];
        r[i] = (uint32_t)t;
        t >>= 32;
    }
    c = (uint32_t)t;
#endif
    return c;
}
// r = E + (O << 32): E[k] sits at limb k, O[k] at limb k+1 (O[15] is always zero)
BCP_DEV void merge_even_odd(uint32_t *r /*16*/, const uint32_t *E, const uint32_t *O) {
    r[0] = E[0];
#if BCP_PTX
    asm("add.cc.u32 %0, %15, %30;\n\t"
        "addc.cc.u32 %1, %16, %31;\n\t"
        "addc.cc.u32 %2, %17, %32;\n\t"
        "addc.cc.u32 %3, %18, %33;\n\t"
        "addc.cc.u32 %4, %19, %34;\n\t"
        "addc.cc.u32 %5, %20, %35;\n\t"
        "addc.cc.u32 %6, %21, %36;\n\t"
        "addc.cc.u32 %7, %22, %37;\n\t"
        "addc.cc.u32 %8, %23, %38;\n\t"
        "addc.cc.u32 %9, %24, %39;\n\t"
        "addc.cc.u32 %10, %25, %40;\n\t"
        "addc.cc.u32 %11, %26, %41;\n\t"
        "addc.cc.u32 %12, %27, %42;\n\t"
        "addc.cc.u32 %13, %28, %43;\n\t"
        "addc.u32 %14, %29, %44;"
        : "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(E[1]), "r"(E[2]), "r"(E[3]), "r"(E[4]), "r"(E[5]), "r"(E[6]), "r"(E[7]), "r"(E[8]), "r"(E[9]), "r"(E[10]), "r"(E[11]), "r"(E[12]), "r"(E[13]), "r"(E[14]), "r"(E[15]), "r"(O[0]), "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]), "r"(O[5]), "r"(O[6]), "r"(O[7]), "r"(O[8]), "r"(O[9]), "r"(O[10]), "r"(O[11]), "r"(O[12]), "r"(O[13]), "r"(O[14]));
#else
    uint64_t t = 0;
    for (int k = 1; k < 16; k++) {
        t += (uint64_t)E[k] + O[k - 1];
        r[k] = (uint32_t)t;
        t >>= 32;
    }
#endif
}
// r = a + b over 16 limbs (no carry out: callers guarantee the sum fits)
BCP_DEV void add16(uint32_t *r, const uint32_t *a, const uint32_t *b) {
#if BCP_PTX
    asm("add.cc.u32 %0, %16, %32;\n\t"
        "addc.cc.u32 %1, %17, %33;\n\t"
        "addc.cc.u32 %2, %18, %34;\n\t"
        "addc.cc.u32 %3, %19, %35;\n\t"
        "addc.cc.u32 %4, %20, %36;\n\t"
        "addc.cc.u32 %5, %21, %37;\n\t"
        "addc.cc.u32 %6, %22, %38;\n\t"
        "addc.cc.u32 %7, %23, %39;\n\t"
        "addc.cc.u32 %8, %24, %40;\n\t"
        "addc.cc.u32 %9, %25, %41;\n\t"
        "addc.cc.u32 %10, %26, %42;\n\t"
        "addc.cc.u32 %11, %27, %43;\n\t"
        "addc.cc.u32 %12, %28, %44;\n\t"
        "addc.cc.u32 %13, %29, %45;\n\t"
        "addc.cc.u32 %14, %30, %46;\n\t"
        "addc.u32 %15, %31, %47;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(a[8]), "r"(a[9]), "r"(a[10]), "r"(a[11]), "r"(a[12]), "r"(a[13]), "r"(a[14]), "r"(a[15]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]), "r"(b[8]), "r"(b[9]), "r"(b[10]), "r"(b[11]), "r"(b[12]), "r"(b[13]), "r"(b[14]), "r"(b[15]));
#else
    uint64_t t = 0;
    for (int k = 0; k < 16; k++) {
        t += (uint64_t)a[k] + b[k];
        r[k] = (uint32_t)t;
        t >>= 32;
    }
#endif
}
// ------------------------------------------------------------------ 256x256 -> 512
BCP_DEV void mul_wide(uint32_t *r /*16*/, const uint32_t *a, const uint32_t *b) {
    uint32_t E[16], O[16]; /* E[k] sits at limb k, O[k] at limb k+1 */
    row_set(E, a[0], a[2], a[4], a[6], b[0]);
    row_set(O, a[1], a[3], a[5], a[7], b[0]);
    BCP_UNROLL
    for (int k = 8; k < 16; k++) { E[k] = 0; O[k] = 0; }
    BCP_UNROLL
    for (int i = 1; i < 8; i++) {
        if (i & 1) {
            row_acc9(O + i - 1, a[0], a[2], a[4], a[6], b[i]);
            if (i < 7) row_acc9(E + i + 1, a[1], a[3], a[5], a[7], b[i]);
            else row_acc8(E + 8, a[1], a[3], a[5], a[7], b[i]);
        } else {
            row_acc9(E + i, a[0], a[2], a[4], a[6], b[i]);
            row_acc9(O + i, a[1], a[3], a[5], a[7], b[i]);
        }
    }
    merge_even_odd(r, E, O);
}

// ------------------------------------------------------------------ 128x128 -> 256 and one Karatsuba level
// same even/odd carry-chain layout as mul_wide, rows of two products: 16 IMAD.WIDE
BCP_DEV void mul4_wide(uint32_t *r /*8*/, const uint32_t *a, const uint32_t *b) {
    uint32_t E[9], O[9];
    wide_mul(E[0], E[1], a[0], b[0]); wide_mul(E[2], E[3], a[2], b[0]);
    wide_mul(O[0], O[1], a[1], b[0]); wide_mul(O[2], O[3], a[3], b[0]);
    BCP_UNROLL
    for (int k = 4; k < 9; k++) { E[k] = 0; O[k] = 0; }
    const uint32_t ev[2] = {a[0], a[2]}, od[2] = {a[1], a[3]};
    row_acc_k<2>(O + 0, ev, b[1]); /* limbs 1,3 */
    row_acc_k<2>(E + 2, od, b[1]); /* limbs 2,4 */
    row_acc_k<2>(E + 2, ev, b[2]); /* limbs 2,4 */
    row_acc_k<2>(O + 2, od, b[2]); /* limbs 3,5 */
    row_acc_k<2>(O + 2, ev, b[3]); /* limbs 3,5 */
    row_acc_k<2>(E + 4, od, b[3]); /* limbs 4,6; the product fits 8 limbs, E[8] stays 0 */
    uint32_t t[8];
    addn_cc<8>(t, E + 1, O); /* t[k] = E[k+1] + O[k]; t[7] = E[8] + O[7] = 0 */
    r[0] = E[0];
    BCP_UNROLL
    for (int k = 1; k < 8; k++) r[k] = t[k - 1];
}
// d = |x - y| over 4 limbs, returns 1 when x < y
BCP_DEV uint32_t absdiff4(uint32_t *d, const uint32_t *x, const uint32_t *y) {
    uint32_t t[4];
    const uint32_t bw = subn_cc<4>(t, x, y);
    const uint32_t mask = 0u - bw;
    uint32_t f[4], z[4] = {0, 0, 0, 0};
    BCP_UNROLL
    for (int i = 0; i < 4; i++) f[i] = t[i] ^ mask;
    addn_cin<4>(d, f, z, bw); /* two's complement negate when the difference was negative */
    return bw;
}
// 256x256 -> 512 with one level of (subtractive) Karatsuba: 48 IMAD.WIDE instead of 64.
// MEASURED SLOWER on B200 (39.8 vs 42.5 M sigverifies/s): the 16 saved IMAD.WIDE (64 pipe cycles) cost ~65 extra ALU
// instructions, longer carry chains and spills at the 128-register budget; kept behind BCP_FE_KARATSUBA (default 0)
// as a recorded experiment, exercised by tests/test_hostsim.py.
//   a*b = z0 + (z0 + z2 + (a_lo - a_hi)(b_hi - b_lo)) * 2^128 + z2 * 2^256,  z0 = a_lo*b_lo, z2 = a_hi*b_hi
BCP_DEV void mul_wide_karatsuba(uint32_t *r /*16*/, const uint32_t *a, const uint32_t *b) {
    uint32_t z0[8], z2[8], m[8], da[4], db[4];
    mul4_wide(z0, a, b);
    mul4_wide(z2, a + 4, b + 4);
    const uint32_t sa = absdiff4(da, a, a + 4);
    const uint32_t sb = absdiff4(db, b + 4, b);
    mul4_wide(m, da, db);
    const uint32_t neg = sa ^ sb, mask = 0u - neg;
    uint32_t mid[9], mx[8];
    mid[8] = addn_cc<8>(mid, z0, z2);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) mx[i] = m[i] ^ mask;
    mid[8] += addn_cin<8>(mid, mid, mx, neg) - neg; /* +m, or +(~m + 1) - 2^256 = -m; the true middle term is >= 0 */
    BCP_UNROLL
    for (int i = 0; i < 4; i++) r[i] = z0[i];
    uint32_t x[8] = {z0[4], z0[5], z0[6], z0[7], z2[0], z2[1], z2[2], z2[3]};
    uint32_t c = addn_cc<8>(r + 4, x, mid);
    uint32_t y[4] = {mid[8], 0, 0, 0};
    addn_cin<4>(r + 12, z2 + 4, y, c);
}

// ------------------------------------------------------------------ 512 -> [0, 2^256) mod p
template <bool SHORT> BCP_DEV void fe_reduce_wide_t(fe &r, const uint32_t *t /*16*/) {
    // first fold: S = L + H*977 + (H << 32), accumulated in place as two interleaved arrays
    uint32_t EV[9], OD[9];
    BCP_UNROLL
    for (int k = 0; k < 8; k++) { EV[k] = t[k]; OD[k] = t[8 + k]; }
    EV[8] = 0; OD[8] = 0;
    row_acc9(EV, t[8], t[10], t[12], t[14], BCP_FOLD);  /* limbs (0,1),(2,3),(4,5),(6,7) += H_even*977        */
    row_acc9(OD, t[9], t[11], t[13], t[15], BCP_FOLD);  /* OD sits at limb 1: (H << 32) + H_odd*977 << 32      */
    uint32_t s[10];
    s[0] = EV[0];
    s[9] = OD[8] + addn_cc<8>(s + 1, EV + 1, OD);       /* S < 2^290, so s[9] <= 3 and s[9]:s[8] < 2^34      */
    // second fold: T = s[9]:s[8];  S' = s[0..7] + T*977 + (T << 32)
    uint64_t m = (uint64_t)s[8] * BCP_FOLD + ((uint64_t)(s[9] * BCP_FOLD) << 32);
    uint64_t a1 = (m >> 32) + s[8];
    uint32_t A[8] = {(uint32_t)m, (uint32_t)a1, s[9] + (uint32_t)(a1 >> 32), 0, 0, 0, 0, 0};
    // T*977 + (T << 32) only reaches limbs 0..2; a carry out of limb 2 (probability about 2^-30) is the only way the
    // upper limbs — or a third fold — can be affected, and then the full chains below run instead
    if (SHORT) {
        uint32_t lo3[3];
        if (__builtin_expect(addn_cc<3>(lo3, s, A) == 0, 1)) {
            r.v[0] = lo3[0]; r.v[1] = lo3[1]; r.v[2] = lo3[2];
            BCP_UNROLL
            for (int i = 3; i < 8; i++) r.v[i] = s[i];
            return;
        }
    }
    uint32_t c = addn_cc<8>(r.v, s, A);
    // third fold: a carry means S' wrapped past 2^256 (== 2^32+977); the wrapped value is < 2^67,
    // so only limbs 0..2 can change and no further wrap is possible
    uint32_t B[3] = {(0u - c) & BCP_FOLD, c, 0}; /* c is 0/1: mask instead of a multiply (keeps the IMAD pipe free) */
    addn_cc<3>(r.v, r.v, B);
}
BCP_DEV void fe_reduce_wide(fe &r, const uint32_t *t /*16*/) { fe_reduce_wide_t<BCP_FE_SHORT_FOLD != 0>(r, t); }

// The multiplication and the squaring are real (non-inlined) device functions taking and returning the
// element BY VALUE: ptxas passes the 8+8 limbs in registers (no local-memory traffic, checked in SASS), a call
// costs a CALL/RET pair plus a few moves, and every kernel holds exactly one copy of each.  The fully
// inlined first version of the point kernel was 800 KB of SASS, later 105 KB with a 45 KB hot loop, and in
// both the top stall reason was instruction fetch (profiles/r01_*); with calls the hot loop fits the
// instruction cache.
#ifndef BCP_FE_CALLS
#define BCP_FE_CALLS 1
#endif
#if BCP_FE_CALLS
#define BCP_FE_FN BCP_DEV_NOINLINE
#else
#define BCP_FE_FN BCP_DEV
#endif
#ifndef BCP_FE_KARATSUBA
#define BCP_FE_KARATSUBA 0
#endif
BCP_FE_FN fe fe_mul_f(fe a, fe b) {
    uint32_t t[16];
    fe r;
#if BCP_FE_KARATSUBA
    mul_wide_karatsuba(t, a.v, b.v);
#else
    mul_wide(t, a.v, b.v);
#endif
    fe_reduce_wide(r, t);
    return r;
}
BCP_DEV void fe_mul(fe &r, const fe &a, const fe &b) { r = fe_mul_f(a, b); }
// a^2 as 512 bits: 28 off-diagonal products (same even/odd carry-chain layout as mul_wide, rows shrink as i
// grows), doubled by a funnel shift, plus the 8 diagonal squares: 36 IMAD.WIDE instead of 64.
BCP_DEV void sqr_wide(uint32_t *r /*16*/, const uint32_t *a) {
    uint32_t E[16], O[16];
    BCP_UNROLL
    for (int k = 0; k < 16; k++) { E[k] = 0; O[k] = 0; }
    // i = 0: a0 * {a1,a3,a5,a7} -> limbs 1,3,5,7 (O 0,2,4,6);  a0 * {a2,a4,a6} -> limbs 2,4,6
    wide_mul(O[0], O[1], a[1], a[0]); wide_mul(O[2], O[3], a[3], a[0]);
    wide_mul(O[4], O[5], a[5], a[0]); wide_mul(O[6], O[7], a[7], a[0]);
    wide_mul(E[2], E[3], a[2], a[0]); wide_mul(E[4], E[5], a[4], a[0]); wide_mul(E[6], E[7], a[6], a[0]);
    { const uint32_t m[3] = {a[2], a[4], a[6]}; row_acc_k<3>(O + 2, m, a[1]); }   /* limbs 3,5,7   */
    { const uint32_t m[3] = {a[3], a[5], a[7]}; row_acc_k<3>(E + 4, m, a[1]); }   /* limbs 4,6,8   */
    { const uint32_t m[3] = {a[3], a[5], a[7]}; row_acc_k<3>(O + 4, m, a[2]); }   /* limbs 5,7,9   */
    { const uint32_t m[2] = {a[4], a[6]}; row_acc_k<2>(E + 6, m, a[2]); }         /* limbs 6,8     */
    { const uint32_t m[2] = {a[4], a[6]}; row_acc_k<2>(O + 6, m, a[3]); }         /* limbs 7,9     */
    { const uint32_t m[2] = {a[5], a[7]}; row_acc_k<2>(E + 8, m, a[3]); }         /* limbs 8,10    */
    { const uint32_t m[2] = {a[5], a[7]}; row_acc_k<2>(O + 8, m, a[4]); }         /* limbs 9,11    */
    { const uint32_t m[1] = {a[6]}; row_acc_k<1>(E + 10, m, a[4]); }              /* limb 10       */
    { const uint32_t m[1] = {a[6]}; row_acc_k<1>(O + 10, m, a[5]); }              /* limb 11       */
    { const uint32_t m[1] = {a[7]}; row_acc_k<1>(E + 12, m, a[5]); }              /* limb 12       */
    { const uint32_t m[1] = {a[7]}; row_acc_k<1>(O + 12, m, a[6]); }              /* limb 13       */
    uint32_t S[16], T[16];
    merge_even_odd(S, E, O);
    T[0] = S[0] << 1;
    BCP_UNROLL
    for (int k = 1; k < 16; k++) T[k] = (S[k] << 1) | (S[k - 1] >> 31);
#if BCP_PTX
    // r = T + the eight diagonal squares, as one carry chain of multiply-adds (8 IMAD.WIDE.U32.X, no separate additions)
    asm("mad.lo.cc.u32 %0, %16, %16, %24;\n\tmadc.hi.cc.u32 %1, %16, %16, %25;\n\t"
        "madc.lo.cc.u32 %2, %17, %17, %26;\n\tmadc.hi.cc.u32 %3, %17, %17, %27;\n\t"
        "madc.lo.cc.u32 %4, %18, %18, %28;\n\tmadc.hi.cc.u32 %5, %18, %18, %29;\n\t"
        "madc.lo.cc.u32 %6, %19, %19, %30;\n\tmadc.hi.cc.u32 %7, %19, %19, %31;\n\t"
        "madc.lo.cc.u32 %8, %20, %20, %32;\n\tmadc.hi.cc.u32 %9, %20, %20, %33;\n\t"
        "madc.lo.cc.u32 %10, %21, %21, %34;\n\tmadc.hi.cc.u32 %11, %21, %21, %35;\n\t"
        "madc.lo.cc.u32 %12, %22, %22, %36;\n\tmadc.hi.cc.u32 %13, %22, %22, %37;\n\t"
        "madc.lo.cc.u32 %14, %23, %23, %38;\n\tmadc.hi.u32 %15, %23, %23, %39;"
        : "=&r"(r[0]), "=&r"(r[1]), "=&r"(r[2]), "=&r"(r[3]), "=&r"(r[4]), "=&r"(r[5]), "=&r"(r[6]), "=&r"(r[7]), "=&r"(r[8]), "=&r"(r[9]),
          "=&r"(r[10]), "=&r"(r[11]), "=&r"(r[12]), "=&r"(r[13]), "=&r"(r[14]), "=&r"(r[15])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(T[0]), "r"(T[1]), "r"(T[2]), "r"(T[3]),
          "r"(T[4]), "r"(T[5]), "r"(T[6]), "r"(T[7]), "r"(T[8]), "r"(T[9]), "r"(T[10]), "r"(T[11]), "r"(T[12]), "r"(T[13]), "r"(T[14]), "r"(T[15]));
#else
    uint32_t D[16];
    for (int i = 0; i < 8; i++) wide_mul(D[2 * i], D[2 * i + 1], a[i], a[i]);
    add16(r, T, D);
#endif
}
BCP_FE_FN fe fe_sqr_f(fe a) {
    uint32_t t[16];
    fe r;
    sqr_wide(t, a.v);
    fe_reduce_wide(r, t);
    return r;
}
BCP_DEV void fe_sqr(fe &r, const fe &a) { r = fe_sqr_f(a); }

// ------------------------------------------------------------------ add / sub / neg (weakly reduced in and out)
// The folds after the first chain almost never propagate: c * (2^32 + 977) lands in limbs 0 and 1, and only a carry (borrow)
// out of limb 1 — probability about 2^-32 — can reach further or wrap a second time.  That case is finished by these
// out-of-line full chains; the inline path is one 8-limb chain plus a 2-limb one.
BCP_DEV_NOINLINE fe fe_add_fold_slow(fe r, uint32_t c) { /* by value: the element travels in registers */
    uint32_t A[8] = {(0u - c) & BCP_FOLD, c, 0, 0, 0, 0, 0, 0};
    uint32_t c2 = addn_cc<8>(r.v, r.v, A);
    // a second wrap needs both inputs within 2^33 of 2^256; afterwards the value is < 2^34
    uint32_t B[2] = {(0u - c2) & BCP_FOLD, c2};
    addn_cc<2>(r.v, r.v, B);
    return r;
}
BCP_DEV_NOINLINE fe fe_sub_fold_slow(fe r, uint32_t bw) {
    uint32_t A[8] = {(0u - bw) & BCP_FOLD, bw, 0, 0, 0, 0, 0, 0};
    uint32_t bw2 = subn_cc<8>(r.v, r.v, A);
    // second borrow: value is now within 2^34 of 2^256, subtracting again cannot borrow
    uint32_t B[2] = {(0u - bw2) & BCP_FOLD, bw2};
    subn_cc<2>(r.v, r.v, B);
    return r;
}
BCP_DEV void fe_add(fe &r, const fe &a, const fe &b) {
    uint32_t c = addn_cc<8>(r.v, a.v, b.v);
    // a + b - 2^256 == a + b - 2^256 + (2^32 + 977) (mod p)
#if BCP_FE_SHORT_FOLD
    uint32_t A2[2] = {(0u - c) & BCP_FOLD, c}, lo2[2];
    if (__builtin_expect(addn_cc<2>(lo2, r.v, A2) == 0, 1)) {
        r.v[0] = lo2[0];
        r.v[1] = lo2[1];
        return;
    }
#endif
    r = fe_add_fold_slow(r, c);
}
BCP_DEV void fe_sub(fe &r, const fe &a, const fe &b) {
    uint32_t bw = subn_cc<8>(r.v, a.v, b.v);
    // a - b + 2^256 == a - b + 2^256 - (2^32 + 977) (mod p)
#if BCP_FE_SHORT_FOLD
    uint32_t A2[2] = {(0u - bw) & BCP_FOLD, bw}, lo2[2];
    if (__builtin_expect(subn_cc<2>(lo2, r.v, A2) == 0, 1)) {
        r.v[0] = lo2[0];
        r.v[1] = lo2[1];
        return;
    }
#endif
    r = fe_sub_fold_slow(r, bw);
}
// the always-full-chain forms (what fe_add / fe_sub were before the short cut): reference for the device self-test
BCP_DEV void fe_add_full(fe &r, const fe &a, const fe &b) {
    uint32_t c = addn_cc<8>(r.v, a.v, b.v);
    r = fe_add_fold_slow(r, c);
}
BCP_DEV void fe_sub_full(fe &r, const fe &a, const fe &b) {
    uint32_t bw = subn_cc<8>(r.v, a.v, b.v);
    r = fe_sub_fold_slow(r, bw);
}
BCP_DEV void fe_set_zero(fe &r) {
    BCP_UNROLL
    for (int i = 0; i < 8; i++) r.v[i] = 0;
}
BCP_DEV void fe_set_u32(fe &r, uint32_t x) {
    fe_set_zero(r);
    r.v[0] = x;
}
BCP_DEV void fe_neg(fe &r, const fe &a) {
    fe z;
    fe_set_zero(z);
    fe_sub(r, z, a);
}
BCP_DEV void fe_dbl(fe &r, const fe &a) { fe_add(r, a, a); }

// ------------------------------------------------------------------ quick forms (the point kernel's hot loop)
// Same results as fe_mul / fe_sqr / fe_add / fe_sub / fe_neg, without their out-of-line branch: the carry that would
// send those to the full chains (out of limb 2 in the reduction, out of limb 1 in an addition: probability about 2^-30
// and 2^-33 per operation, two signatures in a million hit one somewhere) is COUNTED in `bad` and the result is then not
// reduced correctly.  The caller checks `bad` once at the end and redoes the whole computation with the exact forms
// above when it is not zero.  What this buys: three to four instructions per field operation (branch, reconvergence
// pair, flag test) — every instruction of the point kernel costs an issue slot next to the multiplier instructions
// (tools/ubench/pipe_mix.cu).
BCP_DEV void fe_reduce_wide_q(fe &r, const uint32_t *t /*16*/, uint32_t &bad) {
    // first fold as in fe_reduce_wide_t: S = L + H*977 + (H << 32) in two interleaved arrays
    uint32_t EV[9], OD[9];
    BCP_UNROLL
    for (int k = 0; k < 8; k++) { EV[k] = t[k]; OD[k] = t[8 + k]; }
    EV[8] = 0; OD[8] = 0;
    row_acc9(EV, t[8], t[10], t[12], t[14], BCP_FOLD);
    row_acc9(OD, t[9], t[11], t[13], t[15], BCP_FOLD);
    uint32_t s[10];
    s[0] = EV[0];
#if BCP_PTX
    // s[1..8] = EV[1..8] + OD[0..7]; s[9] = OD[8] + carry (<= 3)
    asm("add.cc.u32 %0, %9, %17;\n\taddc.cc.u32 %1, %10, %18;\n\taddc.cc.u32 %2, %11, %19;\n\taddc.cc.u32 %3, %12, %20;\n\t"
        "addc.cc.u32 %4, %13, %21;\n\taddc.cc.u32 %5, %14, %22;\n\taddc.cc.u32 %6, %15, %23;\n\taddc.cc.u32 %7, %16, %24;\n\t"
        "addc.u32 %8, %25, 0;"
        : "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=&r"(s[5]), "=&r"(s[6]), "=&r"(s[7]), "=&r"(s[8]), "=r"(s[9])
        : "r"(EV[1]), "r"(EV[2]), "r"(EV[3]), "r"(EV[4]), "r"(EV[5]), "r"(EV[6]), "r"(EV[7]), "r"(EV[8]), "r"(OD[0]), "r"(OD[1]),
          "r"(OD[2]), "r"(OD[3]), "r"(OD[4]), "r"(OD[5]), "r"(OD[6]), "r"(OD[7]), "r"(OD[8]));
#else
    s[9] = OD[8] + addn_cc<8>(s + 1, EV + 1, OD);
#endif
    // second fold: T = s[9]:s[8] < 2^34;  S' = s[0..7] + T*977 + (T << 32) reaches limbs 0..2 only
    uint32_t mlo, mhi;
#if BCP_PTX
    asm("{\n\t.reg .u64 m;\n\tmul.wide.u32 m, %2, 977;\n\tmov.b64 {%0, %1}, m;\n\t}" : "=r"(mlo), "=r"(mhi) : "r"(s[8])); /* one IMAD.WIDE */
#else
    wide_mul(mlo, mhi, s[8], BCP_FOLD);
#endif
    const uint32_t h = s[9] * BCP_FOLD + mhi; /* < 977 + 3 * 977: no overflow */
#if BCP_PTX
    uint32_t r0, r1, r2;
    asm("{\n\t.reg .u32 a1, a2;\n\t"
        "add.cc.u32 a1, %4, %5;\n\taddc.u32 a2, %6, 0;\n\t"                          /* a2:a1 = h + s8 + (s9 << 32) */
        "add.cc.u32 %0, %7, %10;\n\taddc.cc.u32 %1, %8, a1;\n\taddc.cc.u32 %2, %9, a2;\n\t"
        "addc.u32 %3, %3, 0;\n\t}"                                                      /* the carry out of limb 2 is counted */
        : "=&r"(r0), "=&r"(r1), "=&r"(r2), "+r"(bad)
        : "r"(h), "r"(s[8]), "r"(s[9]), "r"(s[0]), "r"(s[1]), "r"(s[2]), "r"(mlo));
    r.v[0] = r0; r.v[1] = r1; r.v[2] = r2;
#else
    const uint64_t a1 = (uint64_t)h + s[8];
    const uint32_t A[3] = {mlo, (uint32_t)a1, s[9] + (uint32_t)(a1 >> 32)};
    uint32_t lo3[3];
    bad += addn_cc<3>(lo3, s, A);
    r.v[0] = lo3[0]; r.v[1] = lo3[1]; r.v[2] = lo3[2];
#endif
    BCP_UNROLL
    for (int i = 3; i < 8; i++) r.v[i] = s[i];
}
BCP_DEV void fe_mul_q(fe &r, const fe &a, const fe &b, uint32_t &bad) {
    uint32_t t[16];
    mul_wide(t, a.v, b.v);
    fe_reduce_wide_q(r, t, bad);
}
BCP_DEV void fe_sqr_q(fe &r, const fe &a, uint32_t &bad) {
    uint32_t t[16];
    sqr_wide(t, a.v);
    fe_reduce_wide_q(r, t, bad);
}
BCP_DEV void fe_add_q(fe &r, const fe &a, const fe &b, uint32_t &bad) {
#if BCP_PTX
    uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
    asm("{\n\t.reg .u32 c, k;\n\t.reg .pred p;\n\t"
        "add.cc.u32 %0, %9, %17;\n\taddc.cc.u32 %1, %10, %18;\n\taddc.cc.u32 %2, %11, %19;\n\taddc.cc.u32 %3, %12, %20;\n\t"
        "addc.cc.u32 %4, %13, %21;\n\taddc.cc.u32 %5, %14, %22;\n\taddc.cc.u32 %6, %15, %23;\n\taddc.cc.u32 %7, %16, %24;\n\t"
        "addc.u32 c, 0, 0;\n\t"
        "setp.ne.u32 p, c, 0;\n\tselp.u32 k, 977, 0, p;\n\t" /* + c * (2^32 + 977) */
        "add.cc.u32 %0, %0, k;\n\taddc.cc.u32 %1, %1, c;\n\taddc.u32 %8, %8, 0;\n\t}"
        : "=&r"(r0), "=&r"(r1), "=&r"(r2), "=&r"(r3), "=&r"(r4), "=&r"(r5), "=&r"(r6), "=&r"(r7), "+r"(bad)
        : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]), "r"(b.v[0]),
          "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
    r.v[0] = r0; r.v[1] = r1; r.v[2] = r2; r.v[3] = r3; r.v[4] = r4; r.v[5] = r5; r.v[6] = r6; r.v[7] = r7;
#else
    uint32_t t[8];
    const uint32_t c = addn_cc<8>(t, a.v, b.v);
    const uint32_t A2[2] = {(0u - c) & BCP_FOLD, c};
    uint32_t lo2[2];
    bad += addn_cc<2>(lo2, t, A2);
    r.v[0] = lo2[0]; r.v[1] = lo2[1];
    for (int i = 2; i < 8; i++) r.v[i] = t[i];
#endif
}
BCP_DEV void fe_sub_q(fe &r, const fe &a, const fe &b, uint32_t &bad) {
#if BCP_PTX
    uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
    asm("{\n\t.reg .u32 c, k;\n\t"
        "sub.cc.u32 %0, %9, %17;\n\tsubc.cc.u32 %1, %10, %18;\n\tsubc.cc.u32 %2, %11, %19;\n\tsubc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\tsubc.cc.u32 %5, %14, %22;\n\tsubc.cc.u32 %6, %15, %23;\n\tsubc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 c, 0, 0;\n\t"                                  /* 0 or 0xFFFFFFFF (only subc reads a borrow) */
        "and.b32 k, c, 977;\n\tshr.u32 c, c, 31;\n\t"          /* - borrow * (2^32 + 977) */
        "sub.cc.u32 %0, %0, k;\n\tsubc.cc.u32 %1, %1, c;\n\tsubc.u32 c, 0, 0;\n\t"
        "sub.u32 %8, %8, c;\n\t}"                                /* c is 0 or -1: counts the borrow out of limb 1 */
        : "=&r"(r0), "=&r"(r1), "=&r"(r2), "=&r"(r3), "=&r"(r4), "=&r"(r5), "=&r"(r6), "=&r"(r7), "+r"(bad)
        : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]), "r"(b.v[0]),
          "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
    r.v[0] = r0; r.v[1] = r1; r.v[2] = r2; r.v[3] = r3; r.v[4] = r4; r.v[5] = r5; r.v[6] = r6; r.v[7] = r7;
#else
    uint32_t t[8];
    const uint32_t bw = subn_cc<8>(t, a.v, b.v);
    const uint32_t A2[2] = {(0u - bw) & BCP_FOLD, bw};
    uint32_t lo2[2];
    bad += subn_cc<2>(lo2, t, A2);
    r.v[0] = lo2[0]; r.v[1] = lo2[1];
    for (int i = 2; i < 8; i++) r.v[i] = t[i];
#endif
}
BCP_DEV void fe_dbl_q(fe &r, const fe &a, uint32_t &bad) { fe_add_q(r, a, a, bad); }
// r = 8a: shift by three bits, the three bits that leave the top limb come back times 2^32 + 977
BCP_DEV void fe_mul8_q(fe &r, const fe &a, uint32_t &bad) {
    const uint32_t top = a.v[7] >> 29;
    uint32_t s[8];
    BCP_UNROLL
    for (int i = 7; i > 0; i--) s[i] = (a.v[i] << 3) | (a.v[i - 1] >> 29);
    s[0] = a.v[0] << 3;
    const uint32_t A2[2] = {top * BCP_FOLD, top};
    uint32_t lo2[2];
    bad += addn_cc<2>(lo2, s, A2);
    r.v[0] = lo2[0]; r.v[1] = lo2[1];
    BCP_UNROLL
    for (int i = 2; i < 8; i++) r.v[i] = s[i];
}
// r = -a = p - a for a <= p; an element above p (possible for a weakly reduced value, never seen on random data) borrows
// out of the top limb and is counted in `bad`
BCP_DEV void fe_neg_q(fe &r, const fe &a, uint32_t &bad) {
    const uint32_t P[8] = {BCP_P0, BCP_P1, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
    bad += subn_cc<8>(r.v, P, a.v);
}

// canonical representative in [0, p)
BCP_DEV void fe_normalize(fe &a) {
    // a >= p  <=>  a + (2^32 + 977) >= 2^256
    uint32_t A[8] = {BCP_FOLD, 1, 0, 0, 0, 0, 0, 0}, t[8];
    uint32_t c = addn_cc<8>(t, a.v, A);
    BCP_UNROLL
    for (int i = 0; i < 8; i++) a.v[i] = c ? t[i] : a.v[i];
}
// value == 0 (mod p) for a weakly reduced element: a == 0 or a == p
BCP_DEV bool fe_is_zero(const fe &a) {
    uint32_t z = a.v[0] | a.v[1] | a.v[2] | a.v[3] | a.v[4] | a.v[5] | a.v[6] | a.v[7];
    uint32_t q = (a.v[0] ^ BCP_P0) | (a.v[1] ^ BCP_P1) | ~(a.v[2] & a.v[3] & a.v[4] & a.v[5] & a.v[6] & a.v[7]);
    return z == 0 || q == 0;
}
BCP_DEV bool fe_equal(const fe &a, const fe &b) {
    fe d;
    fe_sub(d, a, b);
    return fe_is_zero(d);
}
BCP_DEV bool fe_is_odd_normalized(const fe &a) { return a.v[0] & 1u; }

// big-endian 32 bytes -> limbs; returns false when the value is >= p (fe_set_b32, field_10x26_impl.h:323-338)
BCP_DEV bool fe_set_b32(fe &r, const uint8_t *b) {
    BCP_UNROLL
    for (int i = 0; i < 8; i++) {
        const uint8_t *q = b + 4 * (7 - i);
        r.v[i] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
    }
    bool top = (r.v[2] & r.v[3] & r.v[4] & r.v[5] & r.v[6] & r.v[7]) == 0xFFFFFFFFu;
    bool ge = top && (r.v[1] == 0xFFFFFFFFu || (r.v[1] == BCP_P1 && r.v[0] >= BCP_P0));
    return !ge;
}

// r = a^(2^n)
BCP_DEV void fe_sqrn(fe &r, const fe &a, int n) {
    r = a;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int i = 0; i < n; i++) fe_sqr(r, r);
}
// candidate square root a^((p+1)/4); the caller checks r^2 == a (fe_sqrt, field_impl.h:38-134).
// (p+1)/4 = 2^254 - 2^30 - 244 has the bit pattern 1^223 0 1^22 0000 11 00.  With x_k = a^(2^k - 1) the chain is
//   x2 x3 x6 x9 x11 x22 x44 x88 x176 x220 x223, then ((x223 << 23) x22 << 6) x2 << 2
// i.e. 14 steps "square n times, multiply by a saved value".  It is driven from a table in ONE rolled loop so
// that the kernel holds a single squaring and a single multiplication instance for it (code size, not
// arithmetic, was the top stall reason of the fully unrolled version: see profiles/r01_*).
BCP_DEV void fe_sqrt_candidate(fe &r, const fe &a) {
    // saved values: 0:a 1:x2 2:x3 3:x11 4:x22 5:x44 6:x88 ; 7 = "do not save"/"no multiply"
    fe sv[7];
    sv[0] = a;
    fe cur = a;
#if BCP_PTX
#pragma unroll 1
#endif
    for (int step = 0; step < 14; step++) {
        int nsq, mi, si;
        switch (step) {
            case 0: nsq = 1; mi = 0; si = 1; break;   /* x2  */
            case 1: nsq = 1; mi = 0; si = 2; break;   /* x3  */
            case 2: nsq = 3; mi = 2; si = 7; break;   /* x6  */
            case 3: nsq = 3; mi = 2; si = 7; break;   /* x9  */
            case 4: nsq = 2; mi = 1; si = 3; break;   /* x11 */
            case 5: nsq = 11; mi = 3; si = 4; break;  /* x22 */
            case 6: nsq = 22; mi = 4; si = 5; break;  /* x44 */
            case 7: nsq = 44; mi = 5; si = 6; break;  /* x88 */
            case 8: nsq = 88; mi = 6; si = 7; break;  /* x176 */
            case 9: nsq = 44; mi = 5; si = 7; break;  /* x220 */
            case 10: nsq = 3; mi = 2; si = 7; break;  /* x223 */
            case 11: nsq = 23; mi = 4; si = 7; break;
            case 12: nsq = 6; mi = 1; si = 7; break;
            default: nsq = 2; mi = 7; si = 7; break;
        }
#if BCP_PTX
#pragma unroll 1
#endif
        for (int i = 0; i < nsq; i++) fe_sqr(cur, cur);
        if (mi != 7) fe_mul(cur, cur, sv[mi]);
        if (si != 7) sv[si] = cur;
    }
    r = cur;
}
// a^(p-2) (only used by one-time table generation)
BCP_DEV void fe_inv(fe &r, const fe &a) {
    // p - 2 = 1^223 0 1^22 0000 101101
    fe x2, x3, x6, x9, x11, x22, x44, x88, x176, x220, x223, t;
    fe_sqr(t, a); fe_mul(x2, t, a);
    fe_sqr(t, x2); fe_mul(x3, t, a);
    fe_sqrn(t, x3, 3); fe_mul(x6, t, x3);
    fe_sqrn(t, x6, 3); fe_mul(x9, t, x3);
    fe_sqrn(t, x9, 2); fe_mul(x11, t, x2);
    fe_sqrn(t, x11, 11); fe_mul(x22, t, x11);
    fe_sqrn(t, x22, 22); fe_mul(x44, t, x22);
    fe_sqrn(t, x44, 44); fe_mul(x88, t, x44);
    fe_sqrn(t, x88, 88); fe_mul(x176, t, x88);
    fe_sqrn(t, x176, 44); fe_mul(x220, t, x44);
    fe_sqrn(t, x220, 3); fe_mul(x223, t, x3);
    fe_sqrn(t, x223, 23); fe_mul(t, t, x22);
    fe_sqrn(t, t, 5); fe_mul(t, t, a);
    fe_sqrn(t, t, 3); fe_mul(t, t, x2);
    fe_sqrn(t, t, 2); fe_mul(r, t, a);
}
