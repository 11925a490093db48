// bcp_ecdsa.cu — batched secp256k1 ECDSA verification kernels for sm_100a and their C-ABI entry
// points (include/bcp_gpu.h).  The arithmetic lives in fe.cuh / scalar.cuh / ge.cuh / ecdsa.cuh.
//
// Launch shape
//   k_ecdsa_prep   one thread per K signatures (K = 8, 4 for small batches; thread t takes signatures t, t+T, t+2T, ...
//                  so that the byte loads of a warp stay adjacent), 128 threads per block; one scalar inversion per
//                  thread by division steps (scalar.cuh)
//   k_ecdsa_point  one thread per signature, 256 threads per block, two blocks per SM (128 registers); the per-thread
//                  {1..8}Q tables (768 B: x, y, beta*x per entry) live in global scratch next to the phase-1 records (the first
//                  version kept them in shared memory, which capped occupancy at 12 warps per SM).  The walk uses the
//                  branch-free "quick" field forms (fe.cuh); a signature whose walk left a fold unfinished is walked again
//                  over the exact forms
// A batch is processed in slices of BCP_SLICE signatures (16 full waves) so that scratch is bounded (718 MB).
//   k_gcomb_build  one thread per comb-table entry, once per device at bcp_gpu_init
#include "bcp_ctx.cuh"
#ifndef BCP_POINT_VARIANT
#define BCP_POINT_VARIANT 1 /* 0: field mul/sqr as called functions; 1: fully inlined + one block barrier per step */
#endif
#if BCP_POINT_VARIANT == 1
#define BCP_FE_CALLS 0
/* the block barrier is there for instruction-cache locality only; measured on B200 at 1M signatures (tools/sync_sweep.sh):
 * every step 46.74 M/s, every 2nd 46.95, 3rd 47.07, 4th 47.15, 6th 47.03, 12th 46.71, 24th 46.10 */
#ifndef BCP_STEP_SYNC_EVERY
#define BCP_STEP_SYNC_EVERY 4
#endif
/* a named barrier over the sync_n threads that walk the steps together: in the last, partly filled round of the persistent
 * kernel only the warps that hold signatures take part, the others have left */
#define BCP_STEP_SYNC() do { if (sync_n != 0 && (BCP_STEP_SYNC_EVERY == 1 || (t % BCP_STEP_SYNC_EVERY) == 0)) asm volatile("bar.sync 1, %0;" :: "r"(sync_n) : "memory"); } while (0)
#ifndef BCP_STEP_SYNC_EVERY_V2
#define BCP_STEP_SYNC_EVERY_V2 4 /* iterations of ecdsa_point_phase_v2 between barriers (4 = two windows); B200, 1M: every iteration 51.91 M/s, 2nd 52.04, 4th 52.23 */
#endif
#define BCP_STEP_SYNC_V2() do { if (sync_n != 0 && (BCP_STEP_SYNC_EVERY_V2 == 1 || (t % BCP_STEP_SYNC_EVERY_V2) == 0)) asm volatile("bar.sync 1, %0;" :: "r"(sync_n) : "memory"); } while (0)
#ifndef BCP_POINT_PHASE
#define BCP_POINT_PHASE 2 /* 1: ecdsa_point_phase (one group operation per step, 51.23 M/s); 2: ecdsa_point_phase_v2 (addition + doublings per iteration, 52.23 M/s) */
#endif
#endif
#include "ecdsa.cuh"
#include <cstdlib>
#include <cstring>
#include <vector>

#if BCP_POINT_VARIANT == 1
#ifndef BCP_POINT_THREADS
#define BCP_POINT_THREADS 256
#endif
#ifndef BCP_POINT_BLOCKS_PER_SM
#define BCP_POINT_BLOCKS_PER_SM (512 / BCP_POINT_THREADS)
#endif
#else
#define BCP_POINT_THREADS 128
#define BCP_POINT_BLOCKS_PER_SM 4
#endif
/* signatures per slice: 16 full waves of the point kernel on a 148-SM part (148 SMs x 512 resident threads) */
#define BCP_WAVE ((size_t)148 * 512)
#ifndef BCP_POINT_SMALL_TPB
#define BCP_POINT_SMALL_TPB 128u
#endif
#define BCP_SLICE (16 * BCP_WAVE)
#define BCP_E2E_CHUNK (6 * BCP_WAVE)

__global__ void __launch_bounds__(128) k_gcomb_build(uint32_t *gcomb) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= BCP_GCOMB_WINDOWS * BCP_GCOMB_ENTRIES) return;
    int j = idx / BCP_GCOMB_ENTRIES;
    uint32_t d = idx % BCP_GCOMB_ENTRIES;
    uint32_t *out = gcomb + (size_t)idx * 16;
    if (d == 0) {
        for (int i = 0; i < 16; i++) out[i] = 0;
        return;
    }
    uint32_t e[16];
    gcomb_entry(e, j, d);
    for (int i = 0; i < 16; i++) out[i] = e[i];
}

template <int K>
__global__ void __launch_bounds__(128) k_ecdsa_prep(const uint8_t *__restrict__ msg32, const uint8_t *__restrict__ sig64,
                                                    size_t n, uint32_t flags, uint32_t *__restrict__ rec) {
    const size_t T = (size_t)gridDim.x * blockDim.x;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    // Montgomery's trick over this thread's signatures: pre[j] = s_0 * ... * s_{j-1}.  Only the prefix products
    // are kept (local memory); the second pass re-parses each signature, which is far cheaper than holding
    // K x (r, s, m) per thread.  One inversion (division steps, scalar.cuh) per K signatures.
    scalar pre[K];
    scalar acc;
#pragma unroll
    for (int i = 0; i < 8; i++) acc.v[i] = (i == 0);
    int cnt = 0;
#pragma unroll 1
    for (int j = 0; j < K; j++) {
        size_t i = t + (size_t)j * T;
        if (i >= n) break;
        sig_scalars g;
        sig_load(g, msg32 + 32 * i, sig64 + 64 * i, flags);
        pre[j] = acc;
        sc_mul(acc, acc, g.s);
        cnt++;
    }
    scalar inv;
    sc_inverse_divsteps(inv, acc);
#pragma unroll 1
    for (int j = cnt - 1; j >= 0; j--) {
        size_t i = t + (size_t)j * T;
        sig_scalars g;
        sig_load(g, msg32 + 32 * i, sig64 + 64 * i, flags);
        scalar sinv;
        sc_mul(sinv, inv, pre[j]);
        sc_mul(inv, inv, g.s);
        uint32_t r[BCP_REC_WORDS];
        sig_finish(r, g, sinv);
        uint4 *dst = reinterpret_cast<uint4 *>(rec + i * BCP_REC_WORDS);
#pragma unroll
        for (int q = 0; q < BCP_REC_WORDS / 4; q++) dst[q] = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
    }
}

__global__ void __launch_bounds__(BCP_POINT_THREADS, BCP_POINT_BLOCKS_PER_SM)
    k_ecdsa_point(const uint8_t *__restrict__ pub65, const uint8_t *__restrict__ sig64, size_t n,
                  const uint32_t *__restrict__ rec, uint32_t *__restrict__ qtabs, const uint32_t *__restrict__ gcomb,
                  uint8_t *__restrict__ ok, uint32_t redo_every) {
    const size_t gi = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = gi < n ? gi : n - 1; /* tail threads shadow the last signature (they must reach the barriers) */
    qtable_packed tab;
    tab.base = qtabs + gi * BCP_QTAB_WORDS; /* own slab even when shadowing: scratch is sized for whole blocks */
    // The walk with the branch-free field operations; `bad` counts the folds they left unfinished (about two signatures
    // in a million have one).  Such a signature is walked again with the exact operations, alone: no barrier.
    // redo_every (test hook, BCP_POINT_FORCE_REDO): every redo_every-th signature is treated as if it had one.
    uint32_t bad = 0;
#if BCP_POINT_PHASE == 2
    bool v = ecdsa_point_phase_v2<qtable_packed, false, fe_quick>(rec + i * BCP_REC_WORDS, pub65 + 65 * i, sig64 + 64 * i, tab, gcomb,
                                                                  (int)blockDim.x, nullptr, &bad);
#else
    bool v = ecdsa_point_phase<qtable_packed, false, fe_quick>(rec + i * BCP_REC_WORDS, pub65 + 65 * i, sig64 + 64 * i, tab, gcomb,
                                                               (int)blockDim.x, nullptr, &bad);
#endif
    if (redo_every != 0 && gi % redo_every == 0) bad = 1;
    if (__builtin_expect(bad != 0, 0))
        v = ecdsa_point_phase<qtable_packed, false, fe_exact>(rec + i * BCP_REC_WORDS, pub65 + 65 * i, sig64 + 64 * i, tab, gcomb, 0);
    if (gi < n) ok[gi] = v ? 1 : 0;
}

// Persistent form: the grid is exactly the resident set (BCP_POINT_BLOCKS_PER_SM blocks on every SM); a block CLAIMS
// whole chunks of blockDim.x signatures from a counter until none is left, then takes its share of the remainder, so that
//   * a thread's {1..8}Q slab belongs to the RESIDENT SLOT (block, thread), not to the signature: the slabs of a whole
//     launch are 148 x 512 x 768 B = 58 MB instead of 768 B of scratch per signature;
//   * claiming (not a fixed stride) keeps every SM at two blocks to the end: the two blocks of an SM do not run at the same
//     speed (with a fixed assignment the faster one ran out of rounds early and left the SM half empty for the rest of the
//     launch: ncu showed 13.8 instead of 15.7 active warps per SM and the launch was 6 % slower);
//   * the remainder (less than one chunk per block) is dealt out EVENLY: every block takes per_last signatures (a multiple
//     of 32) and only the warps that hold some walk the steps, so what is left over runs on every SM at a few warps each
//     (and correspondingly faster) instead of as whole 256-thread blocks on a fifth of the chip.
// `counter` must be zero at launch.
__global__ void __launch_bounds__(BCP_POINT_THREADS, BCP_POINT_BLOCKS_PER_SM)
    k_ecdsa_point_rounds(const uint8_t *__restrict__ pub65, const uint8_t *__restrict__ sig64, uint32_t n,
                         const uint32_t *__restrict__ rec, uint32_t *__restrict__ qtabs, const uint32_t *__restrict__ gcomb,
                         uint8_t *__restrict__ ok, uint32_t per_last, uint32_t *__restrict__ counter) {
    // Everything a chunk needs is recomputed from the kernel parameters (constant bank) and the chunk number, which lives
    // in shared memory: the point phase itself uses all 128 registers.
    __shared__ uint32_t s_chunk;
    const uint32_t full = (n / (gridDim.x * blockDim.x)) * gridDim.x; /* whole chunks handed out by the counter */
    qtable_packed tab; /* slabs are packed per block: blockDim.x per block once there is a full chunk, else per_last */
    tab.base = qtabs + ((size_t)blockIdx.x * (full ? blockDim.x : per_last) + threadIdx.x) * BCP_QTAB_WORDS;
#pragma unroll 1
    for (;;) {
        if (threadIdx.x == 0) s_chunk = atomicAdd(counter, 1u);
        __syncthreads(); /* every thread of the block is still here: threads only leave in the last pass */
        uint32_t i, width;
        {
            const uint32_t chunks = (n / (gridDim.x * blockDim.x)) * gridDim.x;
            const bool last = s_chunk >= chunks; /* no whole chunk left: this block's share of the remainder, then out */
            const uint32_t base = last ? chunks * blockDim.x + blockIdx.x * per_last : s_chunk * blockDim.x;
            if (base >= n) return; /* uniform for the block */
            const uint32_t left = last ? (n - base < per_last ? n - base : per_last) : blockDim.x;
            width = ((left + 31u) >> 5) * 32u;
            if (threadIdx.x >= width) return; /* whole warps without a signature leave (last pass only) */
            i = threadIdx.x < left ? base + threadIdx.x : n - 1u; /* lanes past the end shadow the last signature (barriers) */
        }
        uint32_t bad = 0; /* as in k_ecdsa_point: the quick walk, and the exact one behind it when a fold was left unfinished */
        bool v = ecdsa_point_phase_v2<qtable_packed, false, fe_quick>(rec + (size_t)i * BCP_REC_WORDS, pub65 + (size_t)65 * i,
                                                                      sig64 + (size_t)64 * i, tab, gcomb, (int)width, nullptr, &bad);
        if (__builtin_expect(bad != 0, 0))
            v = ecdsa_point_phase<qtable_packed, false, fe_exact>(rec + (size_t)i * BCP_REC_WORDS, pub65 + (size_t)65 * i,
                                                                  sig64 + (size_t)64 * i, tab, gcomb, 0);
        {
            const uint32_t chunks = (n / (gridDim.x * blockDim.x)) * gridDim.x;
            const bool last = s_chunk >= chunks;
            const uint32_t base = last ? chunks * blockDim.x + blockIdx.x * per_last : s_chunk * blockDim.x;
            const uint32_t left = last ? (n - base < per_last ? n - base : per_last) : blockDim.x;
            if (threadIdx.x < left) ok[base + threadIdx.x] = v ? 1 : 0;
            if (last) return;
        }
        __syncthreads(); /* s_chunk is read above until the very end of the chunk: thread 0 may only replace it now */
    }
}

// secp256k1_ec_pubkey_parse for a batch: affine x || y (big-endian) per key, ok byte per key
__global__ void __launch_bounds__(128) k_pubkey_parse(const uint8_t *__restrict__ pub65, size_t n, uint8_t *__restrict__ xy64,
                                                      uint8_t *__restrict__ ok) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    fe x, y;
    bool v = pubkey_parse(x, y, pub65 + 65 * i);
    fe_normalize(x);
    fe_normalize(y);
    uint8_t *o = xy64 + 64 * i;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        uint32_t a = v ? x.v[7 - k] : 0u, b = v ? y.v[7 - k] : 0u;
        o[4 * k] = a >> 24; o[4 * k + 1] = a >> 16; o[4 * k + 2] = a >> 8; o[4 * k + 3] = a;
        o[32 + 4 * k] = b >> 24; o[32 + 4 * k + 1] = b >> 16; o[32 + 4 * k + 2] = b >> 8; o[32 + 4 * k + 3] = b;
    }
    ok[i] = v ? 1 : 0;
}

// R = u2*P + u1*G as an affine point, for the two verify-side libsecp256k1 calls that are not a verification:
//   mode 1  secp256k1_ecdsa_recover (src/secp256k1/src/modules/recovery/main_impl.h:87-121, 170-190): P = the point with
//           x = r (+ n) and the parity of the recovery id (built by the host as a compressed key), u2 = s/r, u1 = -m/r
//   mode 0  secp256k1_ec_pubkey_tweak_add (src/secp256k1/src/secp256k1.c:477-499, eckey_impl.h:63-75): u2 = 1, u1 = tweak
// One thread per item; the walk is the verification's (ecdsa_point_phase), so the GLV split, the {1..8}P table and the
// comb for G are shared code.  Not a hot path in the node (message signatures, BIP32 public derivation): batches of one
// through the compat API, but the kernel takes any n.
__global__ void __launch_bounds__(128) k_ec_lincomb(const uint8_t *__restrict__ pub65, const uint8_t *__restrict__ in_a,
                                                    const uint8_t *__restrict__ msg32, int mode, size_t n,
                                                    uint32_t *__restrict__ qtabs, const uint32_t *__restrict__ gcomb,
                                                    uint8_t *__restrict__ xy64, uint8_t *__restrict__ ok) {
    const size_t gi = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = gi < n ? gi : n - 1; /* tail threads shadow the last item (they must reach the barriers) */
    sig_scalars g;
    scalar sinv;
    BCP_UNROLL
    for (int k = 0; k < 8; k++) { sinv.v[k] = (k == 0); g.r.v[k] = (k == 0); g.s.v[k] = (k == 0); }
    bool valid;
    if (mode == 1) {
        scalar r, s, m;
        const bool ov_r = sc_set_b32(r, in_a + 64 * i), ov_s = sc_set_b32(s, in_a + 64 * i + 32);
        sc_set_b32(m, msg32 + 32 * i); /* reduced mod n silently (main_impl.h:184) */
        valid = !(ov_r || ov_s) && !sc_is_zero(r) && !sc_is_zero(s); /* parse_compact overflow; :97-99 */
        if (!valid) {
            BCP_UNROLL
            for (int k = 0; k < 8; k++) r.v[k] = (k == 0);
        }
        sc_inverse_divsteps(sinv, r);
        g.r = s;            /* u2 = s / r  */
        sc_negate(g.m, m);  /* u1 = -m / r */
    } else {
        valid = !sc_set_b32(g.m, in_a + 32 * i); /* tweak >= n fails (secp256k1.c:487-488) */
    }
    g.valid = valid;
    uint32_t rec[BCP_REC_WORDS];
    sig_finish(rec, g, sinv);
    qtable_packed tab;
    tab.base = qtabs + gi * BCP_QTAB_WORDS;
    gej R;
    const bool fin = ecdsa_point_phase<qtable_packed, true>(rec, pub65 + 65 * i, nullptr, tab, gcomb, (int)blockDim.x, &R);
    if (gi >= n) return;
    fe ax, ay;
    if (fin) gej_to_affine(ax, ay, R);
    uint8_t *o = xy64 + 64 * gi;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t a = fin ? ax.v[7 - k] : 0u, b = fin ? ay.v[7 - k] : 0u;
        o[4 * k] = a >> 24; o[4 * k + 1] = a >> 16; o[4 * k + 2] = a >> 8; o[4 * k + 3] = a;
        o[32 + 4 * k] = b >> 24; o[32 + 4 * k + 1] = b >> 16; o[32 + 4 * k + 2] = b >> 8; o[32 + 4 * k + 3] = b;
    }
    ok[gi] = fin ? 1 : 0;
}

// Device self-test of the field arithmetic's rarely taken paths.  fe_add / fe_sub / fe_reduce_wide finish the folds
// that practically never propagate (probability about 2^-32 per operation) out of line; random test inputs do not reach
// that code, so crafted operands (fe_selftest_data.cuh) are run through the short-cut forms AND the always-full-chain
// forms, on the device, and must agree limb for limb.  The counterpart of ECC_InitSanityCheck (src/key.cpp) at start-up.
#include "fe_selftest_data.cuh"
__global__ void k_fe_selftest(int *fail) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    int bad = 0;
    if (t < BCP_FE_ST_EDGES * BCP_FE_ST_EDGES) {
        fe a, b, r1, r2;
        for (int i = 0; i < 8; i++) { a.v[i] = FE_ST_EDGE[t / BCP_FE_ST_EDGES][i]; b.v[i] = FE_ST_EDGE[t % BCP_FE_ST_EDGES][i]; }
        fe_add(r1, a, b);
        fe_add_full(r2, a, b);
        for (int i = 0; i < 8; i++) bad |= r1.v[i] != r2.v[i];
        fe_sub(r1, a, b);
        fe_sub_full(r2, a, b);
        for (int i = 0; i < 8; i++) bad |= r1.v[i] != r2.v[i];
        uint32_t w[16];
        mul_wide(w, a.v, b.v);
        fe_reduce_wide_t<true>(r1, w);
        fe_reduce_wide_t<false>(r2, w);
        for (int i = 0; i < 8; i++) bad |= r1.v[i] != r2.v[i];
        fe m1 = fe_mul_f(a, b); /* the product path the kernels use */
        for (int i = 0; i < 8; i++) bad |= m1.v[i] != r2.v[i];
        // a*b computed two ways must also agree after the sum of the operands went through add/sub: (a+b)-b == a (mod p)
        fe_add(r1, a, b);
        fe_sub(r1, r1, b);
        fe na = a;
        fe_normalize(r1);
        fe_normalize(na);
        for (int i = 0; i < 8; i++) bad |= r1.v[i] != na.v[i];
        // the quick forms of the point kernel (carry counted, not finished): a count of zero promises the exact value.
        // fail[2..7]: disagreements per form (add, sub, reduce, square, neg, times 8) for the error message
        uint32_t q = 0;
        int d = 0;
        fe_add_q(r1, a, b, q);
        fe_add_full(r2, a, b);
        if (q) atomicAdd(fail + 1, 1);
        else for (int i = 0; i < 8; i++) d |= r1.v[i] != r2.v[i];
        if (d) atomicAdd(fail + 2, 1);
        bad |= d; d = 0; q = 0;
        fe_sub_q(r1, a, b, q);
        fe_sub_full(r2, a, b);
        if (q) atomicAdd(fail + 1, 1);
        else for (int i = 0; i < 8; i++) d |= r1.v[i] != r2.v[i];
        if (d) atomicAdd(fail + 3, 1);
        bad |= d; d = 0; q = 0;
        mul_wide(w, a.v, b.v);
        fe_reduce_wide_q(r1, w, q);
        fe_reduce_wide_t<false>(r2, w);
        if (q) atomicAdd(fail + 1, 1);
        else for (int i = 0; i < 8; i++) d |= r1.v[i] != r2.v[i];
        if (d) atomicAdd(fail + 4, 1);
        bad |= d; d = 0; q = 0;
        sqr_wide(w, a.v);
        fe_reduce_wide_q(r1, w, q);
        fe_reduce_wide_t<false>(r2, w);
        m1 = fe_mul_f(a, a);
        fe_normalize(r2);
        fe_normalize(m1);
        for (int i = 0; i < 8; i++) d |= r2.v[i] != m1.v[i]; /* the squaring against the multiplication */
        if (!q) { fe_normalize(r1); for (int i = 0; i < 8; i++) d |= r1.v[i] != m1.v[i]; }
        if (d) atomicAdd(fail + 5, 1);
        bad |= d; d = 0; q = 0;
        fe_neg_q(r1, a, q);
        fe_neg(r2, a);
        fe_normalize(r1);
        fe_normalize(r2);
        if (q) atomicAdd(fail + 1, 1);
        else for (int i = 0; i < 8; i++) d |= r1.v[i] != r2.v[i];
        if (d) atomicAdd(fail + 6, 1);
        bad |= d; d = 0; q = 0;
        fe_mul8_q(r1, a, q);
        fe_dbl(r2, a); fe_dbl(r2, r2); fe_dbl(r2, r2);
        fe_normalize(r1);
        fe_normalize(r2);
        if (!q) for (int i = 0; i < 8; i++) d |= r1.v[i] != r2.v[i];
        if (d) atomicAdd(fail + 7, 1);
        bad |= d;
    }
    if (t < BCP_FE_ST_WIDE) {
        fe r1, r2;
        uint32_t w[16];
        for (int i = 0; i < 16; i++) w[i] = FE_ST_WIDE[t][i];
        fe_reduce_wide_t<true>(r1, w);
        fe_reduce_wide_t<false>(r2, w);
        for (int i = 0; i < 8; i++) bad |= r1.v[i] != r2.v[i];
        uint32_t q = 0;
        fe_reduce_wide_q(r1, w, q);
        if (q) atomicAdd(fail + 1, 1);
        else for (int i = 0; i < 8; i++) bad |= r1.v[i] != r2.v[i];
    }
    if (t < BCP_FE_ST_EDGES * BCP_FE_ST_EDGES) {
        // the prep kernel's inversion by division steps against a^(n-2), and a * a^-1 == 1, on products of the crafted values
        scalar x, y, i1, i2, one;
        for (int i = 0; i < 8; i++) { x.v[i] = FE_ST_EDGE[t / BCP_FE_ST_EDGES][i] ^ (0x9e3779b9u * (uint32_t)(t + 1)); y.v[i] = FE_ST_EDGE[t % BCP_FE_ST_EDGES][i]; }
        x.v[7] &= 0x7fffffffu; y.v[7] &= 0x7fffffffu; /* below n */
        sc_mul(x, x, y);
        if (t == 0) for (int i = 0; i < 8; i++) x.v[i] = (i == 0);
        if (!sc_is_zero(x)) {
            sc_inverse_divsteps(i1, x);
            sc_inverse(i2, x);
            sc_mul(one, i1, x);
            int d = 0;
            for (int i = 0; i < 8; i++) d |= (i1.v[i] != i2.v[i]) | (one.v[i] != (i == 0 ? 1u : 0u));
            if (d) atomicAdd(fail + 8, 1);
            bad |= d;
        }
    }
    if (bad) atomicAdd(fail, 1);
}
int bcp_fe_selftest(bcp_ctx *c) {
    int *d_fail = nullptr, h_fail[9] = {-1, -1, 0, 0, 0, 0, 0, 0, 0}; /* [0] disagreements; [1] how often a quick form counted a carry; [2..8] per form */
    BCP_CUDA(cudaMalloc(&d_fail, 9 * sizeof(int)));
    BCP_CUDA(cudaMemsetAsync(d_fail, 0, 9 * sizeof(int), c->stream));
    k_fe_selftest<<<(BCP_FE_ST_EDGES * BCP_FE_ST_EDGES + 127) / 128, 128, 0, c->stream>>>(d_fail);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_fail, d_fail, 9 * sizeof(int), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_fail);
    if (e != cudaSuccess) return bcp_cuda_fail(e, "field self-test");
    if (h_fail[0] != 0) {
        bcp_set_error("field arithmetic self-test failed on device %d: %d operand pairs disagree between the short-cut or quick forms "
                      "and the full carry chains (quick add %d, sub %d, reduce %d, square %d, neg %d, times-8 %d; scalar inversion %d)", c->device,
                      h_fail[0], h_fail[2], h_fail[3], h_fail[4], h_fail[5], h_fail[6], h_fail[7], h_fail[8]);
        return BCP_ERR_CUDA;
    }
    // the crafted operands exist to leave limb 1 / limb 2: the quick forms must have counted carries on many of them, and
    // not on all (a form that always counts would send every signature through the exact walk as well)
    const int cases = BCP_FE_ST_EDGES * BCP_FE_ST_EDGES * 5 + BCP_FE_ST_WIDE;
    if (h_fail[1] < 8 || h_fail[1] > cases / 2) {
        bcp_set_error("field arithmetic self-test failed on device %d: the quick forms counted a carry in %d of %d crafted cases",
                      c->device, h_fail[1], cases);
        return BCP_ERR_CUDA;
    }
    return BCP_OK;
}

int bcp_ecdsa_build_tables(bcp_ctx *c) {
    BCP_CUDA(cudaMalloc(&c->gcomb, (size_t)BCP_GCOMB_WORDS * 4));
    int total = BCP_GCOMB_WINDOWS * BCP_GCOMB_ENTRIES;
    k_gcomb_build<<<(total + 127) / 128, 128, 0, c->stream>>>(c->gcomb);
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaStreamSynchronize(c->stream));
    return BCP_OK;
}

// 0 (default): one block per 256 signatures (128 below a wave); 1: batches of more than a wave go through the persistent
// kernel with chunk claiming; 2: every batch does.  Measured on B200 (1 M signatures): 21.24 ms vs 21.61 ms — the
// persistent form needs 58 MB of table scratch instead of 768 B per signature and deals the last partial wave out evenly,
// but its two extra live values cost more in the 128-register point phase than that brings, so it is the memory-saving
// option (BCP_POINT_ROUNDS=1), not the default.
static int point_rounds_mode() {
    static const int mode = getenv("BCP_POINT_ROUNDS") ? atoi(getenv("BCP_POINT_ROUNDS")) : 0;
    return mode;
}
#define BCP_MAX_RESIDENT_BLOCKS (160 * BCP_POINT_BLOCKS_PER_SM) /* upper bound of the persistent grid on any sm_100 part */
// slabs a launch over `cap` signatures (one slice at most, whole blocks) can touch
static size_t point_slabs(size_t cap) {
    const int mode = point_rounds_mode();
    const size_t resident = (size_t)BCP_MAX_RESIDENT_BLOCKS * BCP_POINT_THREADS;
    if (mode == 0) return cap;
    if (cap > BCP_WAVE) return resident;                        /* rounds kernel with at least one full round */
    return cap + (mode == 2 ? (size_t)BCP_MAX_RESIDENT_BLOCKS * 32 : 0); /* sub-wave: per-signature, plus the warp rounding of the even split */
}
// d_scratch: bcp_gpu_ecdsa_scratch_bytes(n) bytes = [records of one slice][tables: one slab per resident thread, or per
// signature of a sub-wave batch]
int bcp_ecdsa_launch(bcp_ctx *c, const uint8_t *d_msg, const uint8_t *d_pub, const uint8_t *d_sig, size_t n,
                         uint32_t flags, uint8_t *d_ok, uint32_t *d_scratch, cudaStream_t st) {
    size_t cap = n < BCP_SLICE ? n : BCP_SLICE;
    cap = (cap + BCP_POINT_THREADS - 1) / BCP_POINT_THREADS * BCP_POINT_THREADS;
    uint32_t *d_rec = d_scratch, *d_tab = d_scratch + cap * BCP_REC_WORDS;
    const int rounds_mode = point_rounds_mode();
    for (size_t off = 0; off < n; off += BCP_SLICE) {
        const size_t m = n - off < BCP_SLICE ? n - off : BCP_SLICE;
        // signatures per prep thread: one inversion is shared by K signatures (Montgomery's trick), but a thread's work is
        // serial.  With the division-step inversion (about a tenth of a^(n-2)) small K wins everywhere; measured on B200
        // (prep + point, tools/r2_run11.sh): 1M signatures K=2 18.97 ms, 4 18.82, 8 18.76, 16 18.77, 32 18.85; 333k: 6.40,
        // 6.37, 6.38, 6.45, 6.53; one block's 41,650: 1.187, 1.188, 1.236, 1.322, 1.486.  (With a^(n-2) it was K=32 at 1M: 19.5.)
        static const int forced = getenv("BCP_PREP_K") ? atoi(getenv("BCP_PREP_K")) : 0;
        // c->prep_throughput: the caller keeps several batches in flight on this GPU (the chain connector's check lanes), so
        // what counts is the total work, not one batch's latency
        const int K = forced ? forced : (m >= 100000 || (c->prep_throughput && m >= 4096)) ? 8 : 4;
        size_t prep_threads = (m + K - 1) / K;
        unsigned prep_blocks = (unsigned)((prep_threads + 127) / 128);
        if (K >= 32) k_ecdsa_prep<32><<<prep_blocks, 128, 0, st>>>(d_msg + 32 * off, d_sig + 64 * off, m, flags, d_rec);
        else if (K >= 16) k_ecdsa_prep<16><<<prep_blocks, 128, 0, st>>>(d_msg + 32 * off, d_sig + 64 * off, m, flags, d_rec);
        else if (K >= 8) k_ecdsa_prep<8><<<prep_blocks, 128, 0, st>>>(d_msg + 32 * off, d_sig + 64 * off, m, flags, d_rec);
        else k_ecdsa_prep<4><<<prep_blocks, 128, 0, st>>>(d_msg + 32 * off, d_sig + 64 * off, m, flags, d_rec);
        // Less than about one wave: smaller thread blocks spread the signatures evenly over the SMs (the kernel ends when
        // the fullest SM does); from a wave up, 256-thread blocks keep the warps of an SM in the same code region.
        static const int forced_tpb = getenv("BCP_POINT_TPB") ? atoi(getenv("BCP_POINT_TPB")) : 0;
        if (rounds_mode == 2 || (rounds_mode == 1 && m > BCP_WAVE)) {
            const unsigned grid = (unsigned)c->sm_count * BCP_POINT_BLOCKS_PER_SM;
            const size_t stride = (size_t)grid * BCP_POINT_THREADS, rem = m % stride;
            uint32_t per_last = (uint32_t)((rem + grid - 1) / grid);
            per_last = (per_last + 31u) & ~31u; /* whole warps; at most BCP_POINT_THREADS because rem < stride */
            uint32_t *d_counter = d_tab + point_slabs(cap) * BCP_QTAB_WORDS; /* in the 256 spare bytes behind the slabs */
            BCP_CUDA(cudaMemsetAsync(d_counter, 0, sizeof(uint32_t), st));
            k_ecdsa_point_rounds<<<grid, BCP_POINT_THREADS, 0, st>>>(d_pub + 65 * off, d_sig + 64 * off, (uint32_t)m, d_rec, d_tab,
                                                                     c->gcomb, d_ok + off, per_last, d_counter);
            continue;
        }
        const unsigned tpb = forced_tpb ? (unsigned)forced_tpb : (m <= BCP_WAVE ? BCP_POINT_SMALL_TPB : BCP_POINT_THREADS);
        unsigned blocks = (unsigned)((m + tpb - 1) / tpb);
        static const unsigned redo_every = getenv("BCP_POINT_FORCE_REDO") ? (unsigned)atoi(getenv("BCP_POINT_FORCE_REDO")) : 0u; /* tests */
        k_ecdsa_point<<<blocks, tpb, 0, st>>>(d_pub + 65 * off, d_sig + 64 * off, m, d_rec, d_tab, c->gcomb,
                                                           d_ok + off, redo_every);
    }
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}

extern "C" {

size_t bcp_gpu_ecdsa_scratch_bytes(size_t n) {
    size_t cap = n < BCP_SLICE ? n : BCP_SLICE;
    cap = (cap + BCP_POINT_THREADS - 1) / BCP_POINT_THREADS * BCP_POINT_THREADS; /* whole blocks */
    return (cap * BCP_REC_WORDS + point_slabs(cap) * BCP_QTAB_WORDS) * 4 + 256;
}

int bcp_gpu_ecdsa_reserve(int device, size_t n) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    return bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(n));
}

int bcp_gpu_ecdsa_verify_batch_dev(int device, const uint8_t *d_msg32, const uint8_t *d_pub65, const uint8_t *d_sig64,
                                   size_t n, uint32_t flags, uint8_t *d_ok, void *stream) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (n && (!d_msg32 || !d_pub65 || !d_sig64 || !d_ok)) {
        bcp_set_error("bcp_gpu_ecdsa_verify_batch_dev: NULL buffer");
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    int rc = bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(n));
    if (rc != BCP_OK) return rc;
    return bcp_ecdsa_launch(c, d_msg32, d_pub65, d_sig64, n, flags, d_ok, reinterpret_cast<uint32_t *>(c->scratch),
                         stream ? (cudaStream_t)stream : c->stream);
}

// secp256k1_ec_pubkey_parse (src/secp256k1/src/secp256k1.c:150-163) for n keys in the 65-byte-slot layout.
int bcp_gpu_pubkey_parse_batch(int device, const uint8_t *pub65, size_t n, uint8_t *xy64, uint8_t *ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (n == 0) return BCP_OK;
    if (!pub65 || !xy64 || !ok) {
        bcp_set_error("bcp_gpu_pubkey_parse_batch: NULL buffer");
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    int rc = bcp_ctx_reserve_stage(c, n * (65 + 64 + 1) + 512);
    if (rc != BCP_OK) return rc;
    uint8_t *d_pub = c->stage, *d_xy = d_pub + ((n * 65 + 255) / 256) * 256, *d_ok = d_xy + n * 64;
    BCP_CUDA(cudaMemcpyAsync(d_pub, pub65, n * 65, cudaMemcpyHostToDevice, c->stream));
    k_pubkey_parse<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(d_pub, n, d_xy, d_ok);
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaMemcpyAsync(xy64, d_xy, n * 64, cudaMemcpyDeviceToHost, c->stream));
    BCP_CUDA(cudaMemcpyAsync(ok, d_ok, n, cudaMemcpyDeviceToHost, c->stream));
    BCP_CUDA(cudaStreamSynchronize(c->stream));
    return BCP_OK;
}

static int lincomb_host(bcp_ctx *c, int mode, const uint8_t *pub65, const uint8_t *in_a, size_t a_bytes, const uint8_t *msg32,
                        size_t n, uint8_t *xy64, uint8_t *ok) {
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t o_pub = 0, o_a = up(n * 65), o_msg = o_a + up(n * a_bytes), o_xy = o_msg + up(n * 32), o_ok = o_xy + up(n * 64);
    int rc = bcp_ctx_reserve_stage(c, o_ok + up(n));
    if (rc != BCP_OK) return rc;
    const size_t blocks = (n + 127) / 128;
    rc = bcp_ctx_reserve_scratch(c, blocks * 128 * BCP_QTAB_WORDS * 4);
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    cudaStream_t st = c->stream;
    BCP_CUDA(cudaMemcpyAsync(d + o_pub, pub65, n * 65, cudaMemcpyHostToDevice, st));
    BCP_CUDA(cudaMemcpyAsync(d + o_a, in_a, n * a_bytes, cudaMemcpyHostToDevice, st));
    if (msg32) BCP_CUDA(cudaMemcpyAsync(d + o_msg, msg32, n * 32, cudaMemcpyHostToDevice, st));
    k_ec_lincomb<<<(unsigned)blocks, 128, 0, st>>>(d + o_pub, d + o_a, d + o_msg, mode, n, reinterpret_cast<uint32_t *>(c->scratch),
                                                   c->gcomb, d + o_xy, d + o_ok);
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaMemcpyAsync(xy64, d + o_xy, n * 64, cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaMemcpyAsync(ok, d + o_ok, n, cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaStreamSynchronize(st));
    return BCP_OK;
}

// secp256k1_ecdsa_recoverable_signature_parse_compact + secp256k1_ecdsa_recover for n items
// (src/secp256k1/src/modules/recovery/main_impl.h:38-58, 87-121, 170-190).
int bcp_gpu_ecdsa_recover_batch(int device, const uint8_t *msg32, const uint8_t *sig64, const uint8_t *recid, size_t n,
                                uint8_t *xy64, uint8_t *ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (n == 0) return BCP_OK;
    if (!msg32 || !sig64 || !recid || !xy64 || !ok) {
        bcp_set_error("bcp_gpu_ecdsa_recover_batch: NULL buffer");
        return BCP_ERR_INVALID;
    }
    // byte-level part on the host: the x coordinate the key is lifted from is r, or r + n when bit 1 of the recovery id is
    // set and r < p - n (:104-110); bit 0 is the parity of y.  A slot whose first byte is 0 never parses: that item fails.
    static const uint8_t N_BE[32] = {0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFE,
                                     0xBA, 0xAE, 0xDC, 0xE6, 0xAF, 0x48, 0xA0, 0x3B, 0xBF, 0xD2, 0x5E, 0x8C, 0xD0, 0x36, 0x41, 0x41};
    static const uint8_t PMN_BE[32] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0x01, 0x45, 0x51, 0x23, 0x19, 0x50, 0xB7, 0x5F, 0xC4,
                                       0x40, 0x2D, 0xA1, 0x72, 0x2F, 0xC9, 0xBA, 0xEE}; /* p - n (ecdsa_impl.h:45-47) */
    std::vector<uint8_t> pub(n * 65, 0);
    for (size_t i = 0; i < n; i++) {
        const uint8_t *r = sig64 + 64 * i;
        uint8_t *slot = &pub[65 * i];
        if (recid[i] > 3) continue;
        memcpy(slot + 1, r, 32);
        if (recid[i] & 2) {
            if (memcmp(r, PMN_BE, 32) >= 0) continue;
            int carry = 0;
            for (int k = 31; k >= 0; k--) {
                const int v = (int)r[k] + (int)N_BE[k] + carry;
                slot[1 + k] = (uint8_t)v;
                carry = v >> 8;
            }
        }
        slot[0] = (recid[i] & 1) ? 0x03 : 0x02;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    return lincomb_host(c, 1, pub.data(), sig64, 64, msg32, n, xy64, ok);
}

// secp256k1_ec_pubkey_tweak_add for n keys in the 65-byte-slot layout (src/secp256k1/src/secp256k1.c:477-499).
int bcp_gpu_pubkey_tweak_add_batch(int device, const uint8_t *pub65, const uint8_t *tweak32, size_t n, uint8_t *xy64,
                                   uint8_t *ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (n == 0) return BCP_OK;
    if (!pub65 || !tweak32 || !xy64 || !ok) {
        bcp_set_error("bcp_gpu_pubkey_tweak_add_batch: NULL buffer");
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    return lincomb_host(c, 0, pub65, tweak32, 32, nullptr, n, xy64, ok);
}

// Host-pointer form: the batch is cut into chunks; chunk k+1's host->device copy overlaps chunk k's
// kernels (two streams, events), verdicts come back per chunk.
int bcp_gpu_ecdsa_verify_batch(int device, const uint8_t *msg32, const uint8_t *pub65, const uint8_t *sig64, size_t n,
                               uint32_t flags, uint8_t *ok) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (n && (!msg32 || !pub65 || !sig64 || !ok)) {
        bcp_set_error("bcp_gpu_ecdsa_verify_batch: NULL buffer");
        return BCP_ERR_INVALID;
    }
    if (n == 0) return BCP_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    // Chunk schedule: a first chunk of one wave (its copy is the only one nothing hides, so it is kept short), then
    // chunks of six waves (measured on B200: 2 waves 39.4, 3: 40.9, 4: 41.8, 6: 42.7, 8: 42.5 M sigverifies/s end to end); a remainder of less than a wave rides with the chunk before it.
    static const size_t forced_waves = getenv("BCP_E2E_WAVES") ? (size_t)atoi(getenv("BCP_E2E_WAVES")) : 0;
    const size_t CH = forced_waves ? forced_waves * BCP_WAVE : BCP_E2E_CHUNK, W1 = BCP_WAVE;
    const size_t per = 32 + 65 + 64 + 1;
    const size_t cap = n < CH + W1 ? n : CH + W1; // largest chunk the schedule can produce
    const size_t slot = ((cap * per + 255) / 256) * 256;
    int rc = bcp_ctx_reserve_stage(c, 2 * slot);
    if (rc != BCP_OK) return rc;
    rc = bcp_ctx_reserve_scratch(c, 2 * bcp_gpu_ecdsa_scratch_bytes(cap));
    if (rc != BCP_OK) return rc;
    cudaStream_t st[2] = {c->stream, c->copy_stream};
    int k = 0;
    for (size_t off = 0; off < n; k ^= 1) {
        size_t m = off == 0 && n > CH ? W1 : CH;
        if (m > n - off) m = n - off;
        if (n - off - m < W1) m = n - off; // swallow a short tail
        uint8_t *base = c->stage + (size_t)k * slot;
        uint8_t *d_msg = base, *d_pub = d_msg + m * 32, *d_sig = d_pub + m * 65, *d_ok = d_sig + m * 64;
        uint32_t *d_rec = reinterpret_cast<uint32_t *>(c->scratch + (size_t)k * bcp_gpu_ecdsa_scratch_bytes(cap));
        // each slot is reused every second chunk: work on a slot is ordered by its own stream
        BCP_CUDA(cudaMemcpyAsync(d_msg, msg32 + off * 32, m * 32, cudaMemcpyHostToDevice, st[k]));
        BCP_CUDA(cudaMemcpyAsync(d_pub, pub65 + off * 65, m * 65, cudaMemcpyHostToDevice, st[k]));
        BCP_CUDA(cudaMemcpyAsync(d_sig, sig64 + off * 64, m * 64, cudaMemcpyHostToDevice, st[k]));
        rc = bcp_ecdsa_launch(c, d_msg, d_pub, d_sig, m, flags, d_ok, d_rec, st[k]);
        if (rc != BCP_OK) return rc;
        BCP_CUDA(cudaMemcpyAsync(ok + off, d_ok, m, cudaMemcpyDeviceToHost, st[k]));
        off += m;
    }
    BCP_CUDA(cudaStreamSynchronize(st[0]));
    BCP_CUDA(cudaStreamSynchronize(st[1]));
    return BCP_OK;
}

} // extern "C"
