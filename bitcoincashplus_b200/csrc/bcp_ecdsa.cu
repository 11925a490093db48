// bcp_ecdsa.cu — batched secp256k1 ECDSA verification kernels for sm_100a and their C-ABI entry
// points (include/bcp_gpu.h).  The arithmetic lives in fe.cuh / scalar.cuh / ge.cuh / ecdsa.cuh.
//
// Launch shape
//   k_ecdsa_prep   one thread per K signatures (K = 8, 16 or 32 by batch size) (thread t takes signatures t, t+T, t+2T, ...
//                  so that the byte loads of a warp stay adjacent), 128 threads per block
//   k_ecdsa_point  one thread per signature, 128 threads per block, four blocks per SM; the per-thread
//                  {1..8}Q tables (768 B: x, y, beta*x per entry) live in global scratch next to the phase-1 records (the first
//                  version kept them in shared memory, which capped occupancy at 12 warps per SM)
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
#define BCP_STEP_SYNC() do { if (BCP_STEP_SYNC_EVERY == 1 || (t % BCP_STEP_SYNC_EVERY) == 0) asm volatile("bar.sync 1, %0;" :: "r"(sync_n) : "memory"); } while (0)
#endif
#include "ecdsa.cuh"
#include <cstdlib>

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
    // K x (r, s, m) per thread.  K = 32: one 334-multiplication inversion per 32 signatures.
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
    sc_inverse(inv, acc);
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
                  uint8_t *__restrict__ ok) {
    const size_t gi = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = gi < n ? gi : n - 1; /* tail threads shadow the last signature (they must reach the barriers) */
    qtable_packed tab;
    tab.base = qtabs + gi * BCP_QTAB_WORDS; /* own slab even when shadowing: scratch is sized for whole blocks */
    bool v = ecdsa_point_phase(rec + i * BCP_REC_WORDS, pub65 + 65 * i, sig64 + 64 * i, tab, gcomb, (int)blockDim.x);
    if (gi < n) ok[gi] = v ? 1 : 0;
}

// Persistent form: the grid is exactly the resident set (BCP_POINT_BLOCKS_PER_SM blocks on every SM) and each block walks
// over the batch in rounds, so that
//   * a thread's {1..8}Q slab belongs to the RESIDENT SLOT (block, thread), not to the signature: the slabs of a whole
//     launch are 148 x 512 x 768 B = 58 MB, written and re-read round after round inside the 126 MB L2, instead of 768 B
//     of scratch per signature streaming through it (848 MB per million signatures, DRAM traffic 24x the algorithmic bytes);
//   * the last, partly filled round is dealt out EVENLY: every block takes per_last signatures (a multiple of 32) and
//     only the warps that hold some walk the steps, so what is left over runs on every SM at a few warps each (and
//     correspondingly faster) instead of as whole 256-thread blocks on a fifth of the chip.
__global__ void __launch_bounds__(BCP_POINT_THREADS, BCP_POINT_BLOCKS_PER_SM)
    k_ecdsa_point_rounds(const uint8_t *__restrict__ pub65, const uint8_t *__restrict__ sig64, uint32_t n,
                         const uint32_t *__restrict__ rec, uint32_t *__restrict__ qtabs, const uint32_t *__restrict__ gcomb,
                         uint8_t *__restrict__ ok, uint32_t per_last) {
    // Everything a round needs is recomputed from the kernel parameters (constant bank) and the round number, so that the
    // only values this outer loop keeps alive across the point phase are `r` and the barrier width — the phase itself uses
    // all 128 registers.
    const uint32_t full = n / (gridDim.x * blockDim.x);
    qtable_packed tab; /* slabs are packed per block: blockDim.x per block once there is a full round, else per_last */
    tab.base = qtabs + ((size_t)blockIdx.x * (full ? blockDim.x : per_last) + threadIdx.x) * BCP_QTAB_WORDS;
#pragma unroll 1
    for (uint32_t r = 0; r <= n / (gridDim.x * blockDim.x); r++) {
        uint32_t gi, i;
        int sync_n = (int)blockDim.x;
        {
            const uint32_t stride = gridDim.x * blockDim.x;
            if (r < n / stride) {
                gi = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
                i = gi;
            } else {
                const uint32_t base = r * stride + blockIdx.x * per_last;
                if (base >= n) break; /* uniform for the block */
                const uint32_t left = n - base < per_last ? n - base : per_last;
                const uint32_t warps = (left + 31u) >> 5;
                if ((threadIdx.x >> 5) >= warps) break; /* whole warps without a signature leave */
                sync_n = (int)(warps * 32u);
                gi = base + threadIdx.x;
                i = threadIdx.x < left ? gi : n - 1u; /* lanes past the end shadow the last signature (they must reach the barriers) */
            }
        }
        bool v = ecdsa_point_phase(rec + (size_t)i * BCP_REC_WORDS, pub65 + (size_t)65 * i, sig64 + (size_t)64 * i, tab, gcomb, sync_n);
        {
            const uint32_t stride = gridDim.x * blockDim.x;
            bool live = true;
            if (r < n / stride) {
                gi = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
            } else {
                const uint32_t base = r * stride + blockIdx.x * per_last;
                const uint32_t left = n - base < per_last ? n - base : per_last;
                gi = base + threadIdx.x;
                live = threadIdx.x < left;
            }
            if (live) ok[gi] = v ? 1 : 0;
        }
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
    }
    if (t < BCP_FE_ST_WIDE) {
        fe r1, r2;
        uint32_t w[16];
        for (int i = 0; i < 16; i++) w[i] = FE_ST_WIDE[t][i];
        fe_reduce_wide_t<true>(r1, w);
        fe_reduce_wide_t<false>(r2, w);
        for (int i = 0; i < 8; i++) bad |= r1.v[i] != r2.v[i];
    }
    if (bad) atomicAdd(fail, 1);
}
int bcp_fe_selftest(bcp_ctx *c) {
    int *d_fail = nullptr, h_fail = -1;
    BCP_CUDA(cudaMalloc(&d_fail, sizeof(int)));
    BCP_CUDA(cudaMemsetAsync(d_fail, 0, sizeof(int), c->stream));
    k_fe_selftest<<<(BCP_FE_ST_EDGES * BCP_FE_ST_EDGES + 127) / 128, 128, 0, c->stream>>>(d_fail);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h_fail, d_fail, sizeof(int), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_fail);
    if (e != cudaSuccess) return bcp_cuda_fail(e, "field self-test");
    if (h_fail != 0) {
        bcp_set_error("field arithmetic self-test failed on device %d: %d operand pairs disagree between the short-cut and the "
                      "full carry chains", c->device, h_fail);
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

// 1 (default): batches of more than a wave go through the persistent rounds kernel; 2: every batch; 0: never (A/B switch)
static int point_rounds_mode() {
    static const int mode = getenv("BCP_POINT_ROUNDS") ? atoi(getenv("BCP_POINT_ROUNDS")) : 1;
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
        // signatures per prep thread: one 334-multiplication inversion is shared by K signatures, but a thread's work is
        // serial, so small batches trade total work for parallelism (measured on B200 with tools/prep_sweep.py:
        // 41,650 signatures K=4 1.96 ms vs K=32 2.25 ms; 227k K=16 best; 1M K=32 best)
        static const int forced = getenv("BCP_PREP_K") ? atoi(getenv("BCP_PREP_K")) : 0;
        const int K = forced ? forced : (m >= 600000 ? 32 : m >= 150000 ? 16 : m >= 100000 ? 8 : 4);
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
            k_ecdsa_point_rounds<<<grid, BCP_POINT_THREADS, 0, st>>>(d_pub + 65 * off, d_sig + 64 * off, (uint32_t)m, d_rec, d_tab,
                                                                     c->gcomb, d_ok + off, per_last);
            continue;
        }
        const unsigned tpb = forced_tpb ? (unsigned)forced_tpb : (m <= BCP_WAVE ? BCP_POINT_SMALL_TPB : BCP_POINT_THREADS);
        unsigned blocks = (unsigned)((m + tpb - 1) / tpb);
        k_ecdsa_point<<<blocks, tpb, 0, st>>>(d_pub + 65 * off, d_sig + 64 * off, m, d_rec, d_tab, c->gcomb,
                                                           d_ok + off);
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
