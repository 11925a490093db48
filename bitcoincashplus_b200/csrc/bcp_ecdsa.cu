// bcp_ecdsa.cu — batched secp256k1 ECDSA verification kernels for sm_100a and their C-ABI entry
// points (include/bcp_gpu.h).  The arithmetic lives in fe.cuh / scalar.cuh / ge.cuh / ecdsa.cuh.
//
// Launch shape
//   k_ecdsa_prep   one thread per BCP_PREP_K signatures (thread t takes signatures t, t+T, t+2T, ...
//                  so that the byte loads of a warp stay adjacent), 128 threads per block
//   k_ecdsa_point  one thread per signature, 128 threads per block, 64 KB dynamic shared memory for
//                  the per-thread {1..8}Q tables (8 entries x 64 B x 128 threads), three blocks per SM
//   k_gcomb_build  one thread per comb-table entry, once per device at bcp_gpu_init
#include "bcp_ctx.cuh"
#include "ecdsa.cuh"

#define BCP_PREP_K 8
#define BCP_POINT_THREADS 128
#define BCP_POINT_SMEM (BCP_QTAB_WORDS * BCP_POINT_THREADS * 4)

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

__global__ void __launch_bounds__(128) k_ecdsa_prep(const uint8_t *__restrict__ msg32, const uint8_t *__restrict__ sig64,
                                                    size_t n, uint32_t flags, uint32_t *__restrict__ rec) {
    const size_t T = (size_t)gridDim.x * blockDim.x;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    // Montgomery's trick over this thread's signatures: pre[j] = s_0 * ... * s_{j-1}
    sig_scalars g[BCP_PREP_K];
    scalar pre[BCP_PREP_K];
    scalar acc;
#pragma unroll
    for (int i = 0; i < 8; i++) acc.v[i] = (i == 0);
    int cnt = 0;
#pragma unroll 1
    for (int j = 0; j < BCP_PREP_K; j++) {
        size_t i = t + (size_t)j * T;
        if (i >= n) break;
        sig_load(g[j], msg32 + 32 * i, sig64 + 64 * i, flags);
        pre[j] = acc;
        sc_mul(acc, acc, g[j].s);
        cnt++;
    }
    scalar inv;
    sc_inverse(inv, acc);
#pragma unroll 1
    for (int j = cnt - 1; j >= 0; j--) {
        size_t i = t + (size_t)j * T;
        scalar sinv;
        sc_mul(sinv, inv, pre[j]);
        sc_mul(inv, inv, g[j].s);
        uint32_t r[BCP_REC_WORDS];
        sig_finish(r, g[j], sinv);
        uint4 *dst = reinterpret_cast<uint4 *>(rec + i * BCP_REC_WORDS);
#pragma unroll
        for (int q = 0; q < BCP_REC_WORDS / 4; q++) dst[q] = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
    }
}

__global__ void __launch_bounds__(BCP_POINT_THREADS, 3)
    k_ecdsa_point(const uint8_t *__restrict__ pub65, const uint8_t *__restrict__ sig64, size_t n,
                  const uint32_t *__restrict__ rec, const uint32_t *__restrict__ gcomb, uint8_t *__restrict__ ok) {
    extern __shared__ uint32_t smem[];
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t r[BCP_REC_WORDS];
    const uint4 *src = reinterpret_cast<const uint4 *>(rec + i * BCP_REC_WORDS);
#pragma unroll
    for (int q = 0; q < BCP_REC_WORDS / 4; q++) {
        uint4 v = src[q];
        r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
    }
    qtable tab;
    tab.base = smem + threadIdx.x;
    tab.stride = BCP_POINT_THREADS;
    bool v = ecdsa_point_phase(r, pub65 + 65 * i, sig64 + 64 * i, tab, gcomb);
    ok[i] = v ? 1 : 0;
}

int bcp_ecdsa_build_tables(bcp_ctx *c) {
    BCP_CUDA(cudaMalloc(&c->gcomb, (size_t)BCP_GCOMB_WORDS * 4));
    int total = BCP_GCOMB_WINDOWS * BCP_GCOMB_ENTRIES;
    k_gcomb_build<<<(total + 127) / 128, 128, 0, c->stream>>>(c->gcomb);
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaFuncSetAttribute(k_ecdsa_point, cudaFuncAttributeMaxDynamicSharedMemorySize, BCP_POINT_SMEM));
    BCP_CUDA(cudaStreamSynchronize(c->stream));
    return BCP_OK;
}

static int launch_verify(bcp_ctx *c, const uint8_t *d_msg, const uint8_t *d_pub, const uint8_t *d_sig, size_t n,
                         uint32_t flags, uint8_t *d_ok, uint32_t *d_rec, cudaStream_t st) {
    if (n == 0) return BCP_OK;
    size_t prep_threads = (n + BCP_PREP_K - 1) / BCP_PREP_K;
    unsigned prep_blocks = (unsigned)((prep_threads + 127) / 128);
    k_ecdsa_prep<<<prep_blocks, 128, 0, st>>>(d_msg, d_sig, n, flags, d_rec);
    unsigned blocks = (unsigned)((n + BCP_POINT_THREADS - 1) / BCP_POINT_THREADS);
    k_ecdsa_point<<<blocks, BCP_POINT_THREADS, BCP_POINT_SMEM, st>>>(d_pub, d_sig, n, d_rec, c->gcomb, d_ok);
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}

extern "C" {

size_t bcp_gpu_ecdsa_scratch_bytes(size_t n) { return n * BCP_REC_WORDS * 4; }

int bcp_gpu_ecdsa_reserve(int device, size_t n) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(device));
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
    BCP_CUDA(cudaSetDevice(device));
    int rc = bcp_ctx_reserve_scratch(c, bcp_gpu_ecdsa_scratch_bytes(n));
    if (rc != BCP_OK) return rc;
    return launch_verify(c, d_msg32, d_pub65, d_sig64, n, flags, d_ok, reinterpret_cast<uint32_t *>(c->scratch),
                         stream ? (cudaStream_t)stream : c->stream);
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
    BCP_CUDA(cudaSetDevice(device));
    const size_t CH = 1u << 18; // signatures per chunk
    const size_t per = 32 + 65 + 64 + 1;
    const size_t chunk = n < CH ? n : CH;
    const size_t slot = ((chunk * per + 255) / 256) * 256;
    int rc = bcp_ctx_reserve_stage(c, 2 * slot);
    if (rc != BCP_OK) return rc;
    rc = bcp_ctx_reserve_scratch(c, 2 * bcp_gpu_ecdsa_scratch_bytes(chunk));
    if (rc != BCP_OK) return rc;
    cudaStream_t st[2] = {c->stream, c->copy_stream};
    int k = 0;
    for (size_t off = 0; off < n; off += chunk, k ^= 1) {
        size_t m = n - off < chunk ? n - off : chunk;
        uint8_t *base = c->stage + (size_t)k * slot;
        uint8_t *d_msg = base, *d_pub = d_msg + m * 32, *d_sig = d_pub + m * 65, *d_ok = d_sig + m * 64;
        uint32_t *d_rec = reinterpret_cast<uint32_t *>(c->scratch + (size_t)k * bcp_gpu_ecdsa_scratch_bytes(chunk));
        // each slot is reused every second chunk: work on a slot is ordered by its own stream
        BCP_CUDA(cudaMemcpyAsync(d_msg, msg32 + off * 32, m * 32, cudaMemcpyHostToDevice, st[k]));
        BCP_CUDA(cudaMemcpyAsync(d_pub, pub65 + off * 65, m * 65, cudaMemcpyHostToDevice, st[k]));
        BCP_CUDA(cudaMemcpyAsync(d_sig, sig64 + off * 64, m * 64, cudaMemcpyHostToDevice, st[k]));
        rc = launch_verify(c, d_msg, d_pub, d_sig, m, flags, d_ok, d_rec, st[k]);
        if (rc != BCP_OK) return rc;
        BCP_CUDA(cudaMemcpyAsync(ok + off, d_ok, m, cudaMemcpyDeviceToHost, st[k]));
    }
    BCP_CUDA(cudaStreamSynchronize(st[0]));
    BCP_CUDA(cudaStreamSynchronize(st[1]));
    return BCP_OK;
}

} // extern "C"
