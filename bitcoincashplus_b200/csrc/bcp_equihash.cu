// bcp_equihash.cu — Equihash solution verification (Blake2b index / XOR-collision check) for sm_100a.
//
// Reference path replaced, per header: CheckEquihashSolution (src/pow.cpp:298-323) ->
// Equihash<N,K>::IsValidSolution (src/crypto/equihash.cpp:724-770) with GetIndicesFromMinimal (:176-190),
// GenerateHash (:51-60), ExpandArray (:62-103), HasCollision (:279-288), IndicesBefore (equihash.h:101),
// DistinctIndices (equihash.tcc:10-21), IsZero (:256-265).  The verdict is the AND of
//   (a) the compact-size prefix says exactly SolutionWidth bytes;
//   (b) for every tree level l = 1..K and every aligned group of 2^l leaves, the XOR of the leaves'
//       l-th CollisionBitLength-bit chunk is zero (HasCollision after the earlier chunks were dropped),
//       and after level K the one remaining chunk XORs to zero over all leaves (IsZero);
//   (c) at every node the left subtree's first index is smaller than the right subtree's
//       (IndicesBefore is a lexicographic memcmp; with (d) the first element decides);
//   (d) all 2^K indices are pairwise distinct (DistinctIndices at a node compares left x right; every
//       pair of leaves meets at exactly one node).
// Only the bool is observable by callers, so the order in which the reference tests (a)-(d) is immaterial.
//
// k_eh_midstate  one thread per header: Blake2b state after the first 128 of the 140 input bytes
//                (personalization "ZcashPoW" || LE32(N) || LE32(K), digest length (512/N)*N/8)
// k_eh_verify    one thread per index (leaf): final Blake2b block (12 header bytes + LE32(index / ipo)),
//                slice, then a shared-memory XOR tree.  512 / 2^K headers per 512-thread block.
#include "bcp_ctx.cuh"
#include "blake2b.cuh"
#include "pow.cuh"
#include <algorithm>
#include <cstring>

template <unsigned N, unsigned K>
struct eh_params {
    static constexpr unsigned cbits = N / (K + 1);
    static constexpr unsigned nidx = 1u << K;
    static constexpr unsigned ipo = 512 / N;
    static constexpr unsigned hout = ipo * N / 8;
    static constexpr unsigned soln = nidx * (cbits + 1) / 8;
    static constexpr unsigned row_words = (N + 31) / 32;
    static constexpr unsigned hdr_len = 140 + (soln < 253 ? 1 : 3) + soln;
};

__device__ __forceinline__ uint64_t ld_le64(const uint8_t *p) {
    uint64_t v = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) v |= (uint64_t)p[i] << (8 * i);
    return v;
}

template <unsigned N, unsigned K>
__global__ void __launch_bounds__(128) k_eh_midstate(const uint8_t *__restrict__ hdrs, size_t stride, size_t count,
                                                     uint64_t *__restrict__ mid) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint8_t *h = hdrs + i * stride;
    uint64_t st[8], m[16];
    // "ZcashPoW" little-endian = 0x576f50687361635a... built bytewise to stay endian-explicit
    const uint64_t pers_lo = 0x576f50687361635aULL; /* 'Z','c','a','s','h','P','o','W' */
    const uint64_t pers_hi = (uint64_t)N | ((uint64_t)K << 32);
    blake2b_init_personal(st, eh_params<N, K>::hout, pers_lo, pers_hi);
#pragma unroll
    for (int j = 0; j < 16; j++) m[j] = ld_le64(h + 8 * j);
    blake2b_compress(st, m, 128, false);
#pragma unroll
    for (int j = 0; j < 8; j++) mid[i * 8 + j] = st[j];
}

// bits [pos, pos+len) of a big-endian bit string held in bytes
__device__ __forceinline__ uint32_t be_bits(const uint8_t *p, unsigned pos, unsigned len) {
    unsigned byte = pos >> 3, sh = pos & 7;
    uint64_t acc = 0;
#pragma unroll
    for (int i = 0; i < 5; i++) acc = (acc << 8) | p[byte + i];
    return (uint32_t)((acc >> (40 - sh - len)) & ((1u << len) - 1));
}

template <unsigned N, unsigned K>
__global__ void __launch_bounds__(512) k_eh_verify(const uint8_t *__restrict__ hdrs, size_t stride, size_t hdr_len,
                                                   size_t count, const uint64_t *__restrict__ mid,
                                                   uint8_t *__restrict__ ok) {
    typedef eh_params<N, K> P;
    constexpr unsigned HPB = 512 / P::nidx; /* headers per block */
    constexpr unsigned RW = P::row_words;
    __shared__ uint32_t rows[512][RW + 1];
    __shared__ uint32_t idxs[512];
    __shared__ uint8_t soln[HPB][P::soln + 8];
    __shared__ uint64_t smid[HPB][8];
    __shared__ uint32_t tail[HPB][4]; /* header bytes 128..139 */
    __shared__ int bad[HPB];
    __shared__ uint32_t seen[HPB][2 * P::nidx]; /* distinctness table, 2 slots per index */

    const unsigned hl = threadIdx.x / P::nidx;   /* header slot in this block */
    const unsigned leaf = threadIdx.x % P::nidx;
    const size_t hidx = (size_t)blockIdx.x * HPB + hl;
    const bool live = hidx < count;
    const uint8_t *h = hdrs + (live ? hidx : 0) * stride;

    // stage the solution bytes, the midstate and the header tail
    if (leaf == 0) {
        int b = 0;
        if (live) {
            // compact size must encode exactly SolutionWidth and the header must be exactly that long
            if (P::soln < 253) b = (h[140] != P::soln);
            else b = !(h[140] == 253 && h[141] == (P::soln & 0xff) && h[142] == (P::soln >> 8));
            if (hdr_len != P::hdr_len) b = 1;
        }
        bad[hl] = b;
    }
    seen[hl][2 * leaf] = 0;
    seen[hl][2 * leaf + 1] = 0;
    if (live) {
        const uint8_t *s = h + (P::hdr_len - P::soln);
        for (unsigned t = leaf; t < P::soln + 8; t += P::nidx) soln[hl][t] = t < P::soln ? s[t] : 0;
        if (leaf < 8) smid[hl][leaf] = mid[hidx * 8 + leaf];
        if (leaf < 3) {
            const uint8_t *q = h + 128 + 4 * leaf;
            tail[hl][leaf] = q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
        }
    }
    __syncthreads();

    // my index, my leaf hash
    uint32_t my = 0;
    if (live) {
        my = be_bits(soln[hl], leaf * (P::cbits + 1), P::cbits + 1);
        idxs[threadIdx.x] = my;
        uint64_t st[8], m[16];
#pragma unroll
        for (int j = 0; j < 8; j++) st[j] = smid[hl][j];
        const uint32_t g = my / P::ipo;
        m[0] = (uint64_t)tail[hl][0] | ((uint64_t)tail[hl][1] << 32);
        m[1] = (uint64_t)tail[hl][2] | ((uint64_t)g << 32);
#pragma unroll
        for (int j = 2; j < 16; j++) m[j] = 0;
        blake2b_compress(st, m, 144, true);
        // slice (my % ipo) * N/8 .. + N/8 of the little-endian digest bytes, repacked big-endian into words
        const unsigned off = (my % P::ipo) * (N / 8);
#pragma unroll
        for (unsigned w = 0; w < RW; w++) {
            uint32_t v = 0;
#pragma unroll
            for (unsigned b = 0; b < 4; b++) {
                unsigned pos = 4 * w + b; /* byte within the slice */
                uint32_t byte = 0;
                if (pos < N / 8) {
                    unsigned q = off + pos;
                    byte = (uint32_t)(st[q >> 3] >> (8 * (q & 7))) & 0xff;
                }
                v = (v << 8) | byte;
            }
            rows[threadIdx.x][w] = v;
        }
    } else {
        idxs[threadIdx.x] = 0;
#pragma unroll
        for (unsigned w = 0; w < RW; w++) rows[threadIdx.x][w] = 0;
    }
    __syncthreads();

    int fail = 0;
    // (d) pairwise distinct (DistinctIndices at every merge of the reference, src/crypto/equihash.cpp:742-748, amounts to
    // "no index twice"): every thread inserts its index into a small open-addressing table of its header in shared
    // memory; meeting its own value there means a duplicate.  2*nidx slots, linear probing, exact — and ~10 instructions
    // per thread where the all-pairs scan took 1.5 * nidx.
    if (live) {
        const uint32_t key = my + 1; /* 0 = empty slot */
        uint32_t slot = (my * 2654435761u) >> (32 - (K + 1));
        for (;;) {
            const uint32_t old = atomicCAS(&seen[hl][slot], 0u, key);
            if (old == 0u) break;
            if (old == key) { fail = 1; break; }
            slot = (slot + 1) & (2 * P::nidx - 1);
        }
    }
    // (b),(c) tree: after level l, slot (hl*nidx + j * 2^l) holds the XOR of group j
#pragma unroll 1
    for (unsigned l = 1; l <= K; l++) {
        const unsigned width = 1u << l, halfw = width >> 1;
        uint32_t x[RW];
        const bool owner = (leaf & (width - 1)) == 0;
        if (owner) {
#pragma unroll
            for (unsigned w = 0; w < RW; w++) x[w] = rows[threadIdx.x][w] ^ rows[threadIdx.x + halfw][w];
            if (live && !(idxs[threadIdx.x] < idxs[threadIdx.x + halfw])) fail = 1;
        }
        __syncthreads();
        if (owner) {
#pragma unroll
            for (unsigned w = 0; w < RW; w++) rows[threadIdx.x][w] = x[w];
            // chunk l-1 (bits [(l-1)*cbits, l*cbits)) of the XOR must vanish (read back from my own row:
            // a run-time word index into registers would force the row into local memory)
            const unsigned pos = (l - 1) * P::cbits;
            const unsigned wi = pos >> 5, sh = pos & 31;
            uint64_t two = ((uint64_t)rows[threadIdx.x][wi] << 32) | (wi + 1 < RW ? rows[threadIdx.x][wi + 1] : 0u);
            uint32_t chunk = (uint32_t)((two >> (64 - sh - P::cbits)) & ((1u << P::cbits) - 1));
            if (live && chunk != 0) fail = 1;
        }
        __syncthreads();
    }
    if (live && leaf == 0) {
        // IsZero over what is left: chunk K
        const unsigned pos = K * P::cbits;
        const unsigned wi = pos >> 5, sh = pos & 31;
        uint64_t two = ((uint64_t)rows[threadIdx.x][wi] << 32) | (wi + 1 < RW ? rows[threadIdx.x][wi + 1] : 0u);
        uint32_t chunk = (uint32_t)((two >> (64 - sh - P::cbits)) & ((1u << P::cbits) - 1));
        if (chunk != 0) fail = 1;
    }
    if (fail) atomicOr(&bad[hl], 1);
    __syncthreads();
    if (live && leaf == 0) ok[hidx] = bad[hl] ? 0 : 1;
}

template <unsigned N, unsigned K>
static int eh_launch(bcp_ctx *c, const uint8_t *d_hdrs, size_t stride, size_t hdr_len, size_t count, uint64_t *d_mid,
                     uint8_t *d_ok, cudaStream_t st) {
    constexpr unsigned HPB = 512 / eh_params<N, K>::nidx;
    k_eh_midstate<N, K><<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d_hdrs, stride, count, d_mid);
    k_eh_verify<N, K><<<(unsigned)((count + HPB - 1) / HPB), 512, 0, st>>>(d_hdrs, stride, hdr_len, count, d_mid, d_ok);
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}
static int eh_dispatch(bcp_ctx *c, unsigned n, unsigned k, const uint8_t *d_hdrs, size_t stride, size_t hdr_len,
                       size_t count, uint64_t *d_mid, uint8_t *d_ok, cudaStream_t st) {
    if (n == 200 && k == 9) return eh_launch<200, 9>(c, d_hdrs, stride, hdr_len, count, d_mid, d_ok, st);
    if (n == 48 && k == 5) return eh_launch<48, 5>(c, d_hdrs, stride, hdr_len, count, d_mid, d_ok, st);
    if (n == 96 && k == 5) return eh_launch<96, 5>(c, d_hdrs, stride, hdr_len, count, d_mid, d_ok, st);
    if (n == 96 && k == 3) return eh_launch<96, 3>(c, d_hdrs, stride, hdr_len, count, d_mid, d_ok, st);
    bcp_set_error("unsupported Equihash parameters (%u,%u): the reference supports (200,9) (48,5) (96,5) (96,3)", n, k);
    return BCP_ERR_INVALID;
}

int bcp_sha256d_strided_dev(bcp_ctx *c, const uint8_t *d_data, size_t stride, size_t len, size_t count, uint8_t *d_out,
                            cudaStream_t st); // bcp_hash.cu

// Per header: CheckProofOfWork(GetHash(), nBits, IsBCPEnabled(nHeight)) (src/validation.cpp:3214-3217, src/pow.cpp:141-161)
// and whether hashPrevBlock equals the hash of the header before it in the batch.
struct pow_limits {
    uint32_t post[8], legacy[8];
};
__global__ void __launch_bounds__(128) k_header_pow(const uint8_t *__restrict__ hdrs, size_t stride, size_t count,
                                                    const uint8_t *__restrict__ hash, pow_limits lim, uint32_t fork_height,
                                                    uint8_t *__restrict__ ok_pow, uint8_t *__restrict__ linked) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint8_t *h = hdrs + i * stride;
    auto le32 = [](const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; };
    const uint32_t height = le32(h + 68), nbits = le32(h + 104);
    uint32_t hw[8];
    for (int j = 0; j < 8; j++) hw[j] = le32(hash + i * 32 + 4 * j);
    ok_pow[i] = pow_check(hw, nbits, height >= fork_height ? lim.post : lim.legacy) ? 1 : 0;
    if (linked) {
        bool same = true;
        if (i > 0)
            for (int j = 0; j < 32; j++) same = same && h[4 + j] == hash[(i - 1) * 32 + j];
        linked[i] = same ? 1 : 0;
    }
}

extern "C" {

BCP_API size_t bcp_gpu_equihash_scratch_bytes(size_t count) { return count * 64 + 256; }

// d_ok: count verdict bytes.  d_scratch: bcp_gpu_equihash_scratch_bytes(count).
BCP_API int bcp_gpu_equihash_verify_batch_dev(int device, unsigned n, unsigned k, const uint8_t *d_hdrs, size_t stride,
                                              size_t hdr_len, size_t count, uint8_t *d_ok, void *d_scratch,
                                              void *stream) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (count == 0) return BCP_OK;
    if (hdr_len > stride || hdr_len < 140) { bcp_set_error("equihash: bad header length/stride"); return BCP_ERR_INVALID; }
    BCP_CUDA(cudaSetDevice(c->device));
    return eh_dispatch(c, n, k, d_hdrs, stride, hdr_len, count, reinterpret_cast<uint64_t *>(d_scratch), d_ok,
                       stream ? (cudaStream_t)stream : c->stream);
}

// Host-pointer form: verdicts (CheckEquihashSolution) and, when hash32 != NULL, the SHA-256d block hash of
// each serialized header (CBlockHeader::GetHash, src/primitives/block.cpp:17-28) in the same call.
BCP_API int bcp_gpu_equihash_verify_batch(int device, unsigned n, unsigned k, const uint8_t *hdrs, size_t stride,
                                          size_t hdr_len, size_t count, uint8_t *ok, uint8_t *hash32) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (count == 0) return BCP_OK;
    if (!hdrs || !ok) { bcp_set_error("bcp_gpu_equihash_verify_batch: NULL buffer"); return BCP_ERR_INVALID; }
    if (hdr_len > stride || hdr_len < 140 || count * stride > 0xFFFFFFF0ull) {
        bcp_set_error("equihash: bad header length/stride/count");
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    size_t bytes = count * stride;
    size_t o_mid = (bytes + 8 + 255) / 256 * 256, o_ok = o_mid + count * 64, o_hash = (o_ok + count + 255) / 256 * 256;
    size_t total = o_hash + count * 32 + 256;
    int rc = bcp_ctx_reserve_stage(c, total);
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    // Chunks on alternating streams: with pinned host buffers chunk i+1's upload runs under chunk i's kernels (a (200,9)
    // header is 1,487 bytes, so 10,000 of them are 15 MB: as long on the bus as in the verify kernel).
    cudaStream_t sts[2] = {c->stream, c->copy_stream};
    const size_t chunk = count <= 2048 ? count : std::max<size_t>(1024, (count + 3) / 4);
    int q = 0;
    for (size_t off = 0; off < count; off += chunk, q ^= 1) {
        const size_t m = std::min(chunk, count - off);
        cudaStream_t st = sts[q];
        BCP_CUDA(cudaMemcpyAsync(d + off * stride, hdrs + off * stride, m * stride, cudaMemcpyHostToDevice, st));
        rc = eh_dispatch(c, n, k, d + off * stride, stride, hdr_len, m, reinterpret_cast<uint64_t *>(d + o_mid + off * 64),
                         d + o_ok + off, st);
        if (rc != BCP_OK) return rc;
        BCP_CUDA(cudaMemcpyAsync(ok + off, d + o_ok + off, m, cudaMemcpyDeviceToHost, st));
        if (hash32) {
            rc = bcp_sha256d_strided_dev(c, d + off * stride, stride, hdr_len, m, d + o_hash + off * 32, st);
            if (rc != BCP_OK) return rc;
            BCP_CUDA(cudaMemcpyAsync(hash32 + off * 32, d + o_hash + off * 32, m * 32, cudaMemcpyDeviceToHost, st));
        }
    }
    BCP_CUDA(cudaStreamSynchronize(sts[0]));
    BCP_CUDA(cudaStreamSynchronize(sts[1]));
    return BCP_OK;
}

// Headers-first pipeline: everything CheckBlockHeader needs (src/validation.cpp:3202-3220) plus the
// in-batch linkage, for `count` serialized post-fork headers in one pass over one upload.
BCP_API int bcp_gpu_header_check_batch(int device, unsigned n, unsigned k, const uint8_t *hdrs, size_t stride, size_t hdr_len,
                                       size_t count, const uint8_t pow_limit32[32], const uint8_t pow_limit_legacy32[32],
                                       uint32_t fork_height, uint8_t *ok_solution, uint8_t *ok_pow, uint8_t *linked,
                                       uint8_t *hash32) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (count == 0) return BCP_OK;
    if (!hdrs || !ok_solution || !ok_pow || !hash32 || !pow_limit32 || !pow_limit_legacy32) {
        bcp_set_error("bcp_gpu_header_check_batch: NULL buffer");
        return BCP_ERR_INVALID;
    }
    if (hdr_len > stride || hdr_len < 140 || count * stride > 0xFFFFFFF0ull) {
        bcp_set_error("header check: bad header length/stride/count");
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    const size_t bytes = count * stride;
    const size_t o_mid = (bytes + 8 + 255) / 256 * 256, o_ok = o_mid + count * 64, o_pow = o_ok + count, o_link = o_pow + count,
                 o_hash = (o_link + count + 255) / 256 * 256, total = o_hash + count * 32 + 256;
    int rc = bcp_ctx_reserve_stage(c, total);
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    cudaStream_t st = c->stream;
    BCP_CUDA(cudaMemcpyAsync(d, hdrs, bytes, cudaMemcpyHostToDevice, st));
    rc = eh_dispatch(c, n, k, d, stride, hdr_len, count, reinterpret_cast<uint64_t *>(d + o_mid), d + o_ok, st);
    if (rc != BCP_OK) return rc;
    rc = bcp_sha256d_strided_dev(c, d, stride, hdr_len, count, d + o_hash, st);
    if (rc != BCP_OK) return rc;
    pow_limits lim;
    for (int j = 0; j < 8; j++) {
        memcpy(&lim.post[j], pow_limit32 + 4 * j, 4); // uint256 memory order is little-endian limbs on this host
        memcpy(&lim.legacy[j], pow_limit_legacy32 + 4 * j, 4);
    }
    k_header_pow<<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d, stride, count, d + o_hash, lim, fork_height, d + o_pow,
                                                                 d + o_link);
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaMemcpyAsync(ok_solution, d + o_ok, count, cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaMemcpyAsync(ok_pow, d + o_pow, count, cudaMemcpyDeviceToHost, st));
    if (linked) BCP_CUDA(cudaMemcpyAsync(linked, d + o_link, count, cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaMemcpyAsync(hash32, d + o_hash, count * 32, cudaMemcpyDeviceToHost, st));
    BCP_CUDA(cudaStreamSynchronize(st));
    return BCP_OK;
}

} // extern "C"
