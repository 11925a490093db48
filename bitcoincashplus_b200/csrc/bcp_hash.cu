// bcp_hash.cu — SHA-256d kernels: transaction ids, generic message batches (block headers) and the
// Merkle root with the CVE-2012-2459 `mutated` flag.  C-ABI in include/bcp_gpu.h.
//
// Reference functions replaced (per block): CTransaction::ComputeHash (src/primitives/transaction.cpp:
// 73-75) over every transaction, ComputeMerkleRoot / BlockMerkleRoot (src/consensus/merkle.cpp:47-150,
// 175-182), CBlockHeader::GetHash (src/primitives/block.cpp:17-28).
//
// k_sha256d_msgs   one thread per message (tx or header).  Full 64-byte blocks are read as aligned
//                  32-bit words and re-aligned + byte-swapped with one PRMT per word; only the tail goes
//                  through byte loads.  Messages of one warp are adjacent in memory, so the lines a warp
//                  touches are shared by its lanes.
// k_merkle_levels  256 nodes per 128-thread block, staged in shared memory, LEVELS tree levels per launch.  Three levels
//                  (128, 64, 32 pairs: four, two, one FULL warp) is the default: with eight levels per launch the five
//                  upper ones ran on 16, 8, 4, 2, 1 lanes of one warp, which costs a warp's full issue time each, and the
//                  kernel sat at 0.31 of the issue rate (0.50 ms per million leaves); the price is more launches (seven
//                  instead of three for a million leaves).  The eight-level form remains for the entry point whose scratch
//                  is a fixed static array (it needs 1/32 of the scratch).
#include "bcp_ctx.cuh"
#include "sha256.cuh"
#include <cstring>

#define MERKLE_THREADS 128
#define MERKLE_NODES 256
#define MERKLE_LEVELS_WIDE 3  /* default: every warp that works is full */
#define MERKLE_LEVELS_DEEP 8  /* least scratch: one node out per 256 in */

// off != nullptr: message i is data[off[i] .. off[i+1]);  off == nullptr: message i is data[i*stride .. +fixed_len)
__global__ void __launch_bounds__(128) k_sha256d_msgs(const uint8_t *__restrict__ data, const uint32_t *__restrict__ off,
                                                      uint32_t stride, uint32_t fixed_len, size_t count,
                                                      uint8_t *__restrict__ out32) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t start = off ? off[i] : (uint32_t)(i * stride);
    const uint32_t len = off ? off[i + 1] - start : fixed_len;
    const uint32_t *aligned = reinterpret_cast<const uint32_t *>(data);
    uint32_t s[8];
    sha256_init(s);
    uint32_t pos = 0;
    const uint32_t sh = start & 3u;
    const uint32_t sel = (sh << 12) | ((sh + 1) << 8) | ((sh + 2) << 4) | (sh + 3);
#pragma unroll 1
    for (; pos + 64 <= len; pos += 64) {
        const uint32_t *a = aligned + ((start + pos) >> 2);
        uint32_t x[17], w[16];
#pragma unroll
        for (int j = 0; j < 17; j++) x[j] = __ldg(a + j); /* may read 3 bytes past the message: buffers are padded */
#pragma unroll
        for (int j = 0; j < 16; j++) w[j] = __byte_perm(x[j], x[j + 1], sel);
        sha256_compress(s, w);
    }
    const uint32_t total = ((len + 9 + 63) / 64) * 64;
#pragma unroll 1
    for (; pos < total; pos += 64) {
        uint32_t w[16];
#pragma unroll
        for (int j = 0; j < 16; j++) w[j] = sha256_msg_word(data + start, len, total, pos + 4 * j);
        sha256_compress(s, w);
    }
    sha256_second_pass(s);
    uint4 *o = reinterpret_cast<uint4 *>(out32 + 32 * i);
    o[0] = make_uint4(__byte_perm(s[0], 0, 0x0123), __byte_perm(s[1], 0, 0x0123), __byte_perm(s[2], 0, 0x0123),
                      __byte_perm(s[3], 0, 0x0123));
    o[1] = make_uint4(__byte_perm(s[4], 0, 0x0123), __byte_perm(s[5], 0, 0x0123), __byte_perm(s[6], 0, 0x0123),
                      __byte_perm(s[7], 0, 0x0123));
}

// in: `count` nodes of tree level `level0` (each spans 2^level0 leaves) of a tree with `nleaves` leaves.
// out: per block the 256 >> LEVELS nodes of level level0 + LEVELS it owns (fewer at the tree's edge, or the root earlier).  A pair (L, R) at level l is
// hashed as SHA256d(L || R); a missing R (odd tail) is replaced by L.  `mutated` is raised when L == R
// for a pair whose right node lies wholly inside the real leaves (merkle.cpp:86; the final sweep :109-140
// never tests).
template <int LEVELS>
__global__ void __launch_bounds__(MERKLE_THREADS)
    k_merkle_levels(const uint32_t *__restrict__ in, size_t count, int level0, size_t nleaves, uint32_t *__restrict__ out,
                    int *__restrict__ mutated) {
    __shared__ uint32_t node[MERKLE_NODES][8 + 1]; /* +1: pad against bank conflicts on the row stride */
    const size_t base = (size_t)blockIdx.x * MERKLE_NODES;
    for (int t = threadIdx.x; t < MERKLE_NODES * 8; t += MERKLE_THREADS) {
        size_t g = base + (t >> 3);
        node[t >> 3][t & 7] = g < count ? __byte_perm(in[g * 8 + (t & 7)], 0, 0x0123) : 0u;
    }
    __syncthreads();
    size_t cnt = count;          /* nodes at the current level, whole tree */
    size_t first = base;         /* global index of this block's node 0 at the current level */
    int mut = 0;
#pragma unroll 1
    for (int l = 0; l < LEVELS && cnt > 1; l++) {
        const int pairs = MERKLE_NODES >> (l + 1);
        uint32_t s[8];
        bool active = false;
        if ((int)threadIdx.x < pairs) {
            size_t gl = first + 2 * threadIdx.x, gr = gl + 1;
            if (gl < cnt) {
                active = true;
                uint32_t w[16];
                const bool has_r = gr < cnt;
                bool same = true;
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    w[k] = node[2 * threadIdx.x][k];
                    w[8 + k] = has_r ? node[2 * threadIdx.x + 1][k] : w[k];
                    same = same && (w[k] == w[8 + k]);
                }
                const size_t span = (size_t)1 << (level0 + l);
                if (has_r && same && (gr + 1) * span <= nleaves) mut = 1;
                sha256d_64(s, w);
            }
        }
        __syncthreads();
        if (active) {
#pragma unroll
            for (int k = 0; k < 8; k++) node[threadIdx.x][k] = s[k];
        }
        __syncthreads();
        cnt = (cnt + 1) >> 1;
        first >>= 1;
    }
    if (mut) atomicOr(mutated, 1);
    // the nodes this block is left with: positions first .. of the level reached (cnt nodes in all)
    constexpr int LEFT = MERKLE_NODES >> LEVELS;
    for (int t = threadIdx.x; t < LEFT * 8; t += MERKLE_THREADS) {
        const size_t g = first + (size_t)(t >> 3);
        if (g < cnt) out[g * 8 + (t & 7)] = __byte_perm(node[t >> 3][t & 7], 0, 0x0123);
    }
}

// scratch words for a tree of n leaves: two halves, each as large as the first launch's output
static size_t merkle_tmp_words_for(size_t n, int levels) {
    const size_t per = MERKLE_NODES >> levels; /* nodes a full block leaves behind */
    return 2 * (((n + MERKLE_NODES - 1) / MERKLE_NODES) * per * 8) + 16;
}
// root of `n` leaves already on the device; nodes ping-pong between two scratch halves
static int merkle_dev(bcp_ctx *c, const uint8_t *d_leaves, size_t n, uint8_t *d_root32, int *d_mut, uint32_t *d_tmp,
                      cudaStream_t st, int levels = MERKLE_LEVELS_WIDE) {
    BCP_CUDA(cudaMemsetAsync(d_mut, 0, sizeof(int), st));
    if (n == 0) {
        BCP_CUDA(cudaMemsetAsync(d_root32, 0, 32, st));
        return BCP_OK;
    }
    const uint32_t *cur = reinterpret_cast<const uint32_t *>(d_leaves);
    size_t cnt = n;
    int level = 0;
    const size_t half = (merkle_tmp_words_for(n, levels) - 16) / 2;
    uint32_t *buf[2] = {d_tmp, d_tmp + half};
    int k = 0;
    while (cnt > 1) {
        unsigned blocks = (unsigned)((cnt + MERKLE_NODES - 1) / MERKLE_NODES);
        if (levels == MERKLE_LEVELS_WIDE) k_merkle_levels<MERKLE_LEVELS_WIDE><<<blocks, MERKLE_THREADS, 0, st>>>(cur, cnt, level, n, buf[k], d_mut);
        else k_merkle_levels<MERKLE_LEVELS_DEEP><<<blocks, MERKLE_THREADS, 0, st>>>(cur, cnt, level, n, buf[k], d_mut);
        cur = buf[k];
        k ^= 1;
        for (int l = 0; l < levels && cnt > 1; l++) cnt = (cnt + 1) >> 1;
        level += levels;
    }
    BCP_CUDA(cudaGetLastError());
    BCP_CUDA(cudaMemcpyAsync(d_root32, cur, 32, cudaMemcpyDeviceToDevice, st));
    return BCP_OK;
}
static size_t merkle_tmp_words(size_t n) { return merkle_tmp_words_for(n, MERKLE_LEVELS_WIDE); }

int bcp_sha256d_offsets_dev(bcp_ctx *c, const uint8_t *d_data, const uint32_t *d_off, size_t count, uint8_t *d_out,
                            cudaStream_t st) {
    if (count == 0) return BCP_OK;
    k_sha256d_msgs<<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d_data, d_off, 0, 0, count, d_out);
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}
int bcp_merkle_dev(bcp_ctx *c, const uint8_t *d_leaves, size_t n, uint8_t *d_root32, int *d_mut, uint32_t *d_tmp,
                   cudaStream_t st) {
    return merkle_dev(c, d_leaves, n, d_root32, d_mut, d_tmp, st);
}
// fewest launches (eight levels each): for a caller that has the result off its critical path and pays per API call
int bcp_merkle_dev_few_launches(bcp_ctx *c, const uint8_t *d_leaves, size_t n, uint8_t *d_root32, int *d_mut, uint32_t *d_tmp,
                                cudaStream_t st) {
    return merkle_dev(c, d_leaves, n, d_root32, d_mut, d_tmp, st, MERKLE_LEVELS_DEEP);
}
size_t bcp_merkle_tmp_words(size_t n) { return merkle_tmp_words(n); }

// fixed-size messages at a fixed stride (serialized block headers)
int bcp_sha256d_strided_dev(bcp_ctx *c, const uint8_t *d_data, size_t stride, size_t len, size_t count, uint8_t *d_out,
                            cudaStream_t st) {
    if (count == 0) return BCP_OK;
    k_sha256d_msgs<<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d_data, nullptr, (uint32_t)stride, (uint32_t)len, count,
                                                                   d_out);
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}

extern "C" {

// Compact-size + transaction framing scan (src/serialize.h:260-300, src/primitives/transaction.h:223-243,
// header layout src/primitives/block.h:49-73).  Sequential byte parsing: stays on the host by design.
BCP_API long bcp_block_index(const uint8_t *blk, size_t len, int legacy_header, uint32_t *tx_off, size_t max_tx) {
    // one scanner for the whole library: the framing scan of bcp_block.cu without its input and output lists
    long ntx = 0;
    const long rc = bcp_block_index_full(blk, len, legacy_header, tx_off, nullptr, max_tx, nullptr, nullptr, nullptr, 0, &ntx,
                                         nullptr, nullptr, nullptr, 0, nullptr, nullptr);
    return rc < 0 ? rc : ntx;
}

BCP_API int bcp_gpu_sha256d_batch_dev(int device, const uint8_t *d_data, const uint32_t *d_off, size_t count,
                                      uint8_t *d_out32, void *stream) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (count == 0) return BCP_OK;
    BCP_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
    k_sha256d_msgs<<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d_data, d_off, 0, 0, count, d_out32);
    BCP_CUDA(cudaGetLastError());
    return BCP_OK;
}

BCP_API size_t bcp_gpu_merkle_scratch_bytes(size_t n) { return merkle_tmp_words(n) * 4 + 64; }

// d_scratch: bcp_gpu_merkle_scratch_bytes(n) bytes; d_root32: 32 bytes; d_mutated: one int
BCP_API int bcp_gpu_merkle_root_dev(int device, const uint8_t *d_leaves32, size_t n, uint8_t *d_root32, int *d_mutated,
                                    void *d_scratch, void *stream) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    BCP_CUDA(cudaSetDevice(c->device));
    return merkle_dev(c, d_leaves32, n, d_root32, d_mutated, reinterpret_cast<uint32_t *>(d_scratch),
                      stream ? (cudaStream_t)stream : c->stream);
}

// ---- allocation-free Merkle root ---------------------------------------------------------------------
// Same tree walk as bcp_gpu_merkle_root_dev, for callers that must not allocate device memory at all:
// the thread that runs ConnectBlock holds the node's main lock, and cudaMalloc / cudaFree synchronise
// the whole device, so a validation thread that allocates stalls every other stream on that GPU; an
// embedder that hands the library a fixed VRAM budget has nothing left to allocate from either.  The
// inner-node ping-pong buffers therefore live in this file-scope device array instead of a
// caller-supplied or context-owned buffer.  Every GPU that loads the library carries the array
// (BCP_MERKLE_STATIC_BYTES, about 256 KiB) until the process exits.
#define BCP_MERKLE_STATIC_WORDS (2 * ((BCP_MERKLE_STATIC_MAX_LEAVES / MERKLE_NODES) * 8) + 16) /* eight levels per launch */
static __device__ __align__(256) uint32_t g_merkle_static_scratch[BCP_MERKLE_STATIC_WORDS];
// In host code the name above is only a handle for the symbol, not an address a kernel can use; the
// device address is looked up once per device (with that device current) and cached here.  The cache is
// per GPU ordinal — every lane of a GPU shares the one array — and has a lock of its own, because two
// lanes of the same GPU hold different context locks.
static uint32_t *g_merkle_static_ptr[BCP_MAX_DEVICES];
static std::mutex g_merkle_static_mu;

BCP_API int bcp_gpu_merkle_root_dev_static(int device, const uint8_t *d_leaves32, size_t n, uint8_t *d_root32,
                                           int *d_mutated, void *stream) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (!d_root32 || !d_mutated || (n && !d_leaves32)) {
        bcp_set_error("bcp_gpu_merkle_root_dev_static: NULL buffer");
        return BCP_ERR_INVALID;
    }
    if (n > BCP_MERKLE_STATIC_MAX_LEAVES) {
        bcp_set_error("bcp_gpu_merkle_root_dev_static: %zu leaves exceed the built-in scratch (max %u leaves); "
                      "use bcp_gpu_merkle_root_dev with caller-provided scratch",
                      n, (unsigned)BCP_MERKLE_STATIC_MAX_LEAVES);
        return BCP_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    uint32_t *scratch = nullptr;
    { /* `device` may carry a lane in its upper bits: the cache is indexed by the GPU ordinal alone */
        std::lock_guard<std::mutex> lk2(g_merkle_static_mu);
        uint32_t *&slot = g_merkle_static_ptr[c->device];
        if (!slot) { /* first call on this GPU: resolve the symbol's device address */
            uint32_t *p = nullptr;
            BCP_CUDA(cudaGetSymbolAddress((void **)&p, g_merkle_static_scratch));
            slot = p;
        }
        scratch = slot;
    }
    // Nothing in the array has to start at zero: each launch writes every node the next launch reads, and
    // the mutated flag (caller memory) is cleared by merkle_dev on every call.  Stale nodes of an earlier,
    // larger tree are never read.
    return merkle_dev(c, d_leaves32, n, d_root32, d_mutated, scratch, stream ? (cudaStream_t)stream : c->stream, MERKLE_LEVELS_DEEP);
}

// Host-pointer forms.  Staging layout in c->stage: [data + pad][offsets][digests][merkle tmp][root][mutated]
static int hash_host(bcp_ctx *c, const uint8_t *data, size_t len, const uint32_t *off, size_t count, bool want_root,
                     uint8_t *out32, uint8_t *root32, int *mutated) {
    size_t o_data = 0, o_off = (len + 8 + 255) / 256 * 256, o_out = o_off + ((count + 1) * 4 + 255) / 256 * 256;
    size_t o_tmp = o_out + (count * 32 + 255) / 256 * 256, o_root = o_tmp + (merkle_tmp_words(count) * 4 + 255) / 256 * 256;
    size_t total = o_root + 256;
    int rc = bcp_ctx_reserve_stage(c, total);
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    cudaStream_t st = c->stream;
    if (count) {
        BCP_CUDA(cudaMemcpyAsync(d + o_data, data, len, cudaMemcpyHostToDevice, st));
        BCP_CUDA(cudaMemcpyAsync(d + o_off, off, (count + 1) * 4, cudaMemcpyHostToDevice, st));
        k_sha256d_msgs<<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d + o_data, reinterpret_cast<uint32_t *>(d + o_off),
                                                                       0, 0, count, d + o_out);
        BCP_CUDA(cudaGetLastError());
        if (out32) BCP_CUDA(cudaMemcpyAsync(out32, d + o_out, count * 32, cudaMemcpyDeviceToHost, st));
    }
    if (want_root) {
        rc = merkle_dev(c, d + o_out, count, d + o_root, reinterpret_cast<int *>(d + o_root + 32),
                        reinterpret_cast<uint32_t *>(d + o_tmp), st);
        if (rc != BCP_OK) return rc;
        uint8_t res[36];
        BCP_CUDA(cudaMemcpyAsync(res, d + o_root, 36, cudaMemcpyDeviceToHost, st));
        BCP_CUDA(cudaStreamSynchronize(st));
        memcpy(root32, res, 32);
        int m;
        memcpy(&m, res + 32, 4);
        if (mutated) *mutated = m ? 1 : 0;
    } else {
        BCP_CUDA(cudaStreamSynchronize(st));
    }
    return BCP_OK;
}

BCP_API int bcp_gpu_sha256d_batch(int device, const uint8_t *data, size_t len, const uint32_t *off, size_t count,
                                  uint8_t *out32) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (count && (!data || !off || !out32)) { bcp_set_error("bcp_gpu_sha256d_batch: NULL buffer"); return BCP_ERR_INVALID; }
    if (count && (off[count] > len || len > 0xFFFFFFF0ull)) { bcp_set_error("bcp_gpu_sha256d_batch: offsets exceed buffer"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    return hash_host(c, data, len, off, count, false, out32, nullptr, nullptr);
}

BCP_API int bcp_gpu_sha256d_txids_merkle(int device, const uint8_t *blk, size_t len, const uint32_t *tx_off, size_t ntx,
                                         uint8_t *txids, uint8_t root32[32], int *mutated) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (!root32 || (ntx && (!blk || !tx_off))) { bcp_set_error("bcp_gpu_sha256d_txids_merkle: NULL buffer"); return BCP_ERR_INVALID; }
    if (ntx && (tx_off[ntx] > len || len > 0xFFFFFFF0ull)) { bcp_set_error("bcp_gpu_sha256d_txids_merkle: offsets exceed buffer"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    return hash_host(c, blk, len, tx_off, ntx, true, txids, root32, mutated);
}

BCP_API int bcp_gpu_merkle_root(int device, const uint8_t *leaves32, size_t n, uint8_t root32[32], int *mutated) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (!root32 || (n && !leaves32)) { bcp_set_error("bcp_gpu_merkle_root: NULL buffer"); return BCP_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    size_t o_tmp = (n * 32 + 255) / 256 * 256, o_root = o_tmp + (merkle_tmp_words(n) * 4 + 255) / 256 * 256;
    int rc = bcp_ctx_reserve_stage(c, o_root + 256);
    if (rc != BCP_OK) return rc;
    uint8_t *d = c->stage;
    if (n) BCP_CUDA(cudaMemcpyAsync(d, leaves32, n * 32, cudaMemcpyHostToDevice, c->stream));
    rc = merkle_dev(c, d, n, d + o_root, reinterpret_cast<int *>(d + o_root + 32), reinterpret_cast<uint32_t *>(d + o_tmp),
                    c->stream);
    if (rc != BCP_OK) return rc;
    uint8_t res[36];
    BCP_CUDA(cudaMemcpyAsync(res, d + o_root, 36, cudaMemcpyDeviceToHost, c->stream));
    BCP_CUDA(cudaStreamSynchronize(c->stream));
    memcpy(root32, res, 32);
    int m;
    memcpy(&m, res + 32, 4);
    if (mutated) *mutated = m ? 1 : 0;
    return BCP_OK;
}

} // extern "C"
