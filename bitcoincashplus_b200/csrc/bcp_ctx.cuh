// bcp_ctx.cuh — per-device context shared by the .cu translation units of libbcp_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/bcp_gpu.h"

#define BCP_MAX_DEVICES 16
#define BCP_MAX_LANES 4 /* independent contexts (streams, staging, scratch) per GPU: id = device | lane << 8 */

struct bcp_ctx {
    bool ready = false;
    int device = -1;
    int sm_count = 0;
    cudaStream_t stream = nullptr;      // work stream of the host-pointer entry points
    cudaStream_t copy_stream = nullptr; // second stream for chunk pipelining
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t *gcomb = nullptr;          // fixed-base comb table (BCP_GCOMB_WORDS words)
    uint8_t *scratch = nullptr;         // phase-1 records and other per-call scratch
    size_t scratch_bytes = 0;
    uint8_t *stage = nullptr;           // device staging for the host-pointer entry points
    size_t stage_bytes = 0;
    std::vector<uint32_t> h_tx_off, h_in_off, h_in_tx, h_in_n; // host scratch of the block framing scan (reused)
    bool block_attrs = false;           // shared-memory opt-in of the block kernels done on this device
    bool prep_throughput = false;       // set while a caller with several batches in flight uses this context (see bcp_ecdsa_launch)
    std::mutex mu;
};

bcp_ctx *bcp_get_ctx(int device);            // nullptr (+ last error set) when not initialised
void bcp_set_error(const char *fmt, ...);
int bcp_cuda_fail(cudaError_t e, const char *what);
int bcp_ctx_reserve_scratch(bcp_ctx *c, size_t bytes);
int bcp_ctx_reserve_stage(bcp_ctx *c, size_t bytes);

// internal cross-file entry points
int bcp_ecdsa_launch(bcp_ctx *c, const uint8_t *d_msg, const uint8_t *d_pub, const uint8_t *d_sig, size_t n, uint32_t flags,
                     uint8_t *d_ok, uint32_t *d_scratch, cudaStream_t st);            // bcp_ecdsa.cu
int bcp_sha256d_offsets_dev(bcp_ctx *c, const uint8_t *d_data, const uint32_t *d_off, size_t count, uint8_t *d_out,
                            cudaStream_t st);                                          // bcp_hash.cu
int bcp_merkle_dev(bcp_ctx *c, const uint8_t *d_leaves, size_t n, uint8_t *d_root32, int *d_mut, uint32_t *d_tmp,
                   cudaStream_t st);                                                   // bcp_hash.cu
int bcp_merkle_dev_few_launches(bcp_ctx *c, const uint8_t *d_leaves, size_t n, uint8_t *d_root32, int *d_mut, uint32_t *d_tmp,
                                cudaStream_t st);                                      // bcp_hash.cu
size_t bcp_merkle_tmp_words(size_t n);                                                 // bcp_hash.cu
int bcp_block_scripts_resident(bcp_ctx *c, const uint8_t *d_blk, const uint32_t *d_txoff, const uint32_t *h_tx_off, size_t ntx,
                               const uint32_t *d_inoff, const uint32_t *d_intx, const uint32_t *d_inn, const uint32_t *h_in_tx,
                               size_t nin, const uint8_t *d_spk, const uint32_t *d_spkoff, const int64_t *d_amt, size_t stage_used,
                               uint8_t *script_err, int *all_ok, cudaStream_t st);     // bcp_block.cu
size_t bcp_block_scripts_resident_bytes(size_t ntx, size_t nin);                       // bcp_block.cu
int bcp_block_check_flags(uint32_t flags, const char *fn);                             // bcp_block.cu

#define BCP_CUDA(call)                                          \
    do {                                                        \
        cudaError_t e__ = (call);                               \
        if (e__ != cudaSuccess) return bcp_cuda_fail(e__, #call); \
    } while (0)
