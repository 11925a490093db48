// bcp_ctx.cu — context life cycle, error reporting, pinned memory, and the integer-pipe peak
// microbenchmark of libbcp_b200.so.  See include/bcp_gpu.h for the contract of each entry point.
#include "bcp_ctx.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>

int bcp_ecdsa_build_tables(bcp_ctx *c); // bcp_ecdsa.cu
int bcp_fe_selftest(bcp_ctx *c);          // bcp_ecdsa.cu

static bcp_ctx g_ctx[BCP_MAX_DEVICES][BCP_MAX_LANES];
static std::mutex g_init_mu;
static thread_local char g_err[512] = "";

void bcp_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
int bcp_cuda_fail(cudaError_t e, const char *what) {
    bcp_set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    cudaGetLastError(); // clear the sticky-less error state
    return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? BCP_ERR_NO_DEVICE : BCP_ERR_CUDA;
}
bcp_ctx *bcp_get_ctx(int id) {
    const int device = id & 0xFF, lane = id >> 8;
    if (id < 0 || device >= BCP_MAX_DEVICES || lane >= BCP_MAX_LANES || !g_ctx[device][lane].ready) {
        bcp_set_error("device %d lane %d not initialised: call bcp_gpu_init first", device, lane);
        return nullptr;
    }
    return &g_ctx[device][lane];
}
int bcp_ctx_reserve_scratch(bcp_ctx *c, size_t bytes) {
    if (bytes <= c->scratch_bytes) return BCP_OK;
    BCP_CUDA(cudaStreamSynchronize(c->stream));
    if (c->scratch) BCP_CUDA(cudaFree(c->scratch));
    c->scratch = nullptr;
    c->scratch_bytes = 0;
    BCP_CUDA(cudaMalloc(&c->scratch, bytes));
    c->scratch_bytes = bytes;
    return BCP_OK;
}
int bcp_ctx_reserve_stage(bcp_ctx *c, size_t bytes) {
    if (bytes <= c->stage_bytes) return BCP_OK;
    BCP_CUDA(cudaStreamSynchronize(c->stream));
    if (c->stage) BCP_CUDA(cudaFree(c->stage));
    c->stage = nullptr;
    c->stage_bytes = 0;
    BCP_CUDA(cudaMalloc(&c->stage, bytes));
    c->stage_bytes = bytes;
    return BCP_OK;
}

extern "C" {

const char *bcp_gpu_last_error(void) { return g_err; }

int bcp_gpu_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        bcp_cuda_fail(e, "cudaGetDeviceCount");
        return 0;
    }
    return n;
}

// 1 when GPU a and GPU b can read and write each other's memory directly (NVLink / NVSwitch or PCIe peer-to-peer), 0 when
// a copy between them would be staged through host memory, negative on error.  a == b is 1.
int bcp_gpu_peer_access(int a, int b) {
    a &= 0xFF; b &= 0xFF;
    if (a == b) return 1;
    int ab = 0, ba = 0;
    if (cudaDeviceCanAccessPeer(&ab, a, b) != cudaSuccess || cudaDeviceCanAccessPeer(&ba, b, a) != cudaSuccess) {
        cudaGetLastError();
        return BCP_ERR_CUDA;
    }
    return ab && ba ? 1 : 0;
}

static int init_locked(int device, int lane);

int bcp_gpu_init(int id) {
    std::lock_guard<std::mutex> lk(g_init_mu);
    const int device = id & 0xFF, lane = id >> 8;
    if (id < 0 || lane >= BCP_MAX_LANES) {
        bcp_set_error("bad device id 0x%x (device | lane << 8, lanes 0..%d)", id, BCP_MAX_LANES - 1);
        return BCP_ERR_INVALID;
    }
    if (lane > 0) { // lane 0 owns the comb table every lane of the GPU reads
        int rc = init_locked(device, 0);
        if (rc != BCP_OK) return rc;
    }
    return init_locked(device, lane);
}

static int init_locked(int device, int lane) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        bcp_set_error("no CUDA device available (%s): this library has no CPU fallback",
                      e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        cudaGetLastError();
        return BCP_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n || device >= BCP_MAX_DEVICES) {
        bcp_set_error("device ordinal %d out of range (0..%d)", device, n - 1);
        return BCP_ERR_NO_DEVICE;
    }
    bcp_ctx *c = &g_ctx[device][lane];
    if (c->ready) return BCP_OK;
    BCP_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    BCP_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        bcp_set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major,
                      prop.minor);
        return BCP_ERR_NO_DEVICE;
    }
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    BCP_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    BCP_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; i++) // ev[3] is the one host threads sleep on (blocking sync): pipelines with more threads than cores
        BCP_CUDA(cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming | (i == 3 ? cudaEventBlockingSync : 0)));
    if (lane == 0) {
        int rc = bcp_fe_selftest(c); // a few hundred crafted operand pairs, ~0.1 ms: refuse to start on a wrong answer
        if (rc != BCP_OK) return rc;
        rc = bcp_ecdsa_build_tables(c);
        if (rc != BCP_OK) return rc;
    } else {
        c->gcomb = g_ctx[device][0].gcomb;
    }
    c->ready = true;
    return BCP_OK;
}

int bcp_gpu_shutdown(int id) { // lane 0 (the plain device ordinal) takes the GPU's other lanes down with it
    std::lock_guard<std::mutex> lk(g_init_mu);
    const int device = id & 0xFF, lane0 = id >> 8;
    if (id < 0 || device >= BCP_MAX_DEVICES || lane0 >= BCP_MAX_LANES) return BCP_OK;
    for (int lane = BCP_MAX_LANES - 1; lane >= lane0; lane--) {
        bcp_ctx *c = &g_ctx[device][lane];
        if (!c->ready) continue;
        if (lane != lane0 && lane0 != 0) continue;
        std::lock_guard<std::mutex> lk2(c->mu);
        cudaSetDevice(device);
        cudaStreamSynchronize(c->stream);
        cudaStreamSynchronize(c->copy_stream);
        if (c->gcomb && lane == 0) cudaFree(c->gcomb);
        if (c->scratch) cudaFree(c->scratch);
        if (c->stage) cudaFree(c->stage);
        for (int i = 0; i < 4; i++) cudaEventDestroy(c->ev[i]);
        cudaStreamDestroy(c->stream);
        cudaStreamDestroy(c->copy_stream);
        c->gcomb = nullptr; c->scratch = nullptr; c->stage = nullptr;
        c->scratch_bytes = c->stage_bytes = 0;
        c->ready = false;
    }
    return BCP_OK;
}

int bcp_gpu_selftest(int device) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    return bcp_fe_selftest(c);
}

void *bcp_gpu_host_alloc(size_t bytes) {
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        bcp_cuda_fail(e, "cudaHostAlloc");
        return nullptr;
    }
    return p;
}
void bcp_gpu_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

// ecdsa_signature_parse_der_lax (src/pubkey.cpp:26-168): byte-level, branchy, stays on the host.
int bcp_der_lax_parse(const uint8_t *input, size_t inputlen, uint8_t sig64[64]) {
    size_t pos = 0, lenbyte, vpos[2] = {0, 0}, vlen[2] = {0, 0};
    memset(sig64, 0, 64);
    if (pos == inputlen || input[pos] != 0x30) return 0;
    pos++;
    if (pos == inputlen) return 0;
    lenbyte = input[pos++];
    if (lenbyte & 0x80) { /* sequence length: skipped, never checked */
        lenbyte -= 0x80;
        if (pos + lenbyte > inputlen) return 0;
        pos += lenbyte;
    }
    for (int k = 0; k < 2; k++) {
        if (pos == inputlen || input[pos] != 0x02) return 0;
        pos++;
        if (pos == inputlen) return 0;
        lenbyte = input[pos++];
        size_t len;
        if (lenbyte & 0x80) {
            lenbyte -= 0x80;
            if (pos + lenbyte > inputlen) return 0;
            while (lenbyte > 0 && input[pos] == 0) { pos++; lenbyte--; }
            if (lenbyte >= sizeof(size_t)) return 0;
            len = 0;
            while (lenbyte > 0) { len = (len << 8) + input[pos]; pos++; lenbyte--; }
        } else {
            len = lenbyte;
        }
        if (len > inputlen - pos) return 0;
        vpos[k] = pos;
        vlen[k] = len;
        pos += len;
    }
    bool overflow = false;
    for (int k = 0; k < 2; k++) {
        size_t p = vpos[k], l = vlen[k];
        while (l > 0 && input[p] == 0) { l--; p++; }
        if (l > 32) overflow = true;
        else memcpy(sig64 + 32 * k + 32 - l, input + p, l);
    }
    if (!overflow) {
        /* r or s >= n: secp256k1_ecdsa_signature_parse_compact fails and the signature is zeroed */
        static const uint8_t N_BE[32] = {0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF,
                                         0xFF, 0xFF, 0xFF, 0xFF, 0xFE, 0xBA, 0xAE, 0xDC, 0xE6, 0xAF, 0x48,
                                         0xA0, 0x3B, 0xBF, 0xD2, 0x5E, 0x8C, 0xD0, 0x36, 0x41, 0x41};
        for (int k = 0; k < 2; k++)
            if (memcmp(sig64 + 32 * k, N_BE, 32) >= 0) overflow = true;
    }
    if (overflow) memset(sig64, 0, 64);
    return 1;
}

} // extern "C"

// ------------------------------------------------------------------ integer-pipe peak
// Each thread keeps 8 independent accumulator chains so the pipe, not latency, is the limit.
template <int WHICH>
__global__ void __launch_bounds__(256) int_peak_kernel(uint32_t *out, uint32_t a, uint32_t b, int iters) {
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = threadIdx.x * 16 + i + a;
    uint32_t m = b | 1u;
    for (int it = 0; it < iters; it++) {
        if (WHICH == 0) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    unsigned long long acc;
                    asm volatile("mov.b64 %0, {%1, %2};" : "=l"(acc) : "r"(x[i]), "r"(x[i + 1]));
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(m), "r"(x[(i + 3) & 15]));
                    asm volatile("mov.b64 {%0, %1}, %2;" : "=r"(x[i]), "=r"(x[i + 1]) : "l"(acc));
                }
            }
        } else if (WHICH == 1) {
#pragma unroll
            for (int k = 0; k < 2; k++) {
                asm volatile(
                    "mad.lo.cc.u32 %0, %16, %17, %0;\n\tmadc.hi.cc.u32 %1, %16, %17, %1;\n\t"
                    "madc.lo.cc.u32 %2, %16, %18, %2;\n\tmadc.hi.cc.u32 %3, %16, %18, %3;\n\t"
                    "madc.lo.cc.u32 %4, %16, %17, %4;\n\tmadc.hi.cc.u32 %5, %16, %17, %5;\n\t"
                    "madc.lo.cc.u32 %6, %16, %18, %6;\n\tmadc.hi.u32 %7, %16, %18, %7;\n\t"
                    "mad.lo.cc.u32 %8, %16, %17, %8;\n\tmadc.hi.cc.u32 %9, %16, %17, %9;\n\t"
                    "madc.lo.cc.u32 %10, %16, %18, %10;\n\tmadc.hi.cc.u32 %11, %16, %18, %11;\n\t"
                    "madc.lo.cc.u32 %12, %16, %17, %12;\n\tmadc.hi.cc.u32 %13, %16, %17, %13;\n\t"
                    "madc.lo.cc.u32 %14, %16, %18, %14;\n\tmadc.hi.u32 %15, %16, %18, %15;"
                    : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]),
                      "+r"(x[8]), "+r"(x[9]), "+r"(x[10]), "+r"(x[11]), "+r"(x[12]), "+r"(x[13]), "+r"(x[14]), "+r"(x[15])
                    : "r"(m), "r"(a), "r"(b));
            }
        } else if (WHICH == 2) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int i = 0; i < 16; i++) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(m));
            }
        } else if (WHICH == 3) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int i = 0; i < 16; i++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(a));
            }
        } else if (WHICH == 4) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int i = 0; i < 16; i++) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(a));
            }
        } else if (WHICH == 5) {
            double *d = reinterpret_cast<double *>(x); /* 8 independent DFMA chains */
            double dm = __longlong_as_double(0x3ff0000000000001LL + m);
#pragma unroll
            for (int k = 0; k < 8; k++) {
#pragma unroll
                for (int i = 0; i < 8; i++) asm volatile("fma.rn.f64 %0, %0, %1, %1;" : "+d"(d[i]) : "d"(dm));
            }
        } else {
            float *f = reinterpret_cast<float *>(x);
            float fm = __int_as_float(0x3f800001 + m);
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int i = 0; i < 16; i++) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(f[i]) : "f"(fm));
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" int bcp_gpu_measure_int_peak(int device, int which, double *ops_per_second) {
    bcp_ctx *c = bcp_get_ctx(device);
    if (!c) return BCP_ERR_NOT_INIT;
    if (!ops_per_second || which < 0 || which > 6) return BCP_ERR_INVALID;
    std::lock_guard<std::mutex> lk(c->mu);
    BCP_CUDA(cudaSetDevice(c->device));
    const int blocks = c->sm_count * 8, threads = 256, iters = 4096;
    uint32_t *out = nullptr;
    BCP_CUDA(cudaMalloc(&out, (size_t)blocks * threads * 4));
    cudaEvent_t e0, e1;
    BCP_CUDA(cudaEventCreate(&e0));
    BCP_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        BCP_CUDA(cudaEventRecord(e0, c->stream));
        if (which == 0) int_peak_kernel<0><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        else if (which == 1) int_peak_kernel<1><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        else if (which == 2) int_peak_kernel<2><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        else if (which == 3) int_peak_kernel<3><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        else if (which == 4) int_peak_kernel<4><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        else if (which == 5) int_peak_kernel<5><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        else int_peak_kernel<6><<<blocks, threads, 0, c->stream>>>(out, 3, 5, iters);
        BCP_CUDA(cudaEventRecord(e1, c->stream));
        BCP_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        BCP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    BCP_CUDA(cudaGetLastError());
    // per thread per iteration: which 0: 4*8 = 32 wide MADs; which 1: 2*8 = 16 fused lo/hi pairs
    // (= 16 IMAD.WIDE); which 2: 64 adds
    double per_iter = which == 0 ? 32.0 : which == 1 ? 16.0 : 64.0; /* which 2,3,4,6: 64 ops; 5: 64 DFMA */
    *ops_per_second = per_iter * iters * (double)blocks * threads / (best * 1e-3);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return BCP_OK;
}
