// pipe_mix.cu — does an ALU instruction issued next to IMAD.WIDE.U32.X instructions cost time on sm_100a?
// Each loop iteration issues 16 IMAD.WIDE.U32.X (four independent 4-long carry chains, the form the field code uses)
// and NA independent ALU instructions of one kind; ptxas schedules them (read the SASS: `cuobjdump -sass`).
// Printed: SMSP cycles per iteration per warp.  64 = the multiplier pipe alone (4 cycles per IMAD.WIDE per SMSP).
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o oracle/_ref/pipe_mix tools/ubench/pipe_mix.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void chain4(uint32_t *x, uint32_t m, uint32_t a, uint32_t b) {
    asm volatile("mad.lo.cc.u32 %0, %8, %9, %0;\n\tmadc.hi.cc.u32 %1, %8, %9, %1;\n\t"
                 "madc.lo.cc.u32 %2, %8, %10, %2;\n\tmadc.hi.cc.u32 %3, %8, %10, %3;\n\t"
                 "madc.lo.cc.u32 %4, %8, %9, %4;\n\tmadc.hi.cc.u32 %5, %8, %9, %5;\n\t"
                 "madc.lo.cc.u32 %6, %8, %10, %6;\n\tmadc.hi.u32 %7, %8, %10, %7;"
                 : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7])
                 : "r"(m), "r"(a), "r"(b));
}
// KIND 1: LOP3, 2: SHF (funnel), 3: IADD3.X carry chain of 8, 4: SEL, 5: plain add (ptxas may pick IADD3 or IMAD.IADD)
template <int KIND> __device__ __forceinline__ void alu8(uint32_t *y, uint32_t s) {
    if (KIND == 1) {
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("xor.b32 %0, %0, %1;" : "+r"(y[j]) : "r"(y[(j + 3) & 7] ));
    } else if (KIND == 2) {
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(y[j]) : "r"(y[(j + 3) & 7]));
    } else if (KIND == 3) {
        asm volatile("add.cc.u32 %0, %0, %8;\n\taddc.cc.u32 %1, %1, %8;\n\taddc.cc.u32 %2, %2, %8;\n\taddc.cc.u32 %3, %3, %8;\n\t"
                     "addc.cc.u32 %4, %4, %8;\n\taddc.cc.u32 %5, %5, %8;\n\taddc.cc.u32 %6, %6, %8;\n\taddc.u32 %7, %7, %8;"
                     : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7]) : "r"(s));
    } else if (KIND == 4) {
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("{.reg .pred p; setp.lt.u32 p, %1, %2; selp.b32 %0, %0, %1, p;}" : "+r"(y[j]) : "r"(y[(j + 3) & 7]), "r"(s));
    } else if (KIND == 5) {
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("add.u32 %0, %0, %1;" : "+r"(y[j]) : "r"(y[(j + 3) & 7]));
    }
}

template <int KIND, int NA8> // NA8 groups of 8 ALU instructions per iteration
__global__ void __launch_bounds__(256, 2) k(uint32_t *out, uint32_t a, uint32_t b, int iters) {
    uint32_t x[32], y[16];
#pragma unroll
    for (int i = 0; i < 32; i++) x[i] = threadIdx.x * 32 + i + a;
#pragma unroll
    for (int i = 0; i < 16; i++) y[i] = x[i] ^ b;
    uint32_t m = b | 1u;
#pragma unroll 4
    for (int it = 0; it < iters; it++) {
        chain4(x, m, a, b);
        if (NA8 > 0) alu8<KIND>(y, m);
        chain4(x + 8, m, a, b);
        if (NA8 > 1) alu8<KIND>(y + 8, m);
        chain4(x + 16, m, a, b);
        if (NA8 > 2) alu8<KIND>(y, m);
        chain4(x + 24, m, a, b);
        if (NA8 > 3) alu8<KIND>(y + 8, m);
        if (NA8 > 4) alu8<KIND>(y, m);
        if (NA8 > 5) alu8<KIND>(y + 8, m);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 32; i++) s += x[i];
#pragma unroll
    for (int i = 0; i < 16; i++) s += y[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int KIND, int NA8> static void run(const char *name, int threads, int sms, double mhz) {
    const int blocks = sms * 2, iters = 20000;
    uint32_t *out;
    cudaMalloc(&out, (size_t)blocks * threads * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0);
        k<KIND, NA8><<<blocks, threads>>>(out, 3, 5, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double warps_per_smsp = 2.0 * threads / 32 / 4;
    const double cyc = best * 1e-3 * mhz * 1e6 / (iters * warps_per_smsp);
    printf("%-10s +%2d ALU  %3d thr/block (%3.1f warps/SMSP): %7.2f SMSP cycles / iteration / warp   [%s]\n", name, 8 * NA8, threads,
           warps_per_smsp, cyc, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main(int argc, char **argv) {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = argc > 1 ? atof(argv[1]) : khz / 1e3;
    printf("%s, %d SMs, clock used for the cycle count: %.0f MHz; every row = 16 IMAD.WIDE.U32[.X] per iteration + the ALU instructions named\n", p.name, p.multiProcessorCount, mhz);
    for (int threads : {256, 128, 64}) {
        run<0, 0>("wide only", threads, p.multiProcessorCount, mhz);
        run<1, 1>("LOP3", threads, p.multiProcessorCount, mhz);
        run<1, 2>("LOP3", threads, p.multiProcessorCount, mhz);
        run<1, 4>("LOP3", threads, p.multiProcessorCount, mhz);
        run<1, 6>("LOP3", threads, p.multiProcessorCount, mhz);
        run<2, 2>("SHF", threads, p.multiProcessorCount, mhz);
        run<2, 4>("SHF", threads, p.multiProcessorCount, mhz);
        run<3, 2>("IADD3.X", threads, p.multiProcessorCount, mhz);
        run<3, 4>("IADD3.X", threads, p.multiProcessorCount, mhz);
        run<3, 6>("IADD3.X", threads, p.multiProcessorCount, mhz);
        run<4, 2>("SETP+SEL", threads, p.multiProcessorCount, mhz);
        run<5, 2>("add", threads, p.multiProcessorCount, mhz);
        run<5, 4>("add", threads, p.multiProcessorCount, mhz);
    }
    return 0;
}
