// bcp_common.cuh — shared macros for the sm_100a device code.
//
// Every arithmetic header in this directory is written once and compiled two ways:
//   * by nvcc for sm_100a (the product): BCP_DEV = __device__ __forceinline__, and the limb
//     primitives are inline PTX (mad.lo.cc / madc.hi.cc chains that ptxas fuses into
//     IMAD.WIDE.U32[.X] — checked with cuobjdump);
//   * by plain g++ for tests/hostsim (a TEST-ONLY harness that lets the CPU-only CI container
//     exercise the very same group law, recoding and verdict logic against the oracle before
//     GPU time is spent).  There the primitives are portable C with identical semantics.
// The product library never contains or calls the host flavour: libbcp_b200.so is built by nvcc
// only and its entry points return BCP_ERR_NO_DEVICE when no GPU is present.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define BCP_DEV __device__ __forceinline__
#define BCP_DEV_M __device__ __forceinline__ /* member functions */
#define BCP_DEV_NOINLINE __device__ __noinline__
#define BCP_CONST __device__ __constant__
#else
#define BCP_DEV static inline
#define BCP_DEV_M inline
#define BCP_DEV_NOINLINE static
#define BCP_CONST static const
#endif

#if defined(__CUDA_ARCH__)
#define BCP_PTX 1
#define BCP_UNROLL _Pragma("unroll")
#else
#define BCP_PTX 0
#define BCP_UNROLL
#endif
