// Shared helpers for libdpvb200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "dpv_b200.h"

namespace dpv {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

// utils/img_utils.py:13 — torch.finfo(float).eps (double eps) used as the fp32 clamp floor
__host__ __device__ constexpr float eps_floor() { return 2.220446049250313e-16f; }

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return DPV_ECUDA;
  }
  count_launch();
  return DPV_OK;
}

// The device's highest kernel scheduling priority.  The latency-bound kernels of the light-curtain loop (a few CTAs,
// hundreds of microseconds of dependent work) are launched with it: when other streams keep the GPU full of
// streaming kernels with tens of thousands of blocks, the block scheduler then hands freed SM resources to these
// CTAs first instead of after the streaming kernel's last wave.
inline int high_priority() {
  static int greatest = [] {
    int least = 0, g = 0;
    if (const char* e = getenv("DPV_PRIORITY"))                      // tuning aid: DPV_PRIORITY=0 launches at default priority
      if (atoi(e) == 0) return 0;
    if (cudaDeviceGetStreamPriorityRange(&least, &g) != cudaSuccess) {
      (void)cudaGetLastError();
      g = 0;
    }
    return g;
  }();
  return greatest;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_high_priority(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                        Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributePriority;
  attr[0].val.priority = high_priority();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define DPV_REQUIRE(cond, ...)      \
  do {                              \
    if (!(cond)) {                  \
      dpv::set_error(__VA_ARGS__);  \
      return DPV_EINVAL;            \
    }                               \
  } while (0)

#define DPV_CUDA(expr)                                              \
  do {                                                              \
    cudaError_t e__ = (expr);                                       \
    if (e__ != cudaSuccess) {                                       \
      dpv::set_error("%s: %s", #expr, cudaGetErrorString(e__));     \
      return DPV_ECUDA;                                             \
    }                                                               \
  } while (0)

// Streaming (read-once / write-once) accesses: keep the 126 MB L2 for data that is re-read.
__device__ __forceinline__ float ld_stream(const float* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(float* p, float v) { __stcs(p, v); }

// p + n elements, opaque to the optimiser: a chain of these stays a chain of 64-bit adds (one IMAD.WIDE each) instead
// of being re-derived from the element index for every plane of an unrolled loop (6 integer instructions per access).
template <typename P>
__device__ __forceinline__ P* bump(P* p, unsigned n) {
  p += n;
  asm("" : "+l"(p));
  return p;
}

// x, with its value hidden from the optimiser (keeps an operand substitution ahead of a division in place)
__device__ __forceinline__ float opaque(float x) {
  asm("" : "+f"(x));
  return x;
}

// 1/x on the MUFU pipe (rcp.approx.ftz: 1 ulp)
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x on the MUFU pipe (ex2.approx.ftz: 2 ulp; denormal results flush to 0)
__device__ __forceinline__ float exp2f_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// log2(x) on the MUFU pipe (lg2.approx.ftz: callers pass x >= eps_floor(), never a denormal)
__device__ __forceinline__ float log2f_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// grid_sample(align_corners=False) un-normalisation, ATen grid_sampler_unnormalize: ((g+1)*N-1)/2.
// The halving is a multiplication by 0.5: the same correctly rounded result as the division for every input (a power
// of two only moves the exponent), without the IEEE division sequence (it was 18 % of ufield_mask's instructions).
__device__ __forceinline__ float gs_unnormalize(float g, int n) {
  return __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(g, 1.f), (float)n), 1.f), 0.5f);
}

}  // namespace dpv
