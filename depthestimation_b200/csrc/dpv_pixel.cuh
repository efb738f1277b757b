// Per-pixel DPV arithmetic on register-resident depth columns (one thread = one pixel, DPT >= D bins
// held in registers).  Shared by the log-softmax, Bayesian-fusion and sense+fuse kernels.
#pragma once
#include "common.cuh"

namespace dpv {

// utils/img_utils.py:408-436 spread_dpv_hack on one pixel: q[d] = sum_{|j|<=HALF} p[d+j] (zero pad),
// then q /= sum_d q.  In place, sliding window (no second register array).
template <int DPT, int HALF>
__device__ __forceinline__ float spread_sum(float (&p)[DPT], int D) {
  float prev[HALF > 0 ? HALF : 1];
#pragma unroll
  for (int j = 0; j < HALF; j++) prev[j] = 0.f;
  float total = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++) {
    if (d < D) {
      float c = p[d];
      float q = 0.f;
#pragma unroll
      for (int j = HALF - 1; j >= 0; j--) q += prev[j];     // p[d-HALF] .. p[d-1]
      q += c;
#pragma unroll
      for (int j = 1; j <= HALF; j++)
        if (d + j < DPT) q += (d + j < D) ? p[d + j] : 0.f;
#pragma unroll
      for (int j = HALF - 1; j > 0; j--) prev[j] = prev[j - 1];
      if (HALF > 0) prev[0] = c;
      p[d] = q;
      total += q;
    }
  }
  return total;
}

template <int DPT, int HALF>
__device__ __forceinline__ void spread_renorm(float (&p)[DPT], int D) {
  const float total = spread_sum<DPT, HALF>(p, D);
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) p[d] = p[d] / total;
}

template <int DPT>
__device__ __forceinline__ bool spread_dispatch(float (&p)[DPT], int D, int N) {
  switch (N) {
    case 1: spread_renorm<DPT, 0>(p, D); return true;
    case 3: spread_renorm<DPT, 1>(p, D); return true;
    case 5: spread_renorm<DPT, 2>(p, D); return true;
    case 7: spread_renorm<DPT, 3>(p, D); return true;
    default: return false;
  }
}

// models/models.py:1097: curr = clamp(exp(final), eps, 1)
template <int DPT>
__device__ __forceinline__ void load_prob_from_log(float (&p)[DPT], const float* __restrict__ src, size_t stride,
                                                   int D) {
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) p[d] = ld_stream(src + (size_t)d * stride);
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) p[d] = fminf(fmaxf(expf(p[d]), eps_floor()), 1.f);
}

// models/models.py:1119-1122: final = log(clamp(curr, eps, 1)); also E[d] of the stored log-DPV
// (dpv_to_depthmap of exp(final)), accumulated from the clamped probability.
template <int DPT>
__device__ __forceinline__ float store_log_and_mean(const float (&p)[DPT], float* __restrict__ dst, size_t stride,
                                                    int D, const float* __restrict__ dc) {
  float mean = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) {
      float c = fminf(fmaxf(p[d], eps_floor()), 1.f);
      mean += dc[d] * c;
      st_stream(dst + (size_t)d * stride, logf(c));
    }
  return mean;
}

}  // namespace dpv
