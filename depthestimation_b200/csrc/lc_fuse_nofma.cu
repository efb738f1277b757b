// (d4) One LC iteration's sensing + Bayesian update in a single pass over the DPV.
//   models/models.py:1064-1122 (sense each curtain, clamp, multiply, renormalise, spread, clamp, log)
//   with lc/light_curtain.py:847-888 + :228-252 (likelihood) and dprocessor.cpp:806-993 (image formation)
//   evaluated per pixel in registers.  Compiled with -fmad=false for the geometry (see lc_sense.cuh).
//
// A thread owns one pixel: its DPV column p[D] and the current curtain's likelihood column g[D] live in
// registers (straight-line code over the bins; at the simulated baseline of 0.2 m the curtain thickness
// grows to metres, so the Gaussian really is dense over the bins and is evaluated for every bin with one
// MUFU ex2 each).  Multiply, box-spread and renormalise are all homogeneous in the DPV, so the
// per-curtain and per-spread renormalisations of the reference collapse into one scale at the end
// (identical up to fp32 rounding; an extra rescale is inserted only if the running sum leaves fp32 range).
// The volume may be kept as log-probability (the reference's format) or as clamped probability between
// the iterations of one lc_process call (flags), which removes exp/log from the inner iterations.
#include "lc_sense.cuh"

namespace dpv {

constexpr int kFuseThreads = 128;
constexpr int kMaxCurtains = 8;

// g[d] = exp(-(dc-z)^2 / den) as ex2((dc-z)^2 * kk), kk = -log2(e)/den  (img_utils.py:183-186)
__device__ __forceinline__ float gauss_bin(float dcv, float z, float kk) {
  const float t = dcv - z;
  return exp2f_approx((t * t) * kk);
}

// sum over the T lanes that share one pixel (T is a power of two; lanes are adjacent)
template <int T>
__device__ __forceinline__ float group_sum(float v, unsigned gmask) {
#pragma unroll
  for (int off = T / 2; off > 0; off >>= 1) v += __shfl_xor_sync(gmask, v, off);
  return v;
}

// One box-sum pass along depth over a column that is split over T lanes (DPT bins each, lane t owns bins
// [t*DPT, (t+1)*DPT)).  q[d] = sum_{|j|<=HALF} p[d+j], zero outside [0,D).  img_utils.py:408-436.
template <int DPT, int T, int HALF>
__device__ __forceinline__ void spread_pass(float (&p)[DPT], int d0, int D, int t, int lane, unsigned gmask) {
  float lo_h[HALF > 0 ? HALF : 1], hi_h[HALF > 0 ? HALF : 1];     // neighbours' edge bins (before modification)
#pragma unroll
  for (int j = 0; j < HALF; j++) {
    float from_prev = 0.f, from_next = 0.f;
    if (T > 1) {
      from_prev = __shfl_sync(gmask, p[DPT - 1 - j], (lane - 1) & 31);     // p[d0-1-j]
      from_next = __shfl_sync(gmask, p[j], (lane + 1) & 31);               // p[d0+DPT+j]
    }
    lo_h[j] = (T > 1 && t > 0) ? from_prev : 0.f;
    hi_h[j] = (T > 1 && t < T - 1 && d0 + DPT + j < D) ? from_next : 0.f;
  }
  float prev[HALF > 0 ? HALF : 1];
#pragma unroll
  for (int j = 0; j < HALF; j++) prev[j] = lo_h[j];                        // prev[j] = p[d-1-j]
#pragma unroll
  for (int d = 0; d < DPT; d++) {
    const float c = p[d];
    float q = 0.f;
#pragma unroll
    for (int j = HALF - 1; j >= 0; j--) q += prev[j];
    q += c;
#pragma unroll
    for (int j = 1; j <= HALF; j++) {
      const float nb = (d + j < DPT) ? ((d0 + d + j < D) ? p[(d + j < DPT) ? d + j : 0] : 0.f) : hi_h[(d + j - DPT) < HALF ? (d + j - DPT) : 0];
      q += nb;
    }
#pragma unroll
    for (int j = HALF - 1; j > 0; j--) prev[j] = prev[j - 1];
    if (HALF > 0) prev[0] = c;
    p[d] = (d0 + d < D) ? q : 0.f;
  }
}

template <int DPT, int T, int HALF>
__global__ void __launch_bounds__(kFuseThreads, 4)
sense_fuse_kernel(LcDeviceView dv, const float* __restrict__ vol_in, const float* __restrict__ depth,
                  const float* __restrict__ planes, const float* __restrict__ noise, int nc, long long npix,
                  const float* __restrict__ d_candi, int D, int spread_iter, int flags,
                  float* __restrict__ vol_out, float* __restrict__ mean_out) {
  __shared__ float dc[DPT * T];
  for (int i = threadIdx.x; i < DPT * T; i += blockDim.x) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long pix = gid / T;
  if (pix >= npix) return;                                   // whole lane groups leave together
  const int lane = threadIdx.x & 31;
  const int t = lane & (T - 1);
  const unsigned gmask = (T == 32 ? 0xffffffffu : ((1u << T) - 1u)) << (lane & ~(T - 1));
  const int d0 = t * DPT;
  const float* dcl = dc + d0;
  const size_t HW = (size_t)dv.H * dv.W;
  const int b = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)b * HW);
  const int v = (int)(hw / dv.W), u = (int)(hw % dv.W);
  const size_t base = ((size_t)b * D + d0) * HW + hw;
  const bool in_prob = flags & 1, out_prob = flags & 2;

  // ---- issue the volume loads first; the curtain geometry below overlaps their latency
  float p[DPT];
#pragma unroll
  for (int d = 0; d < DPT; d++) p[d] = (d0 + d < D) ? ld_stream(vol_in + base + (size_t)d * HW) : 0.f;
  const float gt = depth[pix];
  bool first = true;
  Lik mine;
  mine.z = 0.f; mine.den = 0.f; mine.A = 0.f; mine.Bu = 0.f;

  for (int c = 0; c < nc; c++) {
    // ---- sensing: in rounds of T curtains, lane t of the group senses curtain c + t (all lanes busy, same code)
    if ((c & (T - 1)) == 0) {
      const int myc = c + t;
      if (myc < nc) {
        const int cu = b * nc + myc;
        float Pl[12];
#pragma unroll
        for (int k = 0; k < 12; k++) Pl[k] = __ldg(planes + ((size_t)cu * dv.W + u) * 12 + k);
        const Ret r = sense_pixel(dv, Pl, u, v, gt, noise ? noise + (size_t)cu * HW + hw : nullptr);
        mine = lik_params(r, gt, D);
      }
    }
    Lik L;
    const int owner = (lane & ~(T - 1)) + (c & (T - 1));
    L.z = __shfl_sync(gmask, mine.z, owner);
    L.den = __shfl_sync(gmask, mine.den, owner);
    L.A = __shfl_sync(gmask, mine.A, owner);
    L.Bu = __shfl_sync(gmask, mine.Bu, owner);
    if (first) {                                             // models.py:1097 (after the loads have landed)
#pragma unroll
      for (int d = 0; d < DPT; d++) p[d] = fminf(fmaxf(in_prob ? p[d] : __expf(p[d]), eps_floor()), 1.f);
      first = false;
    }
    // ---- likelihood of this curtain over the depth bins (mixed_model, light_curtain.py:247-252)
    const float kk = -1.4426950408889634f / L.den;           // den = 0 -> -inf: g = 0 (NaN at dc == z) like the reference
    float g[DPT];
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d += 2) {
      g[d] = (d0 + d < D) ? gauss_bin(dcl[d], L.z, kk) : 0.f;
      g[d + 1] = (d0 + d + 1 < D) ? gauss_bin(dcl[d + 1], L.z, kk) : 0.f;
      s0 += g[d];
      s1 += g[d + 1];
    }
    const float sg = group_sum<T>(s0 + s1, gmask);
    // sg == 0 or NaN: g/sg is NaN -> -1 in every bin -> constant mixture -> uniform likelihood -> no-op
    if (!(sg > 0.f)) continue;
    const float As = L.A / sg;
    s0 = 0.f; s1 = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d += 2) {
      g[d] = (d0 + d < D) ? fmaxf(fmaf(g[d], As, L.Bu), 0.f) : 0.f;     // clamp(soft * A + uni * B, 0, inf)
      g[d + 1] = (d0 + d + 1 < D) ? fmaxf(fmaf(g[d + 1], As, L.Bu), 0.f) : 0.f;
      s0 += g[d];
      s1 += g[d + 1];
    }
    const float sm = group_sum<T>(s0 + s1, gmask);
    if (!(sm > 0.f)) continue;                               // 0/0 in every bin -> NaN -> 0 -> eps: uniform again
    const float inv_sm = 1.f / sm;
    s0 = 0.f; s1 = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d += 2) {
      p[d] *= fminf(fmaxf(g[d] * inv_sm, eps_floor()), 1.f);             // models.py:1109-1111
      p[d + 1] *= fminf(fmaxf(g[d + 1] * inv_sm, eps_floor()), 1.f);
      s0 += p[d];
      s1 += p[d + 1];
    }
    const float s = group_sum<T>(s0 + s1, gmask);
    if (s < 1e-12f) {                                        // models.py:1112 — only needed to stay in range
      const float inv = 1.f / s;
#pragma unroll
      for (int d = 0; d < DPT; d++) p[d] *= inv;
    }
  }
  if (first) {
#pragma unroll
    for (int d = 0; d < DPT; d++) p[d] = fminf(fmaxf(in_prob ? p[d] : __expf(p[d]), eps_floor()), 1.f);
  }
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d0 + d >= D) p[d] = 0.f;

  // ---- box spread along depth x spread_iter (img_utils.py:408-436), one renormalisation at the end
  for (int i = 0; i < spread_iter; i++) spread_pass<DPT, T, HALF>(p, d0, D, t, lane, gmask);
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d += 2) { s0 += p[d]; s1 += p[d + 1]; }
  const float inv = 1.f / group_sum<T>(s0 + s1, gmask);

  // ---- clamp, (log), expectation, store (models.py:1119-1122) -------------------------------
  float m0 = 0.f, m1 = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d += 2) {
    const float c0 = fminf(fmaxf(p[d] * inv, eps_floor()), 1.f);
    const float c1 = fminf(fmaxf(p[d + 1] * inv, eps_floor()), 1.f);
    if (d0 + d < D) {
      m0 = fmaf(dcl[d], c0, m0);
      st_stream(vol_out + base + (size_t)d * HW, out_prob ? c0 : __logf(c0));
    }
    if (d0 + d + 1 < D) {
      m1 = fmaf(dcl[d + 1], c1, m1);
      st_stream(vol_out + base + (size_t)(d + 1) * HW, out_prob ? c1 : __logf(c1));
    }
  }
  const float mean = group_sum<T>(m0 + m1, gmask);
  if (mean_out && t == 0) mean_out[pix] = mean;
}

}  // namespace dpv

using namespace dpv;

extern "C" int lc_sense_fuse(const LcDevice* dev, const float* logp_in, const float* depth, const float* planes,
                             const float* noise, int B, int nc, const float* d_candi, int D, int spread_iter,
                             int spread_n, int flags, float* logp_out, float* mean_out, void* stream) {
  DPV_REQUIRE(dev && logp_in && depth && planes && d_candi && logp_out, "lc_sense_fuse: null pointer");
  DPV_REQUIRE(B > 0 && nc >= 0 && nc <= kMaxCurtains && D > 0 && spread_iter >= 0,
              "lc_sense_fuse: bad shape (B=%d nc=%d D=%d)", B, nc, D);
  DPV_REQUIRE(spread_iter == 0 || spread_n == 1 || spread_n == 3 || spread_n == 5 || spread_n == 7,
              "lc_sense_fuse: spread_n=%d not in {1,3,5,7}", spread_n);
  DPV_REQUIRE(D <= 128, "lc_sense_fuse: D=%d > 128 unsupported", D);
  DPV_REQUIRE((flags & ~3) == 0, "lc_sense_fuse: unknown flags %d", flags);
  const long long npix = (long long)B * dev->H * dev->W;
  cudaStream_t st = (cudaStream_t)stream;
  const int half = spread_iter == 0 ? 0 : spread_n / 2;
#define LAUNCH3(DPT, T, HALF)                                                                                      \
  sense_fuse_kernel<DPT, T, HALF><<<(unsigned)((npix * T + kFuseThreads - 1) / kFuseThreads), kFuseThreads, 0, st>>>( \
      dev->view, logp_in, depth, planes, noise, nc, npix, d_candi, D, spread_iter, flags, logp_out, mean_out)
#define LAUNCH(DPT, T)                       \
  do {                                       \
    if (half == 0) LAUNCH3(DPT, T, 0);       \
    else if (half == 1) LAUNCH3(DPT, T, 1);  \
    else if (half == 2) LAUNCH3(DPT, T, 2);  \
    else LAUNCH3(DPT, T, 3);                 \
  } while (0)
  // each pixel's depth column is split over T adjacent lanes, at most 32 bins per lane (register budget)
  if (D <= 16) LAUNCH(16, 1);
  else if (D <= 32) LAUNCH(32, 1);
  else if (D <= 64) LAUNCH(32, 2);
  else LAUNCH(32, 4);
#undef LAUNCH
#undef LAUNCH3
  return check_launch("lc_sense_fuse");
}
