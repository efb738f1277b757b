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

template <int DPT, bool KEEP_G>
__global__ void __launch_bounds__(kFuseThreads)
sense_fuse_kernel(LcDeviceView dv, const float* __restrict__ vol_in, const float* __restrict__ depth,
                  const float* __restrict__ planes, const float* __restrict__ noise, int nc, long long npix,
                  const float* __restrict__ d_candi, int D, int spread_iter, int spread_n, int flags,
                  float* __restrict__ vol_out, float* __restrict__ mean_out) {
  __shared__ float dc[DPT];
  for (int i = threadIdx.x; i < DPT; i += blockDim.x) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const size_t HW = (size_t)dv.H * dv.W;
  const int b = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)b * HW);
  const int v = (int)(hw / dv.W), u = (int)(hw % dv.W);
  const size_t base = (size_t)b * D * HW + hw;
  const bool in_prob = flags & 1, out_prob = flags & 2;

  // ---- issue the volume loads first; the geometry of the first curtain overlaps their latency
  float p[DPT];
#pragma unroll
  for (int d = 0; d < DPT; d++) p[d] = d < D ? ld_stream(vol_in + base + (size_t)d * HW) : 0.f;
  const float gt = depth[pix];
  bool first = true;

  for (int c = 0; c < nc; c++) {
    // ---- sense this curtain at this pixel (bit-exact image formation, lc_sense.cuh) ----------
    const int cu = b * nc + c;
    float Pl[12];
#pragma unroll
    for (int k = 0; k < 12; k++) Pl[k] = __ldg(planes + ((size_t)cu * dv.W + u) * 12 + k);
    const Ret r = sense_pixel(dv, Pl, u, v, gt, noise ? noise + (size_t)cu * HW + hw : nullptr);
    const Lik L = lik_params(r, gt, D);
    if (first) {                                             // models.py:1097 (after the loads have landed)
#pragma unroll
      for (int d = 0; d < DPT; d++)
        if (d < D) p[d] = fminf(fmaxf(in_prob ? p[d] : __expf(p[d]), eps_floor()), 1.f);
      first = false;
    }
    // ---- likelihood of this curtain over the depth bins (mixed_model, light_curtain.py:247-252)
    const float kk = -1.4426950408889634f / L.den;           // den = 0 -> -inf: g = 0 (or NaN at dc == z) like the reference
    float g[KEEP_G ? DPT : 1];
    float sg = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) {
        const float gv = gauss_bin(dc[d], L.z, kk);
        if (KEEP_G) g[d] = gv;
        sg += gv;
      }
    // sg == 0 or NaN: g/sg is NaN -> -1 in every bin -> the mixture is constant -> uniform likelihood ->
    // multiply + renormalise leaves the DPV unchanged
    if (!(sg > 0.f)) continue;
    const float As = L.A / sg;
    float sm = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) {
        const float gv = KEEP_G ? g[d] : gauss_bin(dc[d], L.z, kk);
        const float m = fmaxf(fmaf(gv, As, L.Bu), 0.f);      // clamp(soft * A + uni * B, 0, inf)
        if (KEEP_G) g[d] = m;
        sm += m;
      }
    if (!(sm > 0.f)) continue;                               // 0/0 in every bin -> NaN -> 0 -> eps: uniform again
    const float inv_sm = 1.f / sm;
    float s = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) {
        const float m = KEEP_G ? g[d] : fmaxf(fmaf(gauss_bin(dc[d], L.z, kk), As, L.Bu), 0.f);
        p[d] *= fminf(fmaxf(m * inv_sm, eps_floor()), 1.f);   // models.py:1109-1111
        s += p[d];
      }
    if (s < 1e-12f) {                                        // models.py:1112 — only needed to stay in range
      const float inv = 1.f / s;
#pragma unroll
      for (int d = 0; d < DPT; d++)
        if (d < D) p[d] *= inv;
    }
  }
  if (first) {
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) p[d] = fminf(fmaxf(in_prob ? p[d] : __expf(p[d]), eps_floor()), 1.f);
  }

  // ---- box spread along depth x spread_iter (img_utils.py:408-436), one renormalisation at the end
  for (int i = 0; i < spread_iter; i++) {
    switch (spread_n) {
      case 3: spread_sum<DPT, 1>(p, D); break;
      case 5: spread_sum<DPT, 2>(p, D); break;
      case 7: spread_sum<DPT, 3>(p, D); break;
      default: break;
    }
  }
  float total = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) total += p[d];
  const float inv = 1.f / total;

  // ---- clamp, (log), expectation, store (models.py:1119-1122) -------------------------------
  float mean = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) {
      const float cval = fminf(fmaxf(p[d] * inv, eps_floor()), 1.f);
      mean = fmaf(dc[d], cval, mean);
      st_stream(vol_out + base + (size_t)d * HW, out_prob ? cval : __logf(cval));
    }
  if (mean_out) mean_out[pix] = mean;
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
  const unsigned g = (unsigned)((npix + kFuseThreads - 1) / kFuseThreads);
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(DPT, KG)                                                                                       \
  sense_fuse_kernel<DPT, KG><<<g, kFuseThreads, 0, st>>>(dev->view, logp_in, depth, planes, noise, nc, npix,    \
                                                         d_candi, D, spread_iter, spread_n, flags, logp_out, \
                                                         mean_out)
  if (D <= 16) LAUNCH(16, true);
  else if (D <= 32) LAUNCH(32, true);
  else if (D <= 64) LAUNCH(64, true);
  else LAUNCH(128, false);
#undef LAUNCH
  return check_launch("lc_sense_fuse");
}
