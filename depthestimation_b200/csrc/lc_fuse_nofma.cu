// (d4) One LC iteration's sensing + Bayesian update in a single pass over the DPV.
//   models/models.py:1064-1122 (sense each curtain, clamp, multiply, renormalise, spread, clamp, log)
//   with lc/light_curtain.py:847-888 + :228-252 (likelihood) and dprocessor.cpp:806-993 (image formation)
//   evaluated per pixel in registers.  Compiled with -fmad=false for the geometry (see lc_sense.cuh).
//
// Work per depth bin is kept to the streaming minimum by using two facts about the reference's update:
//  (1) the mixture likelihood of a curtain is the SAME constant in every bin except the few bins within
//      ~14.5 sigma of the sensed surface (exp underflows to exactly 0 elsewhere), and a constant factor
//      cancels in the renormalisation that follows, so only those "window" bins are touched;
//  (2) multiply, box-spread and renormalise are all homogeneous in the DPV, so the per-curtain and
//      per-spread renormalisations collapse into one scale at the end (identical up to fp32 rounding;
//      an extra renormalisation is inserted only where a near-zero background factor could overflow).
// The volume may be kept as log-probability (the reference's format) or as clamped probability between
// the iterations of one lc_process call (flags), which removes exp/log from the inner iterations.
#include "lc_sense.cuh"

namespace dpv {

constexpr int kFuseThreads = 128;
constexpr int kMaxCurtains = 8;

struct Win {              // one curtain's non-constant likelihood bins at this pixel
  int lo, hi;             // candidate bin range (inclusive); lo > hi = likelihood is uniform
  float z, den, A, Bu;    // Lik
  float sg, inv_sm, l0;   // sum of g, 1/sum of mix, clamped background likelihood
};

__device__ __forceinline__ float window_ratio(const Win& w, float dcv) {
  Lik L;
  L.z = w.z; L.den = w.den; L.A = w.A; L.Bu = w.Bu;
  float lk = mix_value(soft_g(dcv, L), w.sg, L) * w.inv_sm;
  if (lk != lk) lk = 0.f;                                    // models.py:1109
  lk = fminf(fmaxf(lk, eps_floor()), 1.f);                   // :1110
  return lk / w.l0;
}

template <int DPT>
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

  // ---- issue the volume loads first; the per-curtain geometry below overlaps their latency
  float p[DPT];
#pragma unroll
  for (int d = 0; d < DPT; d++) p[d] = d < D ? ld_stream(vol_in + base + (size_t)d * HW) : 0.f;

  // ---- per-curtain sensing: window of non-constant likelihood bins ---------------------------
  const float gt = depth[pix];
  Win win[kMaxCurtains];
  int nwin = 0;
  bool risky = false;
  for (int c = 0; c < nc; c++) {
    const int cu = b * nc + c;
    float Pl[12];
#pragma unroll
    for (int k = 0; k < 12; k++) Pl[k] = __ldg(planes + ((size_t)cu * dv.W + u) * 12 + k);
    Ret r = sense_pixel(dv, Pl, u, v, gt, noise ? noise + (size_t)cu * HW + hw : nullptr);
    Lik L = lik_params(r, gt, D);
    // bins with exp(-(dc-z)^2/den) != 0 satisfy (dc-z)^2 <= 105*den: candidate range by bisection on dc[]
    const float rad = sqrtf(105.f * L.den) * 1.0001f + 1e-6f;
    if (!(rad == rad) || !(L.z == L.z)) {
      // NaN parameters: fall back to treating every bin as a candidate (the exact path handles NaN)
      Win& w = win[nwin];
      w.lo = 0; w.hi = D - 1;
      w.z = L.z; w.den = L.den; w.A = L.A; w.Bu = L.Bu;
    } else {
      int lo = 0, hi = D;                                    // first k with dc[k] >= z - rad
      const float a0 = L.z - rad, a1 = L.z + rad;
      while (lo < hi) { int m = (lo + hi) >> 1; if (dc[m] < a0) lo = m + 1; else hi = m; }
      int lo2 = lo, hi2 = D;                                 // first k with dc[k] > z + rad
      while (lo2 < hi2) { int m = (lo2 + hi2) >> 1; if (dc[m] <= a1) lo2 = m + 1; else hi2 = m; }
      Win& w = win[nwin];
      w.lo = lo; w.hi = lo2 - 1;
      w.z = L.z; w.den = L.den; w.A = L.A; w.Bu = L.Bu;
    }
    Win& w = win[nwin];
    if (w.lo > w.hi) continue;                               // every bin has g = 0 -> uniform likelihood
    float sg = 0.f;
    for (int k = w.lo; k <= w.hi; k++) sg += soft_g(dc[k], L);
    if (!(sg > 0.f)) continue;                               // 0/0 -> -1 in every bin -> uniform
    const float m0 = mix_value(0.f, sg, L);                  // bins with g = 0
    float sm = 0.f;
    for (int k = w.lo; k <= w.hi; k++) sm += mix_value(soft_g(dc[k], L), sg, L);
    sm += (float)(D - (w.hi - w.lo + 1)) * m0;
    float l0 = m0 / sm;
    if (l0 != l0) l0 = 0.f;
    l0 = fminf(fmaxf(l0, eps_floor()), 1.f);
    if (!(sm > 0.f)) continue;                               // all-zero / NaN mixture -> every bin clamps to eps
    w.sg = sg; w.inv_sm = 1.f / sm; w.l0 = l0;
    risky |= (l0 < 1e-6f);
    nwin++;
  }

  // ---- DPV column to clamped probability (models.py:1097) -----------------------------------
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) p[d] = fminf(fmaxf(in_prob ? p[d] : __expf(p[d]), eps_floor()), 1.f);

  // ---- multiply the window bins (models.py:1108-1112; background factor cancels on renormalisation)
  for (int c = 0; c < nwin; c++) {
    const Win w = win[c];
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D && d >= w.lo && d <= w.hi) p[d] *= window_ratio(w, dc[d]);
    if (risky) {                                             // keep the magnitudes bounded (rare)
      float s = 0.f;
#pragma unroll
      for (int d = 0; d < DPT; d++)
        if (d < D) s += p[d];
      const float inv = 1.f / s;
#pragma unroll
      for (int d = 0; d < DPT; d++)
        if (d < D) p[d] *= inv;
    }
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
#define LAUNCH(DPT)                                                                                             \
  sense_fuse_kernel<DPT><<<g, kFuseThreads, 0, st>>>(dev->view, logp_in, depth, planes, noise, nc, npix, d_candi, \
                                                     D, spread_iter, spread_n, flags, logp_out, mean_out)
  if (D <= 16) LAUNCH(16);
  else if (D <= 32) LAUNCH(32);
  else if (D <= 64) LAUNCH(64);
  else LAUNCH(128);
#undef LAUNCH
  return check_launch("lc_sense_fuse");
}
