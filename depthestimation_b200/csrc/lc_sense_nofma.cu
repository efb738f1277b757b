// (d4) Light-curtain image formation per pixel and the fused sense + Bayesian-update kernel.
//   external/lcsim/lcsim/src/dprocessor.cpp:806-887 (calculateSurface, per-pixel part),
//   :942-993 (computeDepthHits), util.cpp:47-52 (intersect);
//   lc/light_curtain.py:847-888 (sense_high tail), :228-252 (mapping, mixed_model);
//   utils/img_utils.py:176-195 (gen_soft_label_torch); models/models.py:1096-1122 (fusion).
// Compiled with -fmad=false (see lc_device_nofma.cu).
//
// The reference builds, per curtain, a [D,H,W] likelihood volume (50 MB) through ~8 elementwise
// passes and then makes ~12 more passes for the update.  Here a thread owns one pixel: the curtain
// surface, thickness and return intensity are recomputed from 12 plane coefficients per column, the
// likelihood is evaluated bin by bin in registers and multiplied straight into the register-resident
// DPV column; the volume is read once and written once per LC iteration.
#include "dpv_pixel.cuh"
#include "lc_device.cuh"

namespace dpv {

constexpr int kSenseThreads = 128;

struct Ret {          // one pixel of LCDevice.get_return
  float x, y, z, inten;   // NaN x => no return
  float thick;
};

// util.cpp:47-52
__device__ __forceinline__ void intersect(const float* P, float rx, float ry, float rz, float out[3]) {
  float t = -P[3] / (P[0] * rx + P[1] * ry + P[2]);
  out[0] = rx * t;
  out[1] = ry * t;
  out[2] = rz * t;
}

__device__ __forceinline__ Ret sense_pixel(const LcDeviceView& dv, const float* __restrict__ P, int u, int v,
                                           float zval, const float* noise) {
  Ret r;
  const float rx = dv.nn_x[u], ry = dv.nn_y[v];
  float c[3], bot[3], top[3];
  intersect(P, rx, ry, 1.0f, c);
  float c3 = 0.f;
  if (c[2] < 0 || fabs((double)c[1]) > 3.0) { c[0] = c[1] = c[2] = c3 = NAN; }      // :818-823
  intersect(P + 4, rx, ry, 1.0f, bot);
  intersect(P + 8, rx, ry, 1.0f, top);
  double dx = (double)(bot[0] - top[0]), dy = (double)(bot[1] - top[1]), dz = (double)(bot[2] - top[2]);
  const float unc = (float)sqrt(dx * dx + dy * dy + dz * dz);                        // :884-887
  r.thick = unc;
  if (!(c[0] != c[0]) && !(unc != unc) && !(zval != zval)) {                         // computeDepthHits
    float srange = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    float drange = zval / dv.ztoramap[(size_t)v * dv.W + u];
    float inten = 0.f;
    if (dv.hit_mode == 0) {
      float err = fabsf(drange - srange);
      float color = 0;
      if (err >= 0) color = (float)(1. - 1. * (err / (unc / 2.)));
      if (err < unc / 2.) inten = color;
    } else {
      float err = drange - srange;
      float scale = (float)(err / (unc / 2.));
      if (fabsf(err) > unc / 2.) scale = (float)(((0.f < err) - (err < 0.f)) * 1.);
      int idx = (int)(((scale + 1.) / 2.) * (dv.hit_N - 1));
      idx = min(max(idx, 0), dv.hit_N - 1);
      inten = dv.hit_lut[idx];
    }
    if (noise) {
      inten += *noise;
      inten = fmaxf(0.f, fminf(inten, 1.f));
    }
    c3 = (float)(int)(255 * inten);
  }
  r.x = c[0]; r.y = c[1]; r.z = c[2]; r.inten = c3;
  return r;
}

// Per-pixel, per-curtain parameters of the likelihood (sense_high tail + mixed_model)
struct Lik {
  float z;      // sensed z * mask
  float den;    // 2 * sigma^2
  float A, Bu;  // mixture weight, (1-A) * (1/D)
};

__device__ __forceinline__ Lik lik_params(const Ret& r, float gt, int D) {
  float z = r.z, th = r.thick, it = r.inten;
  if (r.x != r.x) { z = 0.f; it = 0.f; }            // light_curtain.py:847 (whole pixel -> 0)
  if (th != th) th = 0.f;                           // :848 (only where the thickness itself is NaN)
  const float mask = gt > 0.f ? 1.f : 0.f;          // :870
  z = z * mask;
  th = th * mask;
  it = it * mask;
  const float x = it / 255.0f;
  // mapping(), light_curtain.py:228-244 (m = 20, f = 0.5)
  float A;
  if (x > 0.5f) {
    const float cc = (float)(20.0 / ((20.0 * 0.5 + 0.5) * (20.0 * 0.5 + 0.5)));
    const float c1 = (float)(1.0 - 20.0 / ((20.0 * 0.5 + 0.5) * (20.0 * 0.5 + 0.5)));
    A = cc * x + c1;
  } else {
    A = (1.0f / (20.f * (0.025f + x))) * -1.0f + 1.0f;
  }
  Lik L;
  L.z = z;
  const float t10 = th / 10.0f;
  const float var = t10 * t10;
  const float sg = sqrtf(var);
  L.den = 2.f * (sg * sg);
  L.A = A;
  L.Bu = (1.0f / (float)D) * (1.f - A);
  return L;
}

// exp(-|dc - z|^2 / (2 sigma^2)); exact 0 once the argument is below fp32 exp underflow
__device__ __forceinline__ float soft_g(float dc, const Lik& L) {
  float a = fabsf(dc - L.z);
  a = a * a;
  if (a > 105.f * L.den) return 0.f;      // (-a)/den < -104.9: expf rounds to +0
  return expf((-a) / L.den);
}

__device__ __forceinline__ float clamp0_keepnan(float v) { return (v != v) ? v : fmaxf(v, 0.f); }

// mix[d] before the final normalisation: clamp(g/sg * A + (1/D)*(1-A), 0, inf) with NaN -> -1 first
__device__ __forceinline__ float mix_value(float g, float sg, const Lik& L) {
  float n = g / sg;
  if (n != n) n = -1.f;                   // zero_invalid (img_utils.py:189-190)
  return clamp0_keepnan(n * L.A + L.Bu);
}

template <int DPT>
__device__ __forceinline__ void lik_sums(const float* dc, int D, const Lik& L, float& sg, float& sm) {
  sg = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) sg += soft_g(dc[d], L);
  sm = 0.f;
  if (sg > 0.f) {
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) sm += mix_value(soft_g(dc[d], L), sg, L);
  } else {
    // every bin is 0/0 -> -1 (or NaN sum): the mixture is the same constant in every bin
    float m = mix_value(0.f, sg, L);
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) sm += m;
  }
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSenseThreads)
sense_image_kernel(LcDeviceView dv, const float* __restrict__ planes, const float* __restrict__ depth,
                   const float* __restrict__ noise, int nc, long long npix, float* __restrict__ image,
                   float* __restrict__ thick) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const size_t HW = (size_t)dv.H * dv.W;
  const int cu = (int)(pix / (long long)HW);          // curtain index = b*nc + c
  const size_t hw = (size_t)(pix - (long long)cu * HW);
  const int v = (int)(hw / dv.W), u = (int)(hw % dv.W);
  const int b = cu / nc;
  const float* P = planes + ((size_t)cu * dv.W + u) * 12;
  float Pl[12];
#pragma unroll
  for (int k = 0; k < 12; k++) Pl[k] = __ldg(P + k);
  Ret r = sense_pixel(dv, Pl, u, v, depth[(size_t)b * HW + hw], noise ? noise + pix : nullptr);
  float4 o = make_float4(r.x, r.y, r.z, r.inten);
  reinterpret_cast<float4*>(image)[pix] = o;
  thick[pix] = r.thick;
}

template <int DPT>
__global__ void __launch_bounds__(kSenseThreads)
sense_likelihood_kernel(LcDeviceView dv, const float* __restrict__ planes, const float* __restrict__ depth,
                        const float* __restrict__ noise, int nc, long long npix, const float* __restrict__ d_candi,
                        int D, float* __restrict__ lik) {
  __shared__ float dc[DPT];
  for (int i = threadIdx.x; i < DPT; i += blockDim.x) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const size_t HW = (size_t)dv.H * dv.W;
  const int cu = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)cu * HW);
  const int v = (int)(hw / dv.W), u = (int)(hw % dv.W);
  const int b = cu / nc;
  float Pl[12];
#pragma unroll
  for (int k = 0; k < 12; k++) Pl[k] = __ldg(planes + ((size_t)cu * dv.W + u) * 12 + k);
  const float gt = depth[(size_t)b * HW + hw];
  Ret r = sense_pixel(dv, Pl, u, v, gt, noise ? noise + pix : nullptr);
  Lik L = lik_params(r, gt, D);
  float sg, sm;
  lik_sums<DPT>(dc, D, L, sg, sm);
  float* o = lik + (size_t)cu * D * HW + hw;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) st_stream(o + (size_t)d * HW, mix_value(soft_g(dc[d], L), sg, L) / sm);
}

template <int DPT>
__global__ void __launch_bounds__(kSenseThreads)
sense_fuse_kernel(LcDeviceView dv, const float* __restrict__ logp_in, const float* __restrict__ depth,
                  const float* __restrict__ planes, const float* __restrict__ noise, int nc, long long npix,
                  const float* __restrict__ d_candi, int D, int spread_iter, int spread_n,
                  float* __restrict__ logp_out, float* __restrict__ mean_out) {
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
  float p[DPT];
  load_prob_from_log<DPT>(p, logp_in + base, HW, D);             // models.py:1097
  const float gt = depth[pix];
  for (int c = 0; c < nc; c++) {
    const int cu = b * nc + c;
    float Pl[12];
#pragma unroll
    for (int k = 0; k < 12; k++) Pl[k] = __ldg(planes + ((size_t)cu * dv.W + u) * 12 + k);
    Ret r = sense_pixel(dv, Pl, u, v, gt, noise ? noise + (size_t)cu * HW + hw : nullptr);
    Lik L = lik_params(r, gt, D);
    float sg, sm;
    lik_sums<DPT>(dc, D, L, sg, sm);
    float s = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) {
        float lk = mix_value(sg > 0.f ? soft_g(dc[d], L) : 0.f, sg, L) / sm;
        if (lk != lk) lk = 0.f;                                  // models.py:1109
        lk = fminf(fmaxf(lk, eps_floor()), 1.f);                 // :1110
        p[d] *= lk;                                              // :1111
        s += p[d];
      }
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) p[d] = p[d] / s;                                // :1112
  }
  for (int i = 0; i < spread_iter; i++) spread_dispatch<DPT>(p, D, spread_n);     // :1115-1116
  float mean = store_log_and_mean<DPT>(p, logp_out + base, HW, D, dc);            // :1119-1122
  if (mean_out) mean_out[pix] = mean;
}

}  // namespace dpv

using namespace dpv;

extern "C" {

int lc_sense_image(const LcDevice* dev, const float* planes, const float* depth, const float* noise, int B, int nc,
                   float* image, float* thick, void* stream) {
  DPV_REQUIRE(dev && planes && depth && image && thick, "lc_sense_image: null pointer");
  DPV_REQUIRE(B > 0 && nc > 0, "lc_sense_image: bad shape");
  const long long npix = (long long)B * nc * dev->H * dev->W;
  sense_image_kernel<<<(unsigned)((npix + kSenseThreads - 1) / kSenseThreads), kSenseThreads, 0,
                       (cudaStream_t)stream>>>(dev->view, planes, depth, noise, nc, npix, image, thick);
  return check_launch("lc_sense_image");
}

int lc_sense_likelihood(const LcDevice* dev, const float* planes, const float* depth, const float* noise, int B,
                        int nc, const float* d_candi, int D, float* lik, void* stream) {
  DPV_REQUIRE(dev && planes && depth && d_candi && lik, "lc_sense_likelihood: null pointer");
  DPV_REQUIRE(B > 0 && nc > 0 && D > 0, "lc_sense_likelihood: bad shape");
  DPV_REQUIRE(D <= 128, "lc_sense_likelihood: D=%d > 128 unsupported", D);
  const long long npix = (long long)B * nc * dev->H * dev->W;
  const unsigned g = (unsigned)((npix + kSenseThreads - 1) / kSenseThreads);
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(DPT) \
  sense_likelihood_kernel<DPT><<<g, kSenseThreads, 0, st>>>(dev->view, planes, depth, noise, nc, npix, d_candi, D, lik)
  if (D <= 16) LAUNCH(16);
  else if (D <= 32) LAUNCH(32);
  else if (D <= 64) LAUNCH(64);
  else LAUNCH(128);
#undef LAUNCH
  return check_launch("lc_sense_likelihood");
}

int lc_sense_fuse(const LcDevice* dev, const float* logp_in, const float* depth, const float* planes,
                  const float* noise, int B, int nc, const float* d_candi, int D, int spread_iter, int spread_n,
                  float* logp_out, float* mean_out, void* stream) {
  DPV_REQUIRE(dev && logp_in && depth && planes && d_candi && logp_out, "lc_sense_fuse: null pointer");
  DPV_REQUIRE(B > 0 && nc >= 0 && D > 0 && spread_iter >= 0, "lc_sense_fuse: bad shape");
  DPV_REQUIRE(spread_iter == 0 || spread_n == 1 || spread_n == 3 || spread_n == 5 || spread_n == 7,
              "lc_sense_fuse: spread_n=%d not in {1,3,5,7}", spread_n);
  DPV_REQUIRE(D <= 128, "lc_sense_fuse: D=%d > 128 unsupported", D);
  const long long npix = (long long)B * dev->H * dev->W;
  const unsigned g = (unsigned)((npix + kSenseThreads - 1) / kSenseThreads);
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(DPT)                                                                                              \
  sense_fuse_kernel<DPT><<<g, kSenseThreads, 0, st>>>(dev->view, logp_in, depth, planes, noise, nc, npix, d_candi, \
                                                      D, spread_iter, spread_n, logp_out, mean_out)
  if (D <= 16) LAUNCH(16);
  else if (D <= 32) LAUNCH(32);
  else if (D <= 64) LAUNCH(64);
  else LAUNCH(128);
#undef LAUNCH
  return check_launch("lc_sense_fuse");
}

}  // extern "C"
