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
#include "lc_sense.cuh"

namespace dpv {

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

// The tail of sense_high / sense_low (lc/light_curtain.py:869-888) on maps that have already been brought into the
// RGB camera by transformPoints (the rTc != I branch, :850-858): maps [n,3,H,W] = (depth, intensity, thickness).
template <int DPT>
__global__ void __launch_bounds__(kSenseThreads)
likelihood_maps_kernel(const float* __restrict__ maps, const float* __restrict__ depth, int per_depth, long long npix,
                       size_t HW, const float* __restrict__ d_candi, int D, float* __restrict__ lik) {
  __shared__ float dc[DPT];
  for (int i = threadIdx.x; i < DPT; i += blockDim.x) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const size_t cu = (size_t)(pix / (long long)HW), hw = (size_t)(pix - (long long)cu * HW);
  const float* m = maps + cu * 3 * HW + hw;
  Ret r;
  r.x = r.y = 0.f;                                    // not NaN: the maps carry 0 where nothing was sensed
  r.z = m[0];
  r.inten = m[HW];
  r.thick = m[2 * HW];
  const float gt = depth[(cu / per_depth) * HW + hw];
  Lik L = lik_params(r, gt, D);
  float sg, sm;
  lik_sums<DPT>(dc, D, L, sg, sm);
  float* o = lik + cu * D * HW + hw;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) st_stream(o + (size_t)d * HW, mix_value(soft_g(dc[d], L), sg, L) / sm);
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

int lc_likelihood_maps(const float* maps, const float* depth, int B, int nc, int H, int W, const float* d_candi, int D,
                       float* lik, void* stream) {
  DPV_REQUIRE(maps && depth && d_candi && lik, "lc_likelihood_maps: null pointer");
  DPV_REQUIRE(B > 0 && nc > 0 && H > 0 && W > 0 && D > 0, "lc_likelihood_maps: bad shape");
  DPV_REQUIRE(D <= 128, "lc_likelihood_maps: D=%d > 128 unsupported", D);
  const size_t HW = (size_t)H * W;
  const long long npix = (long long)B * nc * (long long)HW;
  const unsigned g = (unsigned)((npix + kSenseThreads - 1) / kSenseThreads);
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(DPT) likelihood_maps_kernel<DPT><<<g, kSenseThreads, 0, st>>>(maps, depth, nc, npix, HW, d_candi, D, lik)
  if (D <= 16) LAUNCH(16);
  else if (D <= 32) LAUNCH(32);
  else if (D <= 64) LAUNCH(64);
  else LAUNCH(128);
#undef LAUNCH
  return check_launch("lc_likelihood_maps");
}

}  // extern "C"
