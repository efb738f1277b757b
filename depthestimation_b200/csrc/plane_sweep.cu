// (a) Fused homography plane-sweep warp + feature cost, and the diagonal feature warp.
//   warping/homography.py:98-135 est_swp_volume_v4, :137-168 warp_feature, :170-198 _back_warp_homo_parallel.
//
// The reference materialises src.repeat(D) ([D,C,h,w], 210 MB at C2), grid_samples it and reduces
// (warp - ref)^2 over C.  Here one thread owns one reference pixel and DC consecutive depth planes:
// the homography coordinates, the 4 bilinear tap offsets and weights of each plane stay in registers,
// the channel loop streams ref once and gathers src through L1/L2 (the 3.3 MB view is L2-resident;
// lanes of a warp are 32 consecutive x at the same plane, so their taps fall on 1-2 lines per row),
// and the only HBM write is the [D,h,w] cost.
#include <cstdlib>

#include "common.cuh"

namespace dpv {

// element strides between the batch items of ref / src / R / t: the inputs may be views into larger tensors
// (models/models.py:607-617 takes the reference and source views out of one [B,T,C,h,w] feature tensor)
struct SweepStrides {
  long long ref_b, src_b, R_b, t_b;
};

struct Taps {
  int o00, o01, o10, o11;
  float w00, w01, w10, w11;
};

// P = K t + (K R ray) d ; u = P.x / (P.z + 1e-10) ; g = (u - c) / c ; bilinear taps with zeros
// padding under align_corners=False (ATen grid_sampler_2d: nw/ne/sw/se weights from ix_se-ix etc.).
__device__ __forceinline__ Taps make_taps(const float kt[3], const float krp[3], float d, float cu, float cv, int h,
                                          int w) {
  float px = __fadd_rn(kt[0], __fmul_rn(krp[0], d));
  float py = __fadd_rn(kt[1], __fmul_rn(krp[1], d));
  float pz = __fadd_rn(kt[2], __fmul_rn(krp[2], d));
  float z = __fadd_rn(pz, 1e-10f);
  float u = __fdiv_rn(px, z), v = __fdiv_rn(py, z);
  float gx = __fdiv_rn(__fsub_rn(u, cu), cu);
  float gy = __fdiv_rn(__fsub_rn(v, cv), cv);
  float ix = gs_unnormalize(gx, w), iy = gs_unnormalize(gy, h);
  float fx0 = floorf(ix), fy0 = floorf(iy);
  float tx = __fsub_rn(ix, fx0), ty = __fsub_rn(iy, fy0);       // ix - ix_nw
  float ux = __fsub_rn(__fadd_rn(fx0, 1.f), ix), uy = __fsub_rn(__fadd_rn(fy0, 1.f), iy);   // ix_se - ix
  Taps T;
  T.w00 = __fmul_rn(ux, uy);   // nw
  T.w01 = __fmul_rn(tx, uy);   // ne
  T.w10 = __fmul_rn(ux, ty);   // sw
  T.w11 = __fmul_rn(tx, ty);   // se
  // NaN / far-out coordinates: every tap invalid (grid_sample yields 0)
  bool finite = (ix > -2.f) && (ix < (float)(w + 1)) && (iy > -2.f) && (iy < (float)(h + 1));
  int x0 = finite ? (int)fx0 : -2, y0 = finite ? (int)fy0 : -2;
  int x1 = x0 + 1, y1 = y0 + 1;
  bool vx0 = x0 >= 0 && x0 < w, vx1 = x1 >= 0 && x1 < w, vy0 = y0 >= 0 && y0 < h, vy1 = y1 >= 0 && y1 < h;
  if (!(vx0 && vy0)) T.w00 = 0.f;
  if (!(vx1 && vy0)) T.w01 = 0.f;
  if (!(vx0 && vy1)) T.w10 = 0.f;
  if (!(vx1 && vy1)) T.w11 = 0.f;
  int cx0 = min(max(x0, 0), w - 1), cx1 = min(max(x1, 0), w - 1);
  int cy0 = min(max(y0, 0), h - 1), cy1 = min(max(y1, 0), h - 1);
  T.o00 = cy0 * w + cx0;
  T.o01 = cy0 * w + cx1;
  T.o10 = cy1 * w + cx0;
  T.o11 = cy1 * w + cx1;
  return T;
}

// The same taps as a 2 x 2 block anchored at (bx, by) with 0 <= bx <= w-2, 0 <= by <= h-2 (needs w, h >= 2): a tap that
// falls outside the image has weight 0 (zeros padding), so when the block's natural anchor (x0, y0) is -1 or
// w-1 / h-1 the anchor moves one pixel inside and the surviving weights move to the other column / row.  Every tap
// then reads anchor + {0,1} + {0,1} * w: two offset registers per plane and immediate offsets instead of four
// offsets (the plane-sweep kernel keeps DC planes' taps in registers: 6 instead of 8 registers per plane).
struct TapsB {
  int o0, o1;                    // offsets of the block's two rows
  float w00, w01, w10, w11;
};

__device__ __forceinline__ TapsB make_taps_block(const float kt[3], const float krp[3], float d, float cu, float cv,
                                                 int h, int w) {
  float px = __fadd_rn(kt[0], __fmul_rn(krp[0], d));
  float py = __fadd_rn(kt[1], __fmul_rn(krp[1], d));
  float pz = __fadd_rn(kt[2], __fmul_rn(krp[2], d));
  float z = __fadd_rn(pz, 1e-10f);
  float u = __fdiv_rn(px, z), v = __fdiv_rn(py, z);
  float gx = __fdiv_rn(__fsub_rn(u, cu), cu);
  float gy = __fdiv_rn(__fsub_rn(v, cv), cv);
  float ix = gs_unnormalize(gx, w), iy = gs_unnormalize(gy, h);
  float fx0 = floorf(ix), fy0 = floorf(iy);
  float tx = __fsub_rn(ix, fx0), ty = __fsub_rn(iy, fy0);
  float ux = __fsub_rn(__fadd_rn(fx0, 1.f), ix), uy = __fsub_rn(__fadd_rn(fy0, 1.f), iy);
  TapsB T;
  T.w00 = __fmul_rn(ux, uy);
  T.w01 = __fmul_rn(tx, uy);
  T.w10 = __fmul_rn(ux, ty);
  T.w11 = __fmul_rn(tx, ty);
  bool finite = (ix > -2.f) && (ix < (float)(w + 1)) && (iy > -2.f) && (iy < (float)(h + 1));
  int x0 = finite ? (int)fx0 : -2, y0 = finite ? (int)fy0 : -2;
  const bool vx0 = x0 >= 0 && x0 < w, vx1 = x0 + 1 >= 0 && x0 + 1 < w;
  const bool vy0 = y0 >= 0 && y0 < h, vy1 = y0 + 1 >= 0 && y0 + 1 < h;
  if (!(vx0 && vy0)) T.w00 = 0.f;
  if (!(vx1 && vy0)) T.w01 = 0.f;
  if (!(vx0 && vy1)) T.w10 = 0.f;
  if (!(vx1 && vy1)) T.w11 = 0.f;
  const int bx = min(max(x0, 0), w - 2), by = min(max(y0, 0), h - 2);
  if (x0 == bx + 1) { T.w01 = T.w00; T.w11 = T.w10; T.w00 = 0.f; T.w10 = 0.f; }       // x0 = w-1: only column x0 lives
  else if (x0 == bx - 1) { T.w00 = T.w01; T.w10 = T.w11; T.w01 = 0.f; T.w11 = 0.f; }  // x0 = -1: only column 0 lives
  if (y0 == by + 1) { T.w10 = T.w00; T.w11 = T.w01; T.w00 = 0.f; T.w01 = 0.f; }
  else if (y0 == by - 1) { T.w00 = T.w10; T.w01 = T.w11; T.w10 = 0.f; T.w11 = 0.f; }
  T.o0 = by * w + bx;
  T.o1 = T.o0 + w;
  return T;
}

__device__ __forceinline__ float tap_sample(const float* __restrict__ plane, const TapsB& T) {
  float a = __ldg(plane + T.o0) * T.w00;
  a += __ldg(plane + T.o0 + 1) * T.w01;
  a += __ldg(plane + T.o1) * T.w10;
  a += __ldg(plane + T.o1 + 1) * T.w11;
  return a;
}

__device__ __forceinline__ float tap_sample(const float* __restrict__ plane, const Taps& T) {
  // ATen order: nw*w + ne*w + sw*w + se*w
  float a = __ldg(plane + T.o00) * T.w00;
  a += __ldg(plane + T.o01) * T.w01;
  a += __ldg(plane + T.o10) * T.w10;
  a += __ldg(plane + T.o11) * T.w11;
  return a;
}

// K[3x3] R[3x3] t[3] -> M = K R (homography.py:121), kt = K t (:119); row-major, k = 0..2 in order.
__device__ __forceinline__ void pose_terms(const float* __restrict__ K, const float* __restrict__ R,
                                           const float* __restrict__ t, float M[9], float kt[3]) {
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      float s = K[i * 3 + 0] * R[0 * 3 + j];
      s = fmaf(K[i * 3 + 1], R[1 * 3 + j], s);
      s = fmaf(K[i * 3 + 2], R[2 * 3 + j], s);
      M[i * 3 + j] = s;
    }
    float s = K[i * 3 + 0] * t[0];
    s = fmaf(K[i * 3 + 1], t[1], s);
    s = fmaf(K[i * 3 + 2], t[2], s);
    kt[i] = s;
  }
}

template <int DC, int METRIC>
__global__ void __launch_bounds__(128)
plane_sweep_kernel(const float* __restrict__ ref, const float* __restrict__ src, int V, int C, int h, int w,
                   const float* __restrict__ K, const float* __restrict__ R, const float* __restrict__ t,
                   const float* __restrict__ rays, long long rays_bstride, const float* __restrict__ center,
                   const float* __restrict__ d_candi, int D, float sigma, float* __restrict__ cost, SweepStrides sd) {
  const int hw = h * w;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  const int d0 = blockIdx.y * DC;
  const int b = blockIdx.z;
  if (pix >= hw) return;
  const float cu = center[b * 2 + 0], cv = center[b * 2 + 1];
  const float* ray = rays + (size_t)b * rays_bstride;
  const float r0 = ray[pix], r1 = ray[hw + pix], r2 = ray[2 * hw + pix];
  float acc[DC];
#pragma unroll
  for (int j = 0; j < DC; j++) acc[j] = 0.f;
  const float* refp = ref + (size_t)b * sd.ref_b + pix;
  for (int v = 0; v < V; v++) {
    float M[9], kt[3], krp[3];
    pose_terms(K + b * 9, R + (size_t)b * sd.R_b + v * 9, t + (size_t)b * sd.t_b + v * 3, M, kt);
#pragma unroll
    for (int i = 0; i < 3; i++) krp[i] = fmaf(M[i * 3 + 2], r2, fmaf(M[i * 3 + 1], r1, M[i * 3 + 0] * r0));
    Taps T[DC];
#pragma unroll
    for (int j = 0; j < DC; j++) T[j] = make_taps(kt, krp, d_candi[min(d0 + j, D - 1)], cu, cv, h, w);
    const float* sp = src + (size_t)b * sd.src_b + (size_t)v * C * hw;
    float dist[DC];
#pragma unroll
    for (int j = 0; j < DC; j++) dist[j] = 0.f;
#pragma unroll 2
    for (int c = 0; c < C; c++) {
      const float r = __ldg(refp + (size_t)c * hw);
      const float* plane = sp + (size_t)c * hw;
#pragma unroll
      for (int j = 0; j < DC; j++) {
        float diff = tap_sample(plane, T[j]) - r;
        dist[j] += METRIC == 0 ? diff * diff : fabsf(diff);
      }
    }
#pragma unroll
    for (int j = 0; j < DC; j++) acc[j] += dist[j] / sigma;       // homography.py:129
  }
#pragma unroll
  for (int j = 0; j < DC; j++)
    if (d0 + j < D) cost[((size_t)b * D + d0 + j) * hw + pix] = acc[j];
}

#ifndef DPV_SWEEP_MINB
#define DPV_SWEEP_MINB 8
#endif

// plane_sweep_kernel on 2 x 2 tap blocks (images of at least 2 x 2 pixels): fewer registers, more resident warps.
// The gathers go through L1 at its wavefront rate (32 consecutive x read ~32 consecutive words of a source row that
// straddle two 128-byte lines), so what helps is resident warps, not work per thread: 4 planes per thread at 64
// registers (32 warps per SM) measured 0.517 ms at C2 against 0.598 ms for 8 planes at 252 registers; staging the
// source footprint of a 32 x 4 tile in shared memory instead was slower (0.76 ms: the copy and the same number of
// wavefronts, now on the shared-memory side of the same pipe).
template <int DC, int METRIC>
__global__ void __launch_bounds__(128, DPV_SWEEP_MINB)
plane_sweep_block_kernel(const float* __restrict__ ref, const float* __restrict__ src, int V, int C, int h, int w,
                   const float* __restrict__ K, const float* __restrict__ R, const float* __restrict__ t,
                   const float* __restrict__ rays, long long rays_bstride, const float* __restrict__ center,
                   const float* __restrict__ d_candi, int D, float sigma, float* __restrict__ cost, SweepStrides sd) {
  const int hw = h * w;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  const int d0 = blockIdx.y * DC;
  const int b = blockIdx.z;
  if (pix >= hw) return;
  const float cu = center[b * 2 + 0], cv = center[b * 2 + 1];
  const float* ray = rays + (size_t)b * rays_bstride;
  const float r0 = ray[pix], r1 = ray[hw + pix], r2 = ray[2 * hw + pix];
  float acc[DC];
#pragma unroll
  for (int j = 0; j < DC; j++) acc[j] = 0.f;
  const float* refp = ref + (size_t)b * sd.ref_b + pix;
  for (int v = 0; v < V; v++) {
    float M[9], kt[3], krp[3];
    pose_terms(K + b * 9, R + (size_t)b * sd.R_b + v * 9, t + (size_t)b * sd.t_b + v * 3, M, kt);
#pragma unroll
    for (int i = 0; i < 3; i++) krp[i] = fmaf(M[i * 3 + 2], r2, fmaf(M[i * 3 + 1], r1, M[i * 3 + 0] * r0));
    TapsB T[DC];
#pragma unroll
    for (int j = 0; j < DC; j++) T[j] = make_taps_block(kt, krp, d_candi[min(d0 + j, D - 1)], cu, cv, h, w);
    const float* sp = src + (size_t)b * sd.src_b + (size_t)v * C * hw;
    float dist[DC];
#pragma unroll
    for (int j = 0; j < DC; j++) dist[j] = 0.f;
#pragma unroll 2
    for (int c = 0; c < C; c++) {
      const float r = __ldg(refp + (size_t)c * hw);
      const float* plane = sp + (size_t)c * hw;
#pragma unroll
      for (int j = 0; j < DC; j++) {
        float diff = tap_sample(plane, T[j]) - r;
        dist[j] += METRIC == 0 ? diff * diff : fabsf(diff);
      }
    }
#pragma unroll
    for (int j = 0; j < DC; j++) acc[j] += dist[j] / sigma;       // homography.py:129
  }
#pragma unroll
  for (int j = 0; j < DC; j++)
    if (d0 + j < D) cost[((size_t)b * D + d0 + j) * hw + pix] = acc[j];
}

// out[b,v,i,y,x] = bilinear(src[b,v,i], H_{d_i} p): thread = (pixel, plane i, view v, frame b).
__global__ void __launch_bounds__(128)
warp_feature_kernel(const float* __restrict__ src, int T_, int C, int h, int w, const float* __restrict__ K,
                    const float* __restrict__ R, const float* __restrict__ t, const float* __restrict__ rays,
                    long long rays_bstride, const float* __restrict__ center, const float* __restrict__ d_candi,
                    int D, float* __restrict__ out) {
  const int hw = h * w;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  const int bv = blockIdx.z;
  const int b = bv / T_;
  if (pix >= hw) return;
  const float* ray = rays + (size_t)b * rays_bstride;
  float M[9], kt[3], krp[3];
  pose_terms(K + b * 9, R + (size_t)bv * 9, t + (size_t)bv * 3, M, kt);
  const float r0 = ray[pix], r1 = ray[hw + pix], r2 = ray[2 * hw + pix];
#pragma unroll
  for (int k = 0; k < 3; k++) krp[k] = fmaf(M[k * 3 + 2], r2, fmaf(M[k * 3 + 1], r1, M[k * 3 + 0] * r0));
  Taps tp = make_taps(kt, krp, d_candi[i], center[b * 2 + 0], center[b * 2 + 1], h, w);
  out[((size_t)bv * D + i) * hw + pix] = tap_sample(src + ((size_t)bv * C + i) * hw, tp);
}

}  // namespace dpv

using namespace dpv;

extern "C" {

int dpv_plane_sweep_strided(const float* ref, long long ref_batch_stride, const float* src, long long src_batch_stride,
                            int B, int V, int C, int h, int w, const float* K, const float* R, long long R_batch_stride,
                            const float* t, long long t_batch_stride, const float* rays, long long rays_batch_stride,
                            const float* center, const float* d_candi, int D, float sigma, int metric, float* cost,
                            void* stream) {
  DPV_REQUIRE(ref && src && K && R && t && rays && center && d_candi && cost, "dpv_plane_sweep: null pointer");
  DPV_REQUIRE(B > 0 && V > 0 && C > 0 && h > 0 && w > 0 && D > 0, "dpv_plane_sweep: bad shape");
  DPV_REQUIRE(B <= 65535, "dpv_plane_sweep: B=%d > 65535", B);
  DPV_REQUIRE(ref_batch_stride >= (long long)C * h * w && src_batch_stride >= (long long)V * C * h * w &&
              R_batch_stride >= 9ll * V && t_batch_stride >= 3ll * V, "dpv_plane_sweep: batch strides overlap the items");
  const SweepStrides sd = {ref_batch_stride, src_batch_stride, R_batch_stride, t_batch_stride};
  // homography.py:84-86: raise Exception('undefined metric for feature distance ...')
  DPV_REQUIRE(metric == 0 || metric == 1, "undefined metric for feature distance ...");
  cudaStream_t st = (cudaStream_t)stream;
#ifndef DPV_SWEEP_DC
#define DPV_SWEEP_DC 4
#endif
  constexpr int DC = DPV_SWEEP_DC;
  dim3 grid((h * w + 127) / 128, (D + DC - 1) / DC, B);
  static const bool block_taps = !(getenv("DPV_SWEEP_BLOCK") && atoi(getenv("DPV_SWEEP_BLOCK")) == 0);   // tuning aid
  if (block_taps && w >= 2 && h >= 2) {
    if (metric == 0)
      plane_sweep_block_kernel<DC, 0><<<grid, 128, 0, st>>>(ref, src, V, C, h, w, K, R, t, rays, rays_batch_stride,
                                                           center, d_candi, D, sigma, cost, sd);
    else
      plane_sweep_block_kernel<DC, 1><<<grid, 128, 0, st>>>(ref, src, V, C, h, w, K, R, t, rays, rays_batch_stride,
                                                           center, d_candi, D, sigma, cost, sd);
  } else if (metric == 0)
    plane_sweep_kernel<DC, 0><<<grid, 128, 0, st>>>(ref, src, V, C, h, w, K, R, t, rays, rays_batch_stride, center,
                                                   d_candi, D, sigma, cost, sd);
  else
    plane_sweep_kernel<DC, 1><<<grid, 128, 0, st>>>(ref, src, V, C, h, w, K, R, t, rays, rays_batch_stride, center,
                                                   d_candi, D, sigma, cost, sd);
  return check_launch("dpv_plane_sweep");
}

int dpv_plane_sweep(const float* ref, const float* src, int B, int V, int C, int h, int w, const float* K,
                    const float* R, const float* t, const float* rays, long long rays_batch_stride,
                    const float* center, const float* d_candi, int D, float sigma, int metric, float* cost,
                    void* stream) {
  return dpv_plane_sweep_strided(ref, (long long)C * h * w, src, (long long)V * C * h * w, B, V, C, h, w, K, R, 9ll * V, t,
                                 3ll * V, rays, rays_batch_stride, center, d_candi, D, sigma, metric, cost, stream);
}

int dpv_warp_feature(const float* src, int B, int T, int C, int h, int w, const float* K, const float* R,
                     const float* t, const float* rays, long long rays_batch_stride, const float* center,
                     const float* d_candi, int D, float* out, void* stream) {
  DPV_REQUIRE(src && K && R && t && rays && center && d_candi && out, "dpv_warp_feature: null pointer");
  DPV_REQUIRE(B > 0 && T > 0 && C > 0 && h > 0 && w > 0 && D > 0, "dpv_warp_feature: bad shape");
  DPV_REQUIRE(C >= D, "dpv_warp_feature: needs C >= D (reference keeps warp[i,i], homography.py:165-166)");
  DPV_REQUIRE((long long)B * T <= 65535 && D <= 65535, "dpv_warp_feature: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((h * w + 127) / 128, D, B * T);
  warp_feature_kernel<<<grid, 128, 0, st>>>(src, T, C, h, w, K, R, t, rays, rays_batch_stride, center, d_candi, D,
                                            out);
  return check_launch("dpv_warp_feature");
}

}  // extern "C"
