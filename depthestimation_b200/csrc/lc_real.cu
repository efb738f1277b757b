// the real-sensor steps either side of the Bayesian fusion.
//   lc_transform_points   external/lcsim/lcsim/src/depth.cpp:143-222 (Depth::transformPoints: project a curtain's
//                         returns into another camera, z-buffer, optional outlier filter), called from
//                         lc/light_curtain.py:751,:831,:1076-1089 (transform_measurement)
//   lc_sweep_sample       models/models.py:924-927 (lc_process_real: for every pixel take the slice of a sweep
//                         volume whose depth is nearest to the planned curtain's and return its intensity)
// One pass over their data each; not on the simulated hot loop.
#include "common.cuh"

namespace dpv {

constexpr int kRealThreads = 256;

static inline unsigned real_blocks(long long n) { return (unsigned)((n + kRealThreads - 1) / kRealThreads); }

struct Mat34 { float intr[9]; float M[16]; };

// Order-preserving map of a non-NaN float onto unsigned integers (and back).
__device__ __forceinline__ unsigned ordered_bits(float z) {
  const unsigned b = __float_as_uint(z);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ordered_float(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// depth.cpp:150-176.  The reference walks the points in order and keeps, per pixel, a point whose z is smaller
// than the stored one (or any point while the pixel still holds 0).  No accepted point has z == 0 (its projection
// divides by z and lands outside the image), so the survivor is the lexicographic minimum of (z, point index):
// one 64-bit atomicMin per point, independent of the order in which threads arrive.
__global__ void __launch_bounds__(kRealThreads)
zbuffer_scatter_kernel(const float* __restrict__ pts, long long n_total, int N, Mat34 m, int width, int height,
                       unsigned long long* __restrict__ keys) {
  const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= n_total) return;
  const int set = (int)(gi / N);
  const unsigned i = (unsigned)(gi - (long long)set * N);
  const float4 p = __ldg(reinterpret_cast<const float4*>(pts) + gi);
  const float q[4] = {p.x, p.y, p.z, 1.0f};
  float c[4];
#pragma unroll
  for (int r = 0; r < 4; r++) {                       // (M_lidar2cam * points^T)^T, one row at a time
    float s = __fmul_rn(m.M[r * 4], q[0]);
#pragma unroll
    for (int k = 1; k < 4; k++) s = __fadd_rn(s, __fmul_rn(m.M[r * 4 + k], q[k]));
    c[r] = s;
  }
  float pr[3];
#pragma unroll
  for (int r = 0; r < 3; r++)
    pr[r] = __fadd_rn(__fadd_rn(__fmul_rn(m.intr[r * 3], c[0]), __fmul_rn(m.intr[r * 3 + 1], c[1])),
                      __fmul_rn(m.intr[r * 3 + 2], c[2]));
  const float pu = __fdiv_rn(pr[0], pr[2]), pv = __fdiv_rn(pr[1], pr[2]);
  // (int)(pu - 0.5) in double; a NaN or out-of-range value converts to INT_MIN on the reference's hosts (cvttsd2si)
  // and is rejected by the bounds test: rejected here before the conversion
  const double du = (double)pu - 0.5, dv = (double)pv - 0.5;
  if (!(du > -2147483649.0 && du < 2147483648.0) || !(dv > -2147483649.0 && dv < 2147483648.0)) return;
  const int u = (int)du, v = (int)dv;
  if (u < 0 || u >= width || v < 0 || v >= height) return;
  const float z = c[2];
  if (z != z) return;
  const unsigned long long key = ((unsigned long long)ordered_bits(z) << 32) | i;
  atomicMin(keys + ((size_t)set * height + v) * width + u, key);
}

// depth.cpp:177-215: the maps of the surviving points (0 where no point fell), then the outlier mask
__global__ void __launch_bounds__(kRealThreads)
zbuffer_resolve_kernel(const unsigned long long* __restrict__ keys, const float* __restrict__ pts,
                       const float* __restrict__ thick, long long npix, int N, size_t HW, float* __restrict__ out) {
  const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= npix) return;
  const size_t set = (size_t)(gi / (long long)HW), hw = (size_t)(gi - (long long)set * HW);
  const unsigned long long key = keys[gi];
  float z = 0.f, it = 0.f, th = 0.f;
  if (key != ~0ull) {
    const unsigned i = (unsigned)(key & 0xffffffffull);
    z = ordered_float((unsigned)(key >> 32));
    it = pts[((size_t)set * N + i) * 4 + 3];
    th = thick[(size_t)set * N + i];
  }
  float* o = out + set * 3 * HW + hw;
  o[0] = z;
  o[HW] = it;
  o[2 * HW] = th;
}

__global__ void __launch_bounds__(kRealThreads)
zbuffer_filter_mask_kernel(const float* __restrict__ maps, long long npix, int width, int height, int off,
                           unsigned char* __restrict__ mask) {
  const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= npix) return;
  const size_t HW = (size_t)width * height;
  const size_t set = (size_t)(gi / (long long)HW), hw = (size_t)(gi - (long long)set * HW);
  const int v = (int)(hw / width), u = (int)(hw - (size_t)v * width);
  unsigned char keep = 1;
  if (v >= off && v < height - off - 1 && u >= off && u < width - off - 1) {
    const float* d = maps + set * 3 * HW;
    const float z = d[hw];
    for (int vv = v - off; vv < v + off + 1; vv++)
      for (int uu = u - off; uu < u + off + 1; uu++) {
        if (vv == v && uu == u) continue;
        const float zn = d[(size_t)vv * width + uu];
        if (zn == 0.f) continue;
        if (__fsub_rn(zn, z) < -1.f) keep = 0;
      }
  }
  mask[gi] = keep;
}

__global__ void __launch_bounds__(kRealThreads)
zbuffer_apply_mask_kernel(const unsigned char* __restrict__ mask, long long npix, size_t HW, float* __restrict__ maps) {
  const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= npix) return;
  const size_t set = (size_t)(gi / (long long)HW), hw = (size_t)(gi - (long long)set * HW);
  const float mk = mask[gi] ? 1.f : 0.f;              // cwiseProduct with the 0/1 mask (keeps NaN and -0 like it)
  float* o = maps + set * 3 * HW + hw;
  o[0] *= mk;
  o[HW] *= mk;
  o[2 * HW] *= mk;
}

// models.py:924-926: inds = argmin_k |sampling_depth - feat_z[k]| (first minimum; a NaN difference wins, the first
// one if several, as torch.argmin does), result = feat_int[inds].  A thread owns one pixel of one sample and walks
// the K slices once for all of the sample's curtains.
template <int NC>
__global__ void __launch_bounds__(kRealThreads)
sweep_sample_kernel(const float* __restrict__ sdepth, const float* __restrict__ feat_z,
                    const float* __restrict__ feat_int, long long npix, int nc, int K, size_t HW,
                    float* __restrict__ out, int32_t* __restrict__ inds_out) {
  const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= npix) return;
  const size_t b = (size_t)(gi / (long long)HW), hw = (size_t)(gi - (long long)b * HW);
  for (int c0 = 0; c0 < nc; c0 += NC) {
    float sd[NC], best[NC];
    int bi[NC];
#pragma unroll
    for (int j = 0; j < NC; j++) {
      sd[j] = (c0 + j < nc) ? sdepth[(b * nc + c0 + j) * HW + hw] : 0.f;
      best[j] = INFINITY;
      bi[j] = -1;
    }
    const float* fz = feat_z + b * K * HW + hw;
    for (int k = 0; k < K; k++) {
      const float z = ld_stream(fz + (size_t)k * HW);
#pragma unroll
      for (int j = 0; j < NC; j++) {
        const float a = fabsf(__fsub_rn(sd[j], z));
        const bool take = bi[j] < 0 || (best[j] == best[j] && (a != a || a < best[j]));
        if (take) { best[j] = a; bi[j] = k; }
      }
    }
#pragma unroll
    for (int j = 0; j < NC; j++)
      if (c0 + j < nc) {
        const size_t o = (b * nc + c0 + j) * HW + hw;
        out[o] = feat_int[(b * K + bi[j]) * HW + hw];
        if (inds_out) inds_out[o] = bi[j];
      }
  }
}

// utils/img_utils.py:269-293 depth_to_pts, written straight into the [N,4] point layout transformPoints consumes
// (lc/light_curtain.py:826-830 builds it with a cat and a transpose): ((u - cx) / fx * d, (v - cy) / fy * d, d, w).
__global__ void __launch_bounds__(kRealThreads)
depth_to_points_kernel(const float* __restrict__ depth, long long npix, int H, int W, float fx, float cx, float fy,
                       float cy, float w4, float4* __restrict__ pts) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const long long hw = i % ((long long)H * W);
  const int v = (int)(hw / W), u = (int)(hw % W);
  const float d = depth[i];
  const float xf = __fdiv_rn(__fsub_rn((float)u, cx), fx), yf = __fdiv_rn(__fsub_rn((float)v, cy), fy);
  pts[i] = make_float4(__fmul_rn(xf, d), __fmul_rn(yf, d), d, w4);
}

// F.interpolate(depth, scale_factor=s, mode='nearest') of the ground-truth depth (models/models.py:986,
// lc/light_curtain.py callers): dst (y, x) <- src (y / s, x / s) for an integer factor.
__global__ void __launch_bounds__(kRealThreads)
upsample_nearest_kernel(const float* __restrict__ in, long long nout, int h, int w, int s, float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nout) return;
  const int W = w * s, H = h * s;
  const int x = (int)(i % W);
  const long long r = i / W;
  const int y = (int)(r % H);
  const long long n = r / H;
  out[i] = in[(n * h + y / s) * w + x / s];
}

}  // namespace dpv

using namespace dpv;

extern "C" {

int dpv_upsample_nearest(const float* in, int n, int h, int w, int scale, float* out, void* stream) {
  DPV_REQUIRE(in && out, "dpv_upsample_nearest: null pointer");
  DPV_REQUIRE(n > 0 && h > 0 && w > 0 && scale >= 1, "dpv_upsample_nearest: bad shape n=%d h=%d w=%d scale=%d", n, h, w, scale);
  const long long nout = (long long)n * h * w * scale * scale;
  upsample_nearest_kernel<<<real_blocks(nout), kRealThreads, 0, (cudaStream_t)stream>>>(in, nout, h, w, scale, out);
  return check_launch("dpv_upsample_nearest");
}

int lc_depth_to_points(const float* depth, int n, int H, int W, const float* intr, float w4, float* pts, void* stream) {
  DPV_REQUIRE(depth && intr && pts, "lc_depth_to_points: null pointer");
  DPV_REQUIRE(n > 0 && H > 0 && W > 0, "lc_depth_to_points: bad shape n=%d H=%d W=%d", n, H, W);
  DPV_REQUIRE(((uintptr_t)pts & 15) == 0, "lc_depth_to_points: pts must be 16-byte aligned");
  const long long npix = (long long)n * H * W;
  depth_to_points_kernel<<<real_blocks(npix), kRealThreads, 0, (cudaStream_t)stream>>>(
      depth, npix, H, W, intr[0], intr[2], intr[4], intr[5], w4, reinterpret_cast<float4*>(pts));
  return check_launch("lc_depth_to_points");
}

size_t lc_transform_points_scratch_bytes(int n, int width, int height, int filtering) {
  const size_t px = (size_t)n * width * height;
  return px * 8 + (filtering ? px : 0);
}

int lc_transform_points(const float* pts, const float* thick, int n, int N, const float* intr, const float* M,
                        int width, int height, int filtering, float* maps, void* scratch, size_t scratch_bytes,
                        void* stream) {
  DPV_REQUIRE(pts && thick && intr && M && maps && scratch, "lc_transform_points: null pointer");
  DPV_REQUIRE(n > 0 && N > 0 && width > 0 && height > 0 && filtering >= 0, "lc_transform_points: bad shape");
  DPV_REQUIRE(((uintptr_t)pts & 15) == 0 && ((uintptr_t)scratch & 7) == 0, "lc_transform_points: pts must be 16-byte, scratch 8-byte aligned");
  DPV_REQUIRE(scratch_bytes >= lc_transform_points_scratch_bytes(n, width, height, filtering),
              "lc_transform_points: scratch of %zu bytes is too small", scratch_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  Mat34 m;
  for (int i = 0; i < 9; i++) m.intr[i] = intr[i];
  for (int i = 0; i < 16; i++) m.M[i] = M[i];
  const size_t HW = (size_t)width * height;
  const long long npix = (long long)n * (long long)HW, npts = (long long)n * N;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(scratch);
  DPV_CUDA(cudaMemsetAsync(keys, 0xff, (size_t)npix * 8, st));
  zbuffer_scatter_kernel<<<real_blocks(npts), kRealThreads, 0, st>>>(pts, npts, N, m, width, height, keys);
  int rc = check_launch("lc_transform_points/scatter");
  if (rc) return rc;
  zbuffer_resolve_kernel<<<real_blocks(npix), kRealThreads, 0, st>>>(keys, pts, thick, npix, N, HW, maps);
  rc = check_launch("lc_transform_points/resolve");
  if (rc || !filtering) return rc;
  unsigned char* mask = reinterpret_cast<unsigned char*>(keys + npix);
  zbuffer_filter_mask_kernel<<<real_blocks(npix), kRealThreads, 0, st>>>(maps, npix, width, height, filtering, mask);
  rc = check_launch("lc_transform_points/filter");
  if (rc) return rc;
  zbuffer_apply_mask_kernel<<<real_blocks(npix), kRealThreads, 0, st>>>(mask, npix, HW, maps);
  return check_launch("lc_transform_points/mask");
}

int lc_sweep_sample(const float* sampling_depth, const float* feat_z, const float* feat_int, int B, int nc, int K,
                    int H, int W, float* out, int32_t* inds, void* stream) {
  DPV_REQUIRE(sampling_depth && feat_z && feat_int && out, "lc_sweep_sample: null pointer");
  DPV_REQUIRE(B > 0 && nc > 0 && K > 0 && H > 0 && W > 0, "lc_sweep_sample: bad shape");
  const size_t HW = (size_t)H * W;
  const long long npix = (long long)B * (long long)HW;
  cudaStream_t st = (cudaStream_t)stream;
  if (nc == 1)
    sweep_sample_kernel<1><<<real_blocks(npix), kRealThreads, 0, st>>>(sampling_depth, feat_z, feat_int, npix, nc, K,
                                                                       HW, out, inds);
  else
    sweep_sample_kernel<4><<<real_blocks(npix), kRealThreads, 0, st>>>(sampling_depth, feat_z, feat_int, npix, nc, K,
                                                                       HW, out, inds);
  return check_launch("lc_sweep_sample");
}

}  // extern "C"
