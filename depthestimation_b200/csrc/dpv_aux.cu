// Rows a17-a19 and (f)1 of SURVEY §8: the callers either side of the fused path.
//   dpv_truth_dpv        utils/img_utils.py:567-582 (gen_dpv_withmask) + :176-195 (soft label, uniform)
//   lc_likelihood_true   lc/light_curtain.py:1106-1125 (gen_lc_dpv_true) + utils/img_utils.py:144-157
//   dpv_inverse_warp     utils/inverse_warp.py:26-69, :174-210
//   dpv_correlation      models/correlation_package/correlation_cuda_kernel.cu:44-112 (forward),
//                        correlation_cuda.cc:21-38 (output shape); twin models/correlation_native.py:6-23
//   dpv_depth_errors     external/deval_lib/src/evaluate_depth.h:20-118 (depthError)
//   dpv_unc_rmse         utils/img_utils.py:371-382 (compute_unc_rmse; its "rmse" is the mean absolute error)
// None of these is on the per-iteration hot loop; they are written for one pass over their data with coalesced
// accesses, not tuned further.
#include "common.cuh"

namespace dpv {

constexpr int kAuxThreads = 256;

static inline unsigned aux_blocks(long long n) { return (unsigned)((n + kAuxThreads - 1) / kAuxThreads); }

// ---------------------------------------------------------------------------------------------
// gen_dpv_withmask: soft label around the true depth where the mask is set, uniform elsewhere
// ---------------------------------------------------------------------------------------------
__global__ void truth_dpv_kernel(const float* __restrict__ dmaps, const float* __restrict__ masks, long long npix,
                                 size_t HW, const float* __restrict__ d_candi, int D, float var,
                                 float* __restrict__ out) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const size_t b = (size_t)(pix / (long long)HW), hw = (size_t)(pix - (long long)b * HW);
  const float z = dmaps[pix], m = masks[pix], mi = 1.f - m;
  const float sig = sqrtf(var);
  const float den = 2.f * (sig * sig);                       // 2 * pow(sigma, 2)
  float sum = 0.f;
  for (int d = 0; d < D; d++) {
    const float a = fabsf(d_candi[d] - z);
    sum += expf(-(a * a) / den);
  }
  const float uni = 1.f / (float)D;
  float* o = out + b * D * HW + hw;
  for (int d = 0; d < D; d++) {
    const float a = fabsf(d_candi[d] - z);
    float t = expf(-(a * a) / den) / sum;
    if (t != t) t = -1.f;                                    // zero_invalid (img_utils.py:189-190)
    const float v = t * m + uni * mi;
    st_stream(o + (size_t)d * HW, (v != v) ? v : fminf(fmaxf(v, eps_floor()), 1.f));   // torch.clamp keeps NaN
  }
}

// ---------------------------------------------------------------------------------------------
// a17: intensity-model likelihood of a real light-curtain return
// ---------------------------------------------------------------------------------------------
__global__ void likelihood_true_kernel(const float* __restrict__ placement, const float* __restrict__ intensity,
                                       const float* __restrict__ inten_sigma, const float* __restrict__ peak,
                                       float peak_const, float noise_sigma, long long npix,
                                       const float* __restrict__ d_candi, int D, float* __restrict__ mean_out,
                                       float* __restrict__ lik_out) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const float z = placement[pix], x = intensity[pix], sg = inten_sigma[pix];
  const float pk = peak ? peak[pix] : peak_const;
  const float var = noise_sigma * noise_sigma;
  const float denom = (float)sqrt(2.0 * 3.141592653589793 * (double)var);      // (2*math.pi*var)**.5 (python floats)
  const float two_var = 2.f * var;
  // pass 1: running maximum / sum of exp for logsumexp; NaN must poison the pixel like torch.logsumexp
  float mx = -INFINITY;
  bool bad = false;
  for (int d = 0; d < D; d++) {
    const float e = fabsf(d_candi[d] - z) / sg;
    const float mi = expf(-(e * e)) * pk;
    const float t = x - mi;
    const float ll = logf(expf(-(t * t) / two_var) / denom);
    bad = bad || (ll != ll);
    mx = fmaxf(mx, ll);
  }
  float se = 0.f;
  for (int d = 0; d < D; d++) {
    const float e = fabsf(d_candi[d] - z) / sg;
    const float mi = expf(-(e * e)) * pk;
    const float t = x - mi;
    const float ll = logf(expf(-(t * t) / two_var) / denom);
    se += expf(ll - mx);
  }
  const float lse = bad ? NAN : (mx + logf(se));
  for (int d = 0; d < D; d++) {
    const float e = fabsf(d_candi[d] - z) / sg;
    const float mi = expf(-(e * e)) * pk;
    const float t = x - mi;
    const float ll = logf(expf(-(t * t) / two_var) / denom);
    if (mean_out) mean_out[(size_t)d * npix + pix] = mi;
    lik_out[(size_t)d * npix + pix] = expf(ll - lse);
  }
}

// ---------------------------------------------------------------------------------------------
// a18: inverse warp (pixel2cam -> cam2pixel -> grid_sample, align_corners=False)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float reflect_coord(float x, float lo2, float hi2) {   // ATen reflect_coordinates
  if (lo2 == hi2) return 0.f;
  const float mn = lo2 / 2.f, span = (hi2 - lo2) / 2.f;
  x = fabsf(x - mn);
  const float extra = fmodf(x, span);
  const int flips = (int)floorf(x / span);
  return (flips % 2 == 0) ? extra + mn : span - extra + mn;
}

// ATen clip_coordinates_set_grad / reflect_coordinates_set_grad: the coordinate after padding and d(out)/d(in)
__device__ __forceinline__ float clip_coord_grad(float x, int size, float* mult) {
  if (x <= 0.f) { *mult = 0.f; return 0.f; }
  const float mx = (float)(size - 1);
  if (x >= mx) { *mult = 0.f; return mx; }
  *mult = 1.f;
  return x;
}
__device__ __forceinline__ float reflect_coord_grad(float x, float lo2, float hi2, float* mult) {
  if (lo2 == hi2) { *mult = 0.f; return 0.f; }
  const float mn = lo2 / 2.f, span = (hi2 - lo2) / 2.f;
  x = x - mn;
  float m = 1.f;
  if (x < 0.f) { m = -1.f; x = -x; }
  const float extra = fmodf(x, span);
  const int flips = (int)floorf(x / span);
  if (flips % 2 == 0) { *mult = m; return extra + mn; }
  *mult = -m;
  return span - extra + mn;
}

// Backward of inverse_warp_kernel (what torch autograd derives for utils/inverse_warp.py:174-210 in the reference:
// grid_sample backward -> cam2pixel :43-69 -> pixel2cam :26-40).  One thread per output pixel: scatters the image
// gradient to its taps, chains the grid gradient through the projection to the depth, and reduces the gradient of
// the 3x4 projection matrix over the warp before one atomicAdd per entry.
__global__ void inverse_warp_backward_kernel(const float* __restrict__ gout, const float* __restrict__ img,
                                             const float* __restrict__ depth, const float* __restrict__ Kinv,
                                             const float* __restrict__ P, int B, int C, int H, int W, int mode,
                                             int padding, float* __restrict__ gimg, float* __restrict__ gdepth,
                                             float* __restrict__ gP) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long npix = (long long)B * H * W;
  const bool active = pix < npix;
  const size_t HW = (size_t)H * W;
  const int b = active ? (int)(pix / (long long)HW) : B - 1;
  float gp[12];
#pragma unroll
  for (int k = 0; k < 12; k++) gp[k] = 0.f;
  if (active) {
    const size_t hw = (size_t)(pix - (long long)b * HW);
    const int y = (int)(hw / W), x = (int)(hw % W);
    const float* ki = Kinv + b * 9;
    const float* p = P + b * 12;
    const float dz = depth[pix];
    const float fx = (float)x, fy = (float)y;
    const float kx = ki[0] * fx + ki[1] * fy + ki[2], ky = ki[3] * fx + ki[4] * fy + ki[5],
                kz = ki[6] * fx + ki[7] * fy + ki[8];
    const float cx = kx * dz, cy = ky * dz, cz = kz * dz;
    const float X = (p[0] * cx + p[1] * cy + p[2] * cz) + p[3];
    const float Y = (p[4] * cx + p[5] * cy + p[6] * cz) + p[7];
    const float Zraw = (p[8] * cx + p[9] * cy + p[10] * cz) + p[11];
    const float Z = fmaxf(Zraw, 1e-3f);
    const float gx = 2.f * (X / Z) / (float)(W - 1) - 1.f, gy = 2.f * (Y / Z) / (float)(H - 1) - 1.f;
    float ix = gs_unnormalize(gx, W), iy = gs_unnormalize(gy, H);
    float mx = 1.f, my = 1.f;                                   // d(ix after padding) / d(ix)
    if (padding == 1) {
      ix = clip_coord_grad(ix, W, &mx);
      iy = clip_coord_grad(iy, H, &my);
    } else if (padding == 2) {
      float m1, m2;
      ix = clip_coord_grad(reflect_coord_grad(ix, -1.f, 2.f * W - 1.f, &m1), W, &m2);
      mx = m1 * m2;
      iy = clip_coord_grad(reflect_coord_grad(iy, -1.f, 2.f * H - 1.f, &m1), H, &m2);
      my = m1 * m2;
    }
    mx *= (float)W / 2.f;                                       // un-normalisation, align_corners=False
    my *= (float)H / 2.f;
    const float* src = img + (size_t)b * C * HW;
    float* gsrc = gimg ? gimg + (size_t)b * C * HW : nullptr;
    const float* go = gout + (size_t)b * C * HW + hw;
    const bool finite = (ix == ix) && (iy == iy) && fabsf(ix) < 1e9f && fabsf(iy) < 1e9f;
    float gix = 0.f, giy = 0.f;
    if (mode == 1) {                                            // nearest: no gradient to the grid
      const float rx = nearbyintf(ix), ry = nearbyintf(iy);
      if (finite && gsrc && rx >= 0.f && rx <= (float)(W - 1) && ry >= 0.f && ry <= (float)(H - 1)) {
        const size_t o = (size_t)((int)ry) * W + (int)rx;
        for (int c = 0; c < C; c++) atomicAdd(gsrc + (size_t)c * HW + o, go[(size_t)c * HW]);
      }
    } else if (finite) {
      const float x0f = floorf(ix), y0f = floorf(iy);
      const float tx = ix - x0f, ty = iy - y0f;
      const int x0 = (int)x0f, y0 = (int)y0f;
      const bool xin0 = x0 >= 0 && x0 < W, xin1 = x0 + 1 >= 0 && x0 + 1 < W;
      const bool yin0 = y0 >= 0 && y0 < H, yin1 = y0 + 1 >= 0 && y0 + 1 < H;
      const float wnw = (1.f - tx) * (1.f - ty), wne = tx * (1.f - ty), wsw = (1.f - tx) * ty, wse = tx * ty;
      const size_t onw = (size_t)y0 * W + x0;
      for (int c = 0; c < C; c++) {
        const float g = go[(size_t)c * HW];
        const float* s = src + (size_t)c * HW;
        float* gs = gsrc ? gsrc + (size_t)c * HW : nullptr;
        if (yin0 && xin0) {
          const float v = s[onw];
          gix -= v * (1.f - ty) * g; giy -= v * (1.f - tx) * g;
          if (gs) atomicAdd(gs + onw, wnw * g);
        }
        if (yin0 && xin1) {
          const float v = s[onw + 1];
          gix += v * (1.f - ty) * g; giy -= v * tx * g;
          if (gs) atomicAdd(gs + onw + 1, wne * g);
        }
        if (yin1 && xin0) {
          const float v = s[onw + W];
          gix -= v * ty * g; giy += v * (1.f - tx) * g;
          if (gs) atomicAdd(gs + onw + W, wsw * g);
        }
        if (yin1 && xin1) {
          const float v = s[onw + W + 1];
          gix += v * ty * g; giy += v * tx * g;
          if (gs) atomicAdd(gs + onw + W + 1, wse * g);
        }
      }
    }
    const float Ggx = mx * gix, Ggy = my * giy;                 // d loss / d (normalised grid)
    const float sx = 2.f / (float)(W - 1), sy = 2.f / (float)(H - 1);
    const float gX = Ggx * sx / Z, gY = Ggy * sy / Z;
    const float gZ = Zraw >= 1e-3f ? -(Ggx * sx * X + Ggy * sy * Y) / (Z * Z) : 0.f;      // clamp(min=1e-3) backward
    if (gdepth) {
      const float gcx = p[0] * gX + p[4] * gY + p[8] * gZ, gcy = p[1] * gX + p[5] * gY + p[9] * gZ,
                  gcz = p[2] * gX + p[6] * gY + p[10] * gZ;
      gdepth[pix] = gcx * kx + gcy * ky + gcz * kz;
    }
    gp[0] = gX * cx; gp[1] = gX * cy; gp[2] = gX * cz; gp[3] = gX;
    gp[4] = gY * cx; gp[5] = gY * cy; gp[6] = gY * cz; gp[7] = gY;
    gp[8] = gZ * cx; gp[9] = gZ * cy; gp[10] = gZ * cz; gp[11] = gZ;
  }
  if (!gP) return;
  const unsigned full = 0xffffffffu;
  const int b0 = __shfl_sync(full, b, 0);
  if (__all_sync(full, b == b0 || !active)) {                   // the usual case: the warp lies inside one image
#pragma unroll
    for (int k = 0; k < 12; k++) {
      float v = gp[k];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(full, v, off);
      if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(gP + b0 * 12 + k, v);
    }
  } else if (active) {
#pragma unroll
    for (int k = 0; k < 12; k++) atomicAdd(gP + b * 12 + k, gp[k]);
  }
}

// mode: 0 bilinear, 1 nearest; padding: 0 zeros, 1 border, 2 reflection
__global__ void inverse_warp_kernel(const float* __restrict__ img, const float* __restrict__ depth,
                                    const float* __restrict__ Kinv, const float* __restrict__ P, int B, int C, int H,
                                    int W, int mode, int padding, float* __restrict__ out,
                                    uint8_t* __restrict__ valid) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long npix = (long long)B * H * W;
  if (pix >= npix) return;
  const size_t HW = (size_t)H * W;
  const int b = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)b * HW);
  const int y = (int)(hw / W), x = (int)(hw % W);
  const float* ki = Kinv + b * 9;
  const float* p = P + b * 12;
  const float dz = depth[pix];
  // cam = (Kinv @ [x, y, 1]) * depth   (inverse_warp.py:36-40)
  const float fx = (float)x, fy = (float)y;
  const float cx = (ki[0] * fx + ki[1] * fy + ki[2]) * dz;
  const float cy = (ki[3] * fx + ki[4] * fy + ki[5]) * dz;
  const float cz = (ki[6] * fx + ki[7] * fy + ki[8]) * dz;
  // pcoords = rot @ cam + tr          (:54-60)
  const float X = (p[0] * cx + p[1] * cy + p[2] * cz) + p[3];
  const float Y = (p[4] * cx + p[5] * cy + p[6] * cz) + p[7];
  const float Z = fmaxf((p[8] * cx + p[9] * cy + p[10] * cz) + p[11], 1e-3f);
  const float gx = 2.f * (X / Z) / (float)(W - 1) - 1.f;      // :65-66
  const float gy = 2.f * (Y / Z) / (float)(H - 1) - 1.f;
  if (valid) valid[pix] = (gx == gx && gy == gy && fmaxf(fabsf(gx), fabsf(gy)) <= 1.f) ? 1 : 0;      // :208
  float ix = gs_unnormalize(gx, W), iy = gs_unnormalize(gy, H);
  if (padding == 1) {
    ix = fminf(fmaxf(ix, 0.f), (float)(W - 1));
    iy = fminf(fmaxf(iy, 0.f), (float)(H - 1));
  } else if (padding == 2) {
    ix = fminf(fmaxf(reflect_coord(ix, -1.f, 2.f * W - 1.f), 0.f), (float)(W - 1));
    iy = fminf(fmaxf(reflect_coord(iy, -1.f, 2.f * H - 1.f), 0.f), (float)(H - 1));
  }
  const float* src = img + (size_t)b * C * HW;
  float* dst = out + (size_t)b * C * HW + hw;
  if (mode == 1) {
    const float rx = nearbyintf(ix), ry = nearbyintf(iy);
    const bool in = rx >= 0.f && rx <= (float)(W - 1) && ry >= 0.f && ry <= (float)(H - 1);
    const size_t o = in ? (size_t)((int)ry) * W + (int)rx : 0;
    for (int c = 0; c < C; c++) dst[(size_t)c * HW] = in ? src[(size_t)c * HW + o] : 0.f;
    return;
  }
  const float x0f = floorf(ix), y0f = floorf(iy);
  const float tx = ix - x0f, ty = iy - y0f;
  const float wnw = (1.f - tx) * (1.f - ty), wne = tx * (1.f - ty), wsw = (1.f - tx) * ty, wse = tx * ty;
  const bool finite = (ix == ix) && (iy == iy) && fabsf(ix) < 1e9f && fabsf(iy) < 1e9f;
  const int x0 = finite ? (int)x0f : -2, y0 = finite ? (int)y0f : -2;
  const bool xin0 = x0 >= 0 && x0 < W, xin1 = x0 + 1 >= 0 && x0 + 1 < W;
  const bool yin0 = y0 >= 0 && y0 < H, yin1 = y0 + 1 >= 0 && y0 + 1 < H;
  for (int c = 0; c < C; c++) {
    const float* s = src + (size_t)c * HW;
    float acc = 0.f;                                           // ATen order: nw, ne, sw, se
    if (yin0 && xin0) acc += s[(size_t)y0 * W + x0] * wnw;
    if (yin0 && xin1) acc += s[(size_t)y0 * W + x0 + 1] * wne;
    if (yin1 && xin0) acc += s[(size_t)(y0 + 1) * W + x0] * wsw;
    if (yin1 && xin1) acc += s[(size_t)(y0 + 1) * W + x0 + 1] * wse;
    dst[(size_t)c * HW] = acc;
  }
}

// ---------------------------------------------------------------------------------------------
// a19: FlowNet correlation, forward.  One thread per output element; the zero padding is index math.
// ---------------------------------------------------------------------------------------------
__global__ void correlation_kernel(const float* __restrict__ in1, const float* __restrict__ in2, int B, int C, int H,
                                   int W, int pad, int ksize, int max_disp, int stride1, int stride2, int outH,
                                   int outW, float* __restrict__ out) {
  const int krad = (ksize - 1) / 2, drad = max_disp / stride2, dsize = 2 * drad + 1;
  const int outC = dsize * dsize;
  const long long n = (long long)B * outC * outH * outW;
  const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= n) return;
  const int ox = (int)(id % outW), oy = (int)((id / outW) % outH);
  const int tc = (int)((id / ((long long)outW * outH)) % outC), b = (int)(id / ((long long)outW * outH * outC));
  const int tj = tc / dsize - drad, ti = tc % dsize - drad;
  const int x1 = ox * stride1 + max_disp - pad, y1 = oy * stride1 + max_disp - pad;   // unpadded coordinates
  const int x2 = x1 + ti * stride2, y2 = y1 + tj * stride2;
  const size_t HW = (size_t)H * W;
  const float* a = in1 + (size_t)b * C * HW;
  const float* c2 = in2 + (size_t)b * C * HW;
  float acc = 0.f;
  for (int j = -krad; j <= krad; j++)
    for (int i = -krad; i <= krad; i++) {
      const int ya = y1 + j, xa = x1 + i, yb = y2 + j, xb = x2 + i;
      if (ya < 0 || ya >= H || xa < 0 || xa >= W || yb < 0 || yb >= H || xb < 0 || xb >= W) continue;
      const size_t oa = (size_t)ya * W + xa, ob = (size_t)yb * W + xb;
      for (int ch = 0; ch < C; ch++) acc += a[(size_t)ch * HW + oa] * c2[(size_t)ch * HW + ob];
    }
  out[id] = acc / (float)(ksize * ksize * C);
}

// a19 backward (correlation_cuda_kernel.cu:116-303, stride1 = 1): one thread per input element gathers, for
// every displacement, the outputs whose kernel window covers it.  No atomics: both gradients are pure gathers.
__global__ void correlation_backward_kernel(const float* __restrict__ gout, const float* __restrict__ in1,
                                            const float* __restrict__ in2, int B, int C, int H, int W, int pad,
                                            int ksize, int max_disp, int stride2, int outH, int outW,
                                            float* __restrict__ gin1, float* __restrict__ gin2) {
  const int krad = (ksize - 1) / 2, drad = max_disp / stride2, dsize = 2 * drad + 1;
  const int outC = dsize * dsize;
  const long long n = (long long)B * C * H * W;
  const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= n) return;
  const int x = (int)(id % W), y = (int)((id / W) % H);
  const int b = (int)(id / ((long long)W * H * C));
  const size_t HW = (size_t)H * W, oHW = (size_t)outH * outW;
  const size_t plane = (size_t)(id / HW) * HW;                     // (b, c) plane of the inputs
  const float* g = gout + (size_t)b * outC * oHW;
  const int py = y + pad, px = x + pad;                            // the reference's padded coordinates
  const float nel = (float)(ksize * ksize * C);
  // d/d in1[y,x]: outputs (j,i) with j + max_disp - krad <= py <= j + max_disp + krad  (:135-159)
  float acc1 = 0.f;
  {
    const int ymin = max(0, py - krad - max_disp), ymax = min(outH - 1, py + krad - max_disp);
    const int xmin = max(0, px - krad - max_disp), xmax = min(outW - 1, px + krad - max_disp);
    if (ymin <= ymax && xmin <= xmax)
      for (int tc = 0; tc < outC; tc++) {
        const int j2 = (tc / dsize - drad) * stride2, i2 = (tc % dsize - drad) * stride2;
        const int y2 = y + j2, x2 = x + i2;
        if (y2 < 0 || y2 >= H || x2 < 0 || x2 >= W) continue;      // the zero frame of rInput2
        const float v2 = in2[plane + (size_t)y2 * W + x2];
        for (int j = ymin; j <= ymax; j++)
          for (int i = xmin; i <= xmax; i++) acc1 += g[(size_t)tc * oHW + (size_t)j * outW + i] * v2;
      }
  }
  gin1[id] = acc1 / nel;
  // d/d in2[y,x]: the same window shifted by the displacement (:254-287)
  float acc2 = 0.f;
  for (int tc = 0; tc < outC; tc++) {
    const int j2 = (tc / dsize - drad) * stride2, i2 = (tc % dsize - drad) * stride2;
    const int ymin = max(0, py - krad - max_disp - j2), ymax = min(outH - 1, py + krad - max_disp - j2);
    const int xmin = max(0, px - krad - max_disp - i2), xmax = min(outW - 1, px + krad - max_disp - i2);
    if (ymin > ymax || xmin > xmax) continue;
    const int y1 = y - j2, x1 = x - i2;
    if (y1 < 0 || y1 >= H || x1 < 0 || x1 >= W) continue;          // the zero frame of rInput1
    const float v1 = in1[plane + (size_t)y1 * W + x1];
    for (int j = ymin; j <= ymax; j++)
      for (int i = xmin; i <= xmax; i++) acc2 += g[(size_t)tc * oHW + (size_t)j * outW + i] * v1;
  }
  gin2[id] = acc2 / nel;
}

// ---------------------------------------------------------------------------------------------
// (f)1: KITTI depth error metrics, one CTA per image pair; sums in double (the reference's sequential
// float sums are only reproducible to ~1e-5 relative anyway)
// ---------------------------------------------------------------------------------------------
__global__ void depth_errors_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int HW,
                                    float* __restrict__ out) {
  const float* p = pred + (size_t)blockIdx.x * HW;
  const float* g = gt + (size_t)blockIdx.x * HW;
  double acc[10];
#pragma unroll
  for (int k = 0; k < 10; k++) acc[k] = 0.0;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) {
    const float dg = g[i], di = p[i];
    if (!(dg >= 0.f)) continue;                               // io_depth.h:99-101
    const float e = fabsf(dg - di);
    const float e2 = e * e;
    const float ei = (float)fabs(1.0 / (double)dg - 1.0 / (double)di);
    const float lg = logf(dg) - logf(di);
    const float el = fabsf(lg);
    acc[0] += e; acc[1] += e2; acc[2] += ei; acc[3] += ei * ei; acc[4] += el; acc[5] += el * el;
    acc[6] += lg; acc[7] += e / dg; acc[8] += e2 / (dg * dg); acc[9] += 1.0;
  }
  __shared__ double sh[10][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 10; k++) {
    double v = acc[k];
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) sh[k][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s[10];
    for (int k = 0; k < 10; k++) {
      s[k] = 0.0;
      for (int w = 0; w < (int)(blockDim.x >> 5); w++) s[k] += sh[k][w];
    }
    float* o = out + (size_t)blockIdx.x * 10;
    const float n = (float)s[9];
    const float sq_log = (float)s[5] / n, ls = (float)s[6];
    o[0] = (float)s[0] / n;
    o[1] = sqrtf((float)s[1] / n);
    o[2] = (float)s[2] / n;
    o[3] = sqrtf((float)s[3] / n);
    o[4] = (float)s[4] / n;
    o[5] = sqrtf(sq_log);
    o[6] = sqrtf(sq_log - (ls * ls / (n * n)));
    o[7] = (float)s[7] / n;
    o[8] = (float)s[8] / n;
    o[9] = n;                                                  // valid ground-truth pixels (0: the reference throws)
  }
}

// compute_unc_rmse: E[d] of both fields per column, ends of the prediction zeroed, NaN columns dropped
__global__ void unc_rmse_kernel(const float* __restrict__ truth, const float* __restrict__ pred, int D, int W,
                                const float* __restrict__ d_candi, float* __restrict__ out) {
  __shared__ float s_sum[32], s_cnt[32];
  float sum = 0.f, cnt = 0.f;
  for (int w = threadIdx.x; w < W; w += blockDim.x) {
    float et = 0.f, ep = 0.f;
    for (int d = 0; d < D; d++) {
      et += d_candi[d] * truth[(size_t)d * W + w];
      ep += d_candi[d] * pred[(size_t)d * W + w];
    }
    if (w == 0 || w == W - 1) ep = 0.f;
    if (et == et && ep == ep) { sum += fabsf(et - ep); cnt += 1.f; }
  }
  for (int off = 16; off > 0; off >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, off);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
  }
  if ((threadIdx.x & 31) == 0) { s_sum[threadIdx.x >> 5] = sum; s_cnt[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, c = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++) { a += s_sum[i]; c += s_cnt[i]; }
    out[0] = a / c;
  }
}

}  // namespace dpv

using namespace dpv;

extern "C" {

int dpv_truth_dpv(const float* dmaps, const float* masks, int B, int H, int W, const float* d_candi, int D,
                  float var, float* out, void* stream) {
  DPV_REQUIRE(dmaps && masks && d_candi && out, "dpv_truth_dpv: null pointer");
  DPV_REQUIRE(B > 0 && H > 0 && W > 0 && D > 0, "dpv_truth_dpv: bad shape B=%d D=%d H=%d W=%d", B, D, H, W);
  const long long npix = (long long)B * H * W;
  truth_dpv_kernel<<<aux_blocks(npix), kAuxThreads, 0, (cudaStream_t)stream>>>(dmaps, masks, npix, (size_t)H * W,
                                                                                d_candi, D, var, out);
  return check_launch("dpv_truth_dpv");
}

int lc_likelihood_true(const float* placement, const float* intensity, const float* inten_sigma, const float* peak,
                       float peak_const, float noise_sigma, long long npix, const float* d_candi, int D,
                       float* mean_out, float* lik_out, void* stream) {
  DPV_REQUIRE(placement && intensity && inten_sigma && d_candi && lik_out, "lc_likelihood_true: null pointer");
  DPV_REQUIRE(npix > 0 && D > 0, "lc_likelihood_true: bad shape npix=%lld D=%d", npix, D);
  likelihood_true_kernel<<<aux_blocks(npix), kAuxThreads, 0, (cudaStream_t)stream>>>(
      placement, intensity, inten_sigma, peak, peak_const, noise_sigma, npix, d_candi, D, mean_out, lik_out);
  return check_launch("lc_likelihood_true");
}

int dpv_inverse_warp(const float* img, const float* depth, const float* Kinv, const float* P, int B, int C, int H,
                     int W, int mode, int padding, float* out, uint8_t* valid, void* stream) {
  DPV_REQUIRE(img && depth && Kinv && P && out, "dpv_inverse_warp: null pointer");
  DPV_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1, "dpv_inverse_warp: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
  DPV_REQUIRE(mode == 0 || mode == 1, "dpv_inverse_warp: mode %d (0 bilinear, 1 nearest)", mode);
  DPV_REQUIRE(padding >= 0 && padding <= 2, "dpv_inverse_warp: padding %d (0 zeros, 1 border, 2 reflection)", padding);
  const long long npix = (long long)B * H * W;
  inverse_warp_kernel<<<aux_blocks(npix), kAuxThreads, 0, (cudaStream_t)stream>>>(img, depth, Kinv, P, B, C, H, W,
                                                                                   mode, padding, out, valid);
  return check_launch("dpv_inverse_warp");
}

int dpv_inverse_warp_backward(const float* grad_out, const float* img, const float* depth, const float* Kinv,
                              const float* P, int B, int C, int H, int W, int mode, int padding, float* grad_img,
                              float* grad_depth, float* grad_P, void* stream) {
  DPV_REQUIRE(grad_out && img && depth && Kinv && P, "dpv_inverse_warp_backward: null pointer");
  DPV_REQUIRE(grad_img || grad_depth || grad_P, "dpv_inverse_warp_backward: no gradient requested");
  DPV_REQUIRE(B > 0 && C > 0 && H > 1 && W > 1, "dpv_inverse_warp_backward: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
  DPV_REQUIRE(mode == 0 || mode == 1, "dpv_inverse_warp_backward: mode %d (0 bilinear, 1 nearest)", mode);
  DPV_REQUIRE(padding >= 0 && padding <= 2, "dpv_inverse_warp_backward: padding %d", padding);
  cudaStream_t st = (cudaStream_t)stream;
  const long long npix = (long long)B * H * W;
  if (grad_img) DPV_CUDA(cudaMemsetAsync(grad_img, 0, (size_t)npix * C * sizeof(float), st));     // scatter target
  if (grad_P) DPV_CUDA(cudaMemsetAsync(grad_P, 0, (size_t)B * 12 * sizeof(float), st));
  inverse_warp_backward_kernel<<<aux_blocks(npix), kAuxThreads, 0, st>>>(grad_out, img, depth, Kinv, P, B, C, H, W,
                                                                         mode, padding, grad_img, grad_depth, grad_P);
  return check_launch("dpv_inverse_warp_backward");
}

int dpv_correlation_shape(int H, int W, int pad_size, int kernel_size, int max_displacement, int stride1,
                          int stride2, int* outC, int* outH, int* outW) {
  DPV_REQUIRE(outC && outH && outW, "dpv_correlation_shape: null pointer");
  DPV_REQUIRE(kernel_size >= 1 && (kernel_size & 1) && stride1 >= 1 && stride2 >= 1 && max_displacement >= 0 &&
                  pad_size >= 0, "dpv_correlation_shape: bad parameters");
  const int krad = (kernel_size - 1) / 2, border = krad + max_displacement;      // correlation_cuda.cc:28-34
  const int d = (max_displacement / stride2) * 2 + 1;
  *outC = d * d;
  const int ph = H + 2 * pad_size - 2 * border, pw = W + 2 * pad_size - 2 * border;
  *outH = (ph + stride1 - 1) / stride1;                                          // ceil
  *outW = (pw + stride1 - 1) / stride1;
  DPV_REQUIRE(*outH > 0 && *outW > 0, "dpv_correlation_shape: empty output (%d x %d)", *outH, *outW);
  return DPV_OK;
}

int dpv_correlation(const float* in1, const float* in2, int B, int C, int H, int W, int pad_size, int kernel_size,
                    int max_displacement, int stride1, int stride2, float* out, void* stream) {
  DPV_REQUIRE(in1 && in2 && out, "dpv_correlation: null pointer");
  DPV_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, "dpv_correlation: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
  int oc, oh, ow;
  const int rc = dpv_correlation_shape(H, W, pad_size, kernel_size, max_displacement, stride1, stride2, &oc, &oh, &ow);
  if (rc != DPV_OK) return rc;
  // the reference reads its zero-padded copy without bound checks: every tap must stay inside the padded frame
  const int krad = (kernel_size - 1) / 2;
  DPV_REQUIRE((oh - 1) * stride1 + 2 * max_displacement + krad < H + 2 * pad_size &&
                  (ow - 1) * stride1 + 2 * max_displacement + krad < W + 2 * pad_size,
              "dpv_correlation: taps leave the padded frame (pad_size %d too small)", pad_size);
  const long long n = (long long)B * oc * oh * ow;
  correlation_kernel<<<aux_blocks(n), kAuxThreads, 0, (cudaStream_t)stream>>>(
      in1, in2, B, C, H, W, pad_size, kernel_size, max_displacement, stride1, stride2, oh, ow, out);
  return check_launch("dpv_correlation");
}

int dpv_correlation_backward(const float* grad_out, const float* in1, const float* in2, int B, int C, int H, int W,
                             int pad_size, int kernel_size, int max_displacement, int stride1, int stride2,
                             float* grad_in1, float* grad_in2, void* stream) {
  DPV_REQUIRE(grad_out && in1 && in2 && grad_in1 && grad_in2, "dpv_correlation_backward: null pointer");
  DPV_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, "dpv_correlation_backward: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
  // correlation_cuda_kernel.cu:130-131,486: the reference launches one block per INPUT pixel but places it at
  // blockIdx * stride1, so its backward is only defined for stride1 = 1 (it writes out of bounds otherwise)
  DPV_REQUIRE(stride1 == 1, "dpv_correlation_backward: stride1 = %d (the reference's backward is defined for 1)", stride1);
  int oc, oh, ow;
  const int rc = dpv_correlation_shape(H, W, pad_size, kernel_size, max_displacement, stride1, stride2, &oc, &oh, &ow);
  if (rc != DPV_OK) return rc;
  const int krad = (kernel_size - 1) / 2;
  DPV_REQUIRE((oh - 1) * stride1 + 2 * max_displacement + krad < H + 2 * pad_size &&
                  (ow - 1) * stride1 + 2 * max_displacement + krad < W + 2 * pad_size,
              "dpv_correlation_backward: taps leave the padded frame (pad_size %d too small)", pad_size);
  const long long n = (long long)B * C * H * W;
  correlation_backward_kernel<<<aux_blocks(n), kAuxThreads, 0, (cudaStream_t)stream>>>(
      grad_out, in1, in2, B, C, H, W, pad_size, kernel_size, max_displacement, stride2, oh, ow, grad_in1, grad_in2);
  return check_launch("dpv_correlation_backward");
}

int dpv_depth_errors(const float* pred, const float* gt, int n_images, int H, int W, float* out, void* stream) {
  DPV_REQUIRE(pred && gt && out, "dpv_depth_errors: null pointer");
  DPV_REQUIRE(n_images > 0 && H > 0 && W > 0, "dpv_depth_errors: bad shape n=%d H=%d W=%d", n_images, H, W);
  depth_errors_kernel<<<n_images, 1024, 0, (cudaStream_t)stream>>>(pred, gt, H * W, out);
  return check_launch("dpv_depth_errors");
}

int dpv_unc_rmse(const float* field_truth, const float* field_pred, int D, int W, const float* d_candi, float* out,
                 void* stream) {
  DPV_REQUIRE(field_truth && field_pred && d_candi && out, "dpv_unc_rmse: null pointer");
  DPV_REQUIRE(D > 0 && W > 0, "dpv_unc_rmse: bad shape D=%d W=%d", D, W);
  unc_rmse_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(field_truth, field_pred, D, W, d_candi, out);
  return check_launch("dpv_unc_rmse");
}

}  // extern "C"
