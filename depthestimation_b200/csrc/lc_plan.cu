// (d2) planner input preparation and (d3) the dynamic-programming curtain planner.
//   lc/light_curtain.py:646-708 plan_default (up to the planner calls)
//   external/lcsim/lcsim/src/planning.cpp:14-126, 162-217, 255-264
#include <cmath>
#include <vector>

#include "lc_device.cuh"

namespace dpv {

constexpr int kPrepCols = 32;
constexpr int kPrepWarps = 4;
constexpr int kWarpWarps = 16;      // plan_prep_warp_kernel: one row per warp and pass (short dependent chains)

// ---------------------------------------------------------------------------------------------
// P1 (column-local): side fix -> NaN columns -> bilinear row upsample D->R -> 5-tap depth mean ->
// per-column min-max normalise -> NaN->eps.   light_curtain.py:654-681, 79-85, 207-218
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPrepCols* kPrepWarps)
plan_prep_columns_kernel(const float* __restrict__ uf, int D, int W, int R, int flags, float* __restrict__ field_pre) {
  extern __shared__ float sm[];
  float* u = sm;                        // [D][32]
  float* f0 = sm + (size_t)D * kPrepCols;        // [R][32]
  float* f1 = f0 + (size_t)R * kPrepCols;        // [R][32]
  __shared__ int has_nan[kPrepCols];
  __shared__ int has_bad[kPrepCols + 4];        // column (with a halo of 2) holds a NaN or an infinity after :658-662
  __shared__ float lo_s[kPrepWarps][kPrepCols], hi_s[kPrepWarps][kPrepCols];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int x = blockIdx.x * kPrepCols + lane;
  const bool xin = x < W;
  // field[:,0] = field[:,1]; field[:,-1] = field[:,-2]   (:654-655)
  int sx = x;
  if (x == 0) sx = 1;
  if (x == W - 1) sx = W - 2;
  if (W == 2) sx = 1;
  if (threadIdx.x < kPrepCols) has_nan[threadIdx.x] = 0;
  if (threadIdx.x < kPrepCols + 4) has_bad[threadIdx.x] = 0;
  __syncthreads();
  bool nan_here = false, inf_here = false;
  for (int d = warp; d < D; d += kPrepWarps) {
    float v = xin ? uf[((size_t)b * D + d) * W + sx] : 0.f;
    u[d * kPrepCols + lane] = v;
    nan_here |= (v != v);
    inf_here |= (fabsf(v) == INFINITY);
  }
  if (nan_here) has_nan[lane] = 1;
  if (inf_here || (nan_here && !(flags & 1))) has_bad[lane + 2] = 1;
  if (lane < 4) {                               // the two columns either side of this block's 32
    const int hx = blockIdx.x * kPrepCols + (lane < 2 ? lane - 2 : kPrepCols + lane - 2);
    if (hx >= 0 && hx < W) {
      int hsx = hx;
      if (hx == 0) hsx = 1;
      if (hx == W - 1) hsx = W - 2;
      if (W == 2) hsx = 1;
      bool bad = false;
      for (int d = warp; d < D; d += kPrepWarps) {
        const float v = uf[((size_t)b * D + d) * W + hsx];
        bad |= (fabsf(v) == INFINITY) || ((v != v) && !(flags & 1));
      }
      if (bad) has_bad[lane < 2 ? lane : kPrepCols + lane] = 1;
    }
  }
  __syncthreads();
  if ((flags & 1) && has_nan[lane]) {                           // :658-662 (plan_default), :517-521 (plan_m1)
    for (int d = warp; d < D; d += kPrepWarps) u[d * kPrepCols + lane] = (d == D - 1) ? 1.0f : eps_floor();
  }
  __syncthreads();
  // F.interpolate(size=[R,W], bilinear, align_corners=False): rows only (W keeps its size)
  const float scale = (float)D / (float)R;
  for (int r = warp; r < R; r += kPrepWarps) {
    float s = fmaxf(scale * ((float)r + 0.5f) - 0.5f, 0.f);
    int i0 = (int)s;
    int i1 = i0 + (i0 < D - 1 ? 1 : 0);
    float l1 = s - (float)i0, l0 = 1.f - l1;
    f0[r * kPrepCols + lane] = l0 * u[i0 * kPrepCols + lane] + l1 * u[i1 * kPrepCols + lane];
  }
  __syncthreads();
  // conv2d with a 5x5 kernel whose middle COLUMN is 1/5 => 5-tap mean along the range axis, zero pad
  float lo = INFINITY, hi = -INFINITY;
  bool any_nan = false;
  for (int r = warp; r < R; r += kPrepWarps) {
    float s = 0.f;
#pragma unroll
    for (int j = -2; j <= 2; j++) {
      int rr = r + j;
      if (rr >= 0 && rr < R) s += 0.2f * f0[rr * kPrepCols + lane];
    }
    // the reference's kernel is 5x5 with zeros outside its middle column, and 0 * (NaN or inf) = NaN: a column
    // within two of a non-finite one turns NaN as a whole
    if (has_bad[lane] | has_bad[lane + 1] | has_bad[lane + 3] | has_bad[lane + 4]) s = NAN;
    f1[r * kPrepCols + lane] = s;
    lo = fminf(lo, s);
    hi = fmaxf(hi, s);
    any_nan |= (s != s);
  }
  if (any_nan) lo = hi = NAN;                                    // torch.min/max propagate NaN
  lo_s[warp][lane] = lo;
  hi_s[warp][lane] = hi;
  __syncthreads();
  lo = lo_s[0][lane];
  hi = hi_s[0][lane];
#pragma unroll
  for (int k = 1; k < kPrepWarps; k++) {
    float a = lo_s[k][lane], c = hi_s[k][lane];
    lo = (a != a || lo != lo) ? NAN : fminf(lo, a);
    hi = (c != c || hi != hi) ? NAN : fmaxf(hi, c);
  }
  const float span = hi - lo;
  for (int r = warp; r < R; r += kPrepWarps) {
    float v = f1[r * kPrepCols + lane];
    if (flags & 2) v = (v - lo) / span;                           // normalize :207-210
    if ((flags & 4) && v != v) v = eps_floor();                   // :681
    if (xin) field_pre[((size_t)b * R + r) * W + x] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// P1 for an LC camera whose width differs from the field's (FieldWarp.preprocess :79-85 resizes BOTH axes):
// side fix -> NaN columns -> bilinear [D,Win] -> [R,Wout] -> 5-tap depth mean.  Normalisation comes after the
// RGB->LC warp (:676-681), in field_normalize_kernel.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int side_fixed(int c, int W) {          // field[:,0]=field[:,1]; field[:,-1]=field[:,-2]
  if (c == 0 || W == 2) return 1;
  if (c == W - 1) return W - 2;
  return c;
}

__device__ __forceinline__ void resize_taps(int x, float scale, int Win, int& x0, int& x1, float& l0, float& l1) {
  const float s = fmaxf(scale * ((float)x + 0.5f) - 0.5f, 0.f);    // ATen area_pixel_compute_source_index
  x0 = min((int)s, Win - 1);
  x1 = x0 + (x0 < Win - 1 ? 1 : 0);
  if (scale == 1.f) x1 = x0;                     // ATen leaves an axis that keeps its size alone (no 0 * NaN from x+1)
  l1 = s - (float)x0;
  l0 = 1.f - l1;
}

__global__ void __launch_bounds__(kPrepCols* kPrepWarps)
plan_prep_resize_kernel(const float* __restrict__ uf, int D, int Win, int R, int Wout, int flags,
                        float* __restrict__ out) {
  extern __shared__ float sm[];
  float* u = sm;                                 // [D][32] x-interpolated columns
  float* f0 = sm + (size_t)D * kPrepCols;        // [R][32]
  __shared__ int nan0[kPrepCols], nan1[kPrepCols];
  __shared__ int has_bad[kPrepCols + 4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int x = blockIdx.x * kPrepCols + lane;
  const bool xin = x < Wout;
  const float* f = uf + (size_t)b * D * Win;
  const float scale_w = (float)Win / (float)Wout;
  int x0, x1;
  float lx0, lx1;
  resize_taps(xin ? x : 0, scale_w, Win, x0, x1, lx0, lx1);
  const int c0 = side_fixed(x0, Win), c1 = side_fixed(x1, Win);
  if (threadIdx.x < kPrepCols) nan0[threadIdx.x] = nan1[threadIdx.x] = 0;
  if (threadIdx.x < kPrepCols + 4) has_bad[threadIdx.x] = 0;
  __syncthreads();
  bool n0 = false, n1 = false, i0 = false, i1 = false;
  for (int d = warp; d < D; d += kPrepWarps) {
    const float a = f[(size_t)d * Win + c0], c = f[(size_t)d * Win + c1];
    n0 |= (a != a);
    n1 |= (c != c);
    i0 |= (fabsf(a) == INFINITY);
    i1 |= (fabsf(c) == INFINITY);
  }
  if (n0) nan0[lane] = 1;
  if (n1) nan1[lane] = 1;
  if (xin && (i0 || i1 || ((n0 || n1) && !(flags & 1)))) has_bad[lane + 2] = 1;
  if (lane < 4) {                                // the two output columns either side of this block's 32
    const int hx = blockIdx.x * kPrepCols + (lane < 2 ? lane - 2 : kPrepCols + lane - 2);
    if (hx >= 0 && hx < Wout) {
      int h0, h1;
      float t0, t1;
      resize_taps(hx, scale_w, Win, h0, h1, t0, t1);
      h0 = side_fixed(h0, Win);
      h1 = side_fixed(h1, Win);
      bool bad = false;
      for (int d = warp; d < D; d += kPrepWarps) {
        const float a = f[(size_t)d * Win + h0], c = f[(size_t)d * Win + h1];
        bad |= (fabsf(a) == INFINITY) || (fabsf(c) == INFINITY) || ((a != a || c != c) && !(flags & 1));
      }
      if (bad) has_bad[lane < 2 ? lane : kPrepCols + lane] = 1;
    }
  }
  __syncthreads();
  const bool r0 = (flags & 1) && nan0[lane], r1 = (flags & 1) && nan1[lane];        // :658-662
  for (int d = warp; d < D; d += kPrepWarps) {
    const float fill = (d == D - 1) ? 1.0f : eps_floor();
    const float a = r0 ? fill : f[(size_t)d * Win + c0];
    const float c = r1 ? fill : f[(size_t)d * Win + c1];
    u[d * kPrepCols + lane] = lx0 * a + lx1 * c;
  }
  __syncthreads();
  const float scale = (float)D / (float)R;
  for (int r = warp; r < R; r += kPrepWarps) {
    float s = fmaxf(scale * ((float)r + 0.5f) - 0.5f, 0.f);
    int j0 = (int)s;
    int j1 = j0 + (j0 < D - 1 ? 1 : 0);
    float l1 = s - (float)j0, l0 = 1.f - l1;
    f0[r * kPrepCols + lane] = l0 * u[j0 * kPrepCols + lane] + l1 * u[j1 * kPrepCols + lane];
  }
  __syncthreads();
  const bool spread = has_bad[lane] | has_bad[lane + 1] | has_bad[lane + 3] | has_bad[lane + 4];
  for (int r = warp; r < R; r += kPrepWarps) {
    float s = 0.f;
#pragma unroll
    for (int j = -2; j <= 2; j++) {
      int rr = r + j;
      if (rr >= 0 && rr < R) s += 0.2f * f0[rr * kPrepCols + lane];
    }
    if (spread) s = NAN;                         // same 5x5-kernel spread as plan_prep_columns_kernel
    if (xin) out[((size_t)b * R + r) * Wout + x] = s;
  }
}

// normalize (:207-210) and NaN -> eps (:681) of an [R,W] field, one column per lane.
__global__ void __launch_bounds__(kPrepCols* kPrepWarps)
field_normalize_kernel(const float* __restrict__ in, int W, int R, int flags, float* __restrict__ out) {
  __shared__ float lo_s[kPrepWarps][kPrepCols], hi_s[kPrepWarps][kPrepCols];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int x = blockIdx.x * kPrepCols + lane;
  const bool xin = x < W;
  const float* f = in + (size_t)b * R * W;
  float lo = INFINITY, hi = -INFINITY;
  bool any_nan = false;
  for (int r = warp; r < R; r += kPrepWarps) {
    const float s = xin ? f[(size_t)r * W + x] : 0.f;
    lo = fminf(lo, s);
    hi = fmaxf(hi, s);
    any_nan |= (s != s);
  }
  if (any_nan) lo = hi = NAN;
  lo_s[warp][lane] = lo;
  hi_s[warp][lane] = hi;
  __syncthreads();
  lo = lo_s[0][lane];
  hi = hi_s[0][lane];
#pragma unroll
  for (int k = 1; k < kPrepWarps; k++) {
    float a = lo_s[k][lane], c = hi_s[k][lane];
    lo = (a != a || lo != lo) ? NAN : fminf(lo, a);
    hi = (c != c || hi != hi) ? NAN : fmaxf(hi, c);
  }
  const float span = hi - lo;
  if (!xin) return;
  for (int r = warp; r < R; r += kPrepWarps) {
    float v = f[(size_t)r * W + x];
    if (flags & 2) v = (v - lo) / span;
    if ((flags & 4) && v != v) v = eps_floor();
    out[((size_t)b * R + r) * W + x] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// P2: polar z->range warp (FieldWarp.warp :47-58, bilinear grid_sample, zeros, align_corners=False)
// -> per-column argmax (:698) -> left / right peak split (:700-703) -> invert(x,p) (:221-225).
// ---------------------------------------------------------------------------------------------
struct InvertSteps {
  int n;
  float p[8];
  float den[8];   // (float)(p^p (1-p)^(1-p)) evaluated in double like the Python expression
};

__device__ __forceinline__ float pow_like_torch(float x, float e) {
  if (e == 0.5f) return sqrtf(x);          // ATen pow(tensor, scalar) specialisations
  if (e == 1.0f) return x;
  if (e == 2.0f) return x * x;
  if (e == 0.0f) return 1.f;
  return powf(x, e);
}

__global__ void __launch_bounds__(kPrepCols* kWarpWarps)
plan_prep_warp_kernel(const float* __restrict__ field_pre, const float* __restrict__ flow, int W, int R,
                      InvertSteps steps, float* __restrict__ umaps, int32_t* __restrict__ argmax_out) {
  extern __shared__ float sm[];
  float* fr = sm;                      // [R][32] warped field
  __shared__ int amax[kPrepCols];
  __shared__ float pbest[kWarpWarps][kPrepCols];
  __shared__ int pidx[kWarpWarps][kPrepCols];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int x = blockIdx.x * kPrepCols + lane;
  const bool xin = x < W;
  const float* f = field_pre + (size_t)b * R * W;
  const float ystep = (float)(2.0 / (double)(R - 1)), xstep = (float)(2.0 / (double)(W - 1));
  for (int r = warp; r < R; r += kWarpWarps) {
    float val = 0.f;
    if (xin) {
      const float fx = flow[((size_t)r * W + x) * 2 + 0], fy = flow[((size_t)r * W + x) * 2 + 1];
      float gx = __fsub_rn(__fadd_rn(-1.f, __fmul_rn((float)x, xstep)), __fmul_rn(fx, xstep));
      float gy = __fsub_rn(__fadd_rn(-1.f, __fmul_rn((float)r, ystep)), __fmul_rn(fy, ystep));
      float ix = gs_unnormalize(gx, W), iy = gs_unnormalize(gy, R);
      if (ix > -2.f && ix < (float)(W + 1) && iy > -2.f && iy < (float)(R + 1)) {
        float fx0 = floorf(ix), fy0 = floorf(iy);
        int x0 = (int)fx0, y0 = (int)fy0;
        float tx = ix - fx0, ty = iy - fy0;
        float ux = (fx0 + 1.f) - ix, uy = (fy0 + 1.f) - iy;
        bool vx0 = x0 >= 0 && x0 < W, vx1 = x0 + 1 >= 0 && x0 + 1 < W;
        bool vy0 = y0 >= 0 && y0 < R, vy1 = y0 + 1 >= 0 && y0 + 1 < R;
        if (vx0 && vy0) val += f[(size_t)y0 * W + x0] * (ux * uy);
        if (vx1 && vy0) val += f[(size_t)y0 * W + x0 + 1] * (tx * uy);
        if (vx0 && vy1) val += f[(size_t)(y0 + 1) * W + x0] * (ux * ty);
        if (vx1 && vy1) val += f[(size_t)(y0 + 1) * W + x0 + 1] * (tx * ty);
      }
    }
    fr[r * kPrepCols + lane] = val;
  }
  __syncthreads();
  // torch.max(dim=0): first index of the maximum (a NaN wins, the first NaN if several).  Every warp reduces its
  // rows (ascending), warp 0 merges the partial results with the index as tie-break.
  {
    float best = -INFINITY;
    int bi = 0x7fffffff;
    bool have = false;
    for (int r = warp; r < R; r += kWarpWarps) {
      const float v = fr[r * kPrepCols + lane];
      if (!have || v > best || (v != v && best == best)) { best = v; bi = r; have = true; }
    }
    pbest[warp][lane] = best;
    pidx[warp][lane] = have ? bi : 0x7fffffff;
  }
  __syncthreads();
  if (warp == 0) {
    float best = pbest[0][lane];
    int bi = pidx[0][lane];
    for (int k = 1; k < kWarpWarps; k++) {
      const float v = pbest[k][lane];
      const int vi = pidx[k][lane];
      if (vi == 0x7fffffff) continue;
      const bool vnan = v != v, bnan = best != best;
      bool take;
      if (vnan || bnan) take = vnan && (!bnan || vi < bi);        // NaN beats numbers; earliest NaN
      else take = v > best || (v == best && vi < bi);
      if (take) { best = v; bi = vi; }
    }
    amax[lane] = bi;
    if (xin && argmax_out) argmax_out[(size_t)b * W + x] = bi;
  }
  __syncthreads();
  if (!xin) return;
  const int pk = amax[lane];
  const int nmaps = 1 + 2 * steps.n;
  float* out = umaps + (size_t)b * nmaps * R * W;
  // invert(x, p) = x^p (1 - x)^(1 - p) / (p^p (1 - p)^(1 - p))  (light_curtain.py:221-225).  Each row is 1 on one side
  // of the peak (left_field[rows < idx] = 1, right_field[rows >= idx] = 1) and the field on the other, so one of the
  // two inverted values per row is invert(1, p) -- the same expression on constants, taken once per step -- and only
  // the other needs the two powers.
  for (int r = warp; r < R; r += kWarpWarps) out[(size_t)r * W + x] = fr[r * kPrepCols + lane];
  for (int s = 0; s < steps.n; s++) {
    const float p = steps.p[s], q = 1.f - p, den = steps.den[s];
    const float inv_one = (pow_like_torch(1.f, p) * pow_like_torch(1.f - 1.f, q)) / den;
    float* outl = out + (size_t)(1 + 2 * s) * R * W + x;
    float* outr = out + (size_t)(2 + 2 * s) * R * W + x;
    for (int r = warp; r < R; r += kWarpWarps) {
      const float v = fr[r * kPrepCols + lane];
      const float iv = (pow_like_torch(v, p) * pow_like_torch(1.f - v, q)) / den;
      const bool left_is_one = r < pk;            // then right = v; else left = v, right = 1
      outl[(size_t)r * W] = left_is_one ? inv_one : iv;
      outr[(size_t)r * W] = left_is_one ? iv : inv_one;
    }
  }
}

}  // namespace dpv

using namespace dpv;

extern "C" int lc_plan_prepare(const float* uf, int B, int D, int W, int R, const float* flow, const float* steps,
                               int nsteps, float* umaps, int32_t* argmax, float* field_pre, void* stream) {
  DPV_REQUIRE(uf && flow && umaps && field_pre, "lc_plan_prepare: null pointer (field_pre scratch is required)");
  DPV_REQUIRE(B > 0 && D > 1 && W >= 4 && R > 1, "lc_plan_prepare: bad shape");
  DPV_REQUIRE(nsteps >= 0 && nsteps <= 8 && (nsteps == 0 || steps), "lc_plan_prepare: nsteps=%d not in [0,8]", nsteps);
  DPV_REQUIRE(B <= 65535, "lc_plan_prepare: B too large");
  cudaStream_t st = (cudaStream_t)stream;
  InvertSteps is;
  is.n = nsteps;
  for (int i = 0; i < nsteps; i++) {
    double p = (double)steps[i];
    is.p[i] = (float)p;
    is.den[i] = (float)(std::pow(p, p) * std::pow(1.0 - p, 1.0 - p));
  }
  dim3 grid((W + kPrepCols - 1) / kPrepCols, B);
  size_t sm1 = (size_t)(D + 2 * R) * kPrepCols * sizeof(float);
  size_t sm2 = (size_t)R * kPrepCols * sizeof(float);
  DPV_REQUIRE(sm1 <= 200 * 1024, "lc_plan_prepare: D + 2R too large for shared memory");
  // static shared memory of the kernels counts against the 48 KB default too: opt in with some margin
  if (sm1 > 40 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(plan_prep_columns_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
  if (sm2 > 40 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(plan_prep_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
  launch_high_priority(plan_prep_columns_kernel, grid, dim3(kPrepCols * kPrepWarps), sm1, st, uf, D, W, R, 1 | 2 | 4,
                       field_pre);
  int rc = check_launch("lc_plan_prepare/columns");
  if (rc) return rc;
  launch_high_priority(plan_prep_warp_kernel, grid, dim3(kPrepCols * kWarpWarps), sm2, st, field_pre, flow, W, R, is,
                       umaps, argmax);
  return check_launch("lc_plan_prepare/warp");
}

// The planning field alone, for the planners that do not run the DP on the three default maps
// (plan_sweep / plan_empty light_curtain.py:396-503, plan_m1 :505-552): side fix -> [NaN columns] -> row upsample ->
// 5-tap mean -> [normalise] -> [NaN -> eps] -> polar warp.  flags = LC_FIELD_*.
extern "C" int lc_plan_field(const float* uf, int B, int D, int W, int R, const float* flow, int flags,
                             float* field_range, float* field_pre, void* stream) {
  DPV_REQUIRE(uf && flow && field_range && field_pre, "lc_plan_field: null pointer (field_pre scratch is required)");
  DPV_REQUIRE(B > 0 && D > 1 && W >= 4 && R > 1, "lc_plan_field: bad shape");
  DPV_REQUIRE((flags & ~7) == 0 && B <= 65535, "lc_plan_field: unknown flags %d or B too large", flags);
  cudaStream_t st = (cudaStream_t)stream;
  InvertSteps is;
  is.n = 0;
  dim3 grid((W + kPrepCols - 1) / kPrepCols, B);
  size_t sm1 = (size_t)(D + 2 * R) * kPrepCols * sizeof(float);
  size_t sm2 = (size_t)R * kPrepCols * sizeof(float);
  DPV_REQUIRE(sm1 <= 200 * 1024, "lc_plan_field: D + 2R too large for shared memory");
  if (sm1 > 40 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(plan_prep_columns_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
  if (sm2 > 40 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(plan_prep_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
  plan_prep_columns_kernel<<<grid, kPrepCols * kPrepWarps, sm1, st>>>(uf, D, W, R, flags, field_pre);
  int rc = check_launch("lc_plan_field/columns");
  if (rc) return rc;
  plan_prep_warp_kernel<<<grid, kPrepCols * kWarpWarps, sm2, st>>>(field_pre, flow, W, R, is, field_range, nullptr);
  return check_launch("lc_plan_field/warp");
}

// plan_default / plan_sweep input preparation when the LC camera is posed or sized differently from the RGB camera
// (light_curtain.py:665-685 with the :675-676 branch taken): resize [D,Win] -> [R,Wout] + 5-tap mean ->
// FieldWarp.transformZTheta warp by flow_T (:118-155; skipped when flow_T is null: rTc = I, sizes differ) ->
// [normalise] -> [NaN -> eps] -> polar warp by flow -> split / invert.  nsteps = 0 gives the polar field alone
// (umaps [B,1,R,Wout]).  scratch: 2 * B * R * Wout floats.
extern "C" int lc_plan_prepare_transformed(const float* uf, int B, int D, int Win, int R, int Wout,
                                           const float* flow_T, const float* flow, const float* steps, int nsteps,
                                           int flags, float* umaps, int32_t* argmax, float* field_pre,
                                           float* scratch, void* stream) {
  DPV_REQUIRE(uf && flow && umaps && field_pre && scratch, "lc_plan_prepare_transformed: null pointer");
  DPV_REQUIRE(B > 0 && D > 1 && Win >= 4 && Wout >= 4 && R > 1, "lc_plan_prepare_transformed: bad shape");
  DPV_REQUIRE(nsteps >= 0 && nsteps <= 8 && (nsteps == 0 || steps), "lc_plan_prepare_transformed: nsteps=%d not in [0,8]",
              nsteps);
  DPV_REQUIRE((flags & ~7) == 0 && B <= 65535, "lc_plan_prepare_transformed: unknown flags %d or B too large", flags);
  cudaStream_t st = (cudaStream_t)stream;
  InvertSteps none, is;
  none.n = 0;
  is.n = nsteps;
  for (int i = 0; i < nsteps; i++) {
    double p = (double)steps[i];
    is.p[i] = (float)p;
    is.den[i] = (float)(std::pow(p, p) * std::pow(1.0 - p, 1.0 - p));
  }
  dim3 grid((Wout + kPrepCols - 1) / kPrepCols, B);
  size_t sm1 = (size_t)(D + R) * kPrepCols * sizeof(float);
  size_t sm2 = (size_t)R * kPrepCols * sizeof(float);
  DPV_REQUIRE(sm1 <= 200 * 1024, "lc_plan_prepare_transformed: D + R too large for shared memory");
  if (sm1 > 40 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(plan_prep_resize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
  if (sm2 > 40 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(plan_prep_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
  float* s0 = scratch;
  float* s1 = scratch + (size_t)B * R * Wout;
  plan_prep_resize_kernel<<<grid, kPrepCols * kPrepWarps, sm1, st>>>(uf, D, Win, R, Wout, flags, s0);
  int rc = check_launch("lc_plan_prepare_transformed/resize");
  if (rc) return rc;
  const float* before_norm = s0;
  if (flow_T) {
    plan_prep_warp_kernel<<<grid, kPrepCols * kWarpWarps, sm2, st>>>(s0, flow_T, Wout, R, none, s1, nullptr);
    rc = check_launch("lc_plan_prepare_transformed/transform");
    if (rc) return rc;
    before_norm = s1;
  }
  field_normalize_kernel<<<grid, kPrepCols * kPrepWarps, 0, st>>>(before_norm, Wout, R, flags, field_pre);
  rc = check_launch("lc_plan_prepare_transformed/normalize");
  if (rc) return rc;
  plan_prep_warp_kernel<<<grid, kPrepCols * kWarpWarps, sm2, st>>>(field_pre, flow, Wout, R, is, umaps,
                                                                    nsteps ? argmax : nullptr);
  return check_launch("lc_plan_prepare_transformed/warp");
}
