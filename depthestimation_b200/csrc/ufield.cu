// (d1) DPV -> top-down uncertainty field (UF).  utils/img_utils.py:475-565 gen_ufield.
//
// The reference makes ~6 full passes over the 50 MB volume (2 nearest grid_samples, 2 expectations,
// exp * mask, sum).  Here the row-shifted grid_samples collapse into integer index functions
// (mask_back(y,x) is a pure function of mean[y',x'] at one source pixel), the
// expectation comes from the kernel that last wrote the DPV, and only the rows of a column that are
// inside the road band are read from the volume: one block = 32 columns x all D planes.
#include "common.cuh"

namespace dpv {

struct UfArgs {
  const float* dpv;
  const float* mean;
  const float* mask;
  const float* d_candi;
  const float* K_up;
  float* uf;
  float* depth_zero;
  int D, H, W;
  DpvUfieldParams prm;
};

// convert_flowfield (img_utils.py:343-349) + grid_sample(mode='nearest', align_corners=False):
// source index for destination index i under a constant flow; -1 = out of bounds (zero fill).
__device__ __forceinline__ int nearest_src(int i, int n, float flow) {
  const float step = (float)(2.0 / (double)(n - 1));
  float g = __fsub_rn(__fadd_rn(-1.f, __fmul_rn((float)i, step)), __fmul_rn(flow, step));
  float pix = gs_unnormalize(g, n);
  float r = rintf(pix);
  if (!(r >= 0.f && r < (float)n)) return -1;
  return (int)r;
}

struct UfCtx {
  const float* mean;   // [H,W] of this frame
  const float* mask;   // nullable
  int H, W, pshift;
  float fy, cy, ystart, yend, maxd1, mind, fill;
};

// depth_s at (r,c): expectation of the row-shifted DPV (img_utils.py:509-514)
__device__ __forceinline__ float shifted_depth(const UfCtx& c, int r, int col, float* ms) {
  int rs = r, cs = col;
  if (c.pshift != 0) {
    rs = nearest_src(r, c.H, (float)c.pshift);
    cs = nearest_src(col, c.W, 0.f);
  }
  if (rs < 0 || cs < 0) {
    *ms = 0.f;
    return c.fill;
  }
  *ms = c.mask ? c.mask[(size_t)rs * c.W + cs] : 1.f;
  return c.mean[(size_t)rs * c.W + cs];
}

// forward band mask at (r,c) before quash (img_utils.py:516-530)
__device__ __forceinline__ float band_mask(const UfCtx& c, int r, int col, float* depth_s) {
  float ms;
  float z = shifted_depth(c, r, col, &ms);
  *depth_s = z;
  float yf = __fdiv_rn(__fsub_rn((float)r, c.cy), c.fy);
  float Y = __fmul_rn(yf, z);
  bool out = (Y > c.yend) || (Y < c.ystart) || (z > c.maxd1) || (z < c.mind);
  float m = out ? 0.f : 1.f;
  return c.mask ? m * ms : m;
}

__device__ __forceinline__ void atomic_min_pos(float* addr, float v) {
  atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));   // valid for non-negative floats
}

constexpr int kUfCols = 32;
constexpr int kUfWarps = 8;

__global__ void __launch_bounds__(kUfCols* kUfWarps)
ufield_kernel(UfArgs a) {
  extern __shared__ unsigned char smem_raw[];
  const int H = a.H, W = a.W, D = a.D;
  unsigned char* mb = smem_raw;                                   // [H][32] mask_back
  int* rows = reinterpret_cast<int*>(smem_raw + (size_t)H * kUfCols);   // active row list [H]
  __shared__ int count[kUfCols];
  __shared__ float cmin_x[kUfCols], cmin_c[kUfCols];
  __shared__ int nrows;
  __shared__ float fill_s;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int x = blockIdx.x * kUfCols + lane;
  const bool xin = x < W;
  if (threadIdx.x == 0) {
    nrows = 0;
    float s = 0.f;
    if (a.prm.bv_log)
      for (int d = 0; d < D; d++) s += a.d_candi[d];              // zero-filled log-DPV => p = 1 in every bin
    fill_s = s;
  }
  if (threadIdx.x < kUfCols) {
    count[threadIdx.x] = 0;
    cmin_x[threadIdx.x] = 3.0e38f;
    cmin_c[threadIdx.x] = 3.0e38f;
  }
  __syncthreads();

  UfCtx c;
  c.mean = a.mean + (size_t)b * H * W;
  c.mask = a.mask ? a.mask + (size_t)b * H * W : nullptr;
  c.H = H; c.W = W; c.pshift = a.prm.pshift;
  const float* K = a.K_up + b * 9;
  c.fy = K[4]; c.cy = K[5];
  c.ystart = a.prm.ystart; c.yend = a.prm.yend;
  c.maxd1 = a.prm.maxd - 1.f; c.mind = a.prm.mind;
  c.fill = fill_s;

  // column that mask_back(.,x) looks at (inverse shift samples the forward mask at (ri[y], cs[x]))
  const int xc = !xin ? -1 : (c.pshift != 0 ? nearest_src(x, W, 0.f) : x);

  if (a.prm.quash) {                                              // img_utils.py:532-539
    float mx = 3.0e38f, mc = 3.0e38f;
    for (int r = warp; r < H; r += kUfWarps) {
      if (xin) {
        float z;
        float m = band_mask(c, r, x, &z);
        float cl = z * m;
        if (cl == 0.f) cl = 1000.f;
        mx = fminf(mx, cl);
      }
      if (xc >= 0 && xc != x) {
        float z;
        float m = band_mask(c, r, xc, &z);
        float cl = z * m;
        if (cl == 0.f) cl = 1000.f;
        mc = fminf(mc, cl);
      }
    }
    if (xin) atomic_min_pos(&cmin_x[lane], mx);
    if (xc >= 0 && xc != x) atomic_min_pos(&cmin_c[lane], mc);
    __syncthreads();
    if (xc == x) cmin_c[lane] = cmin_x[lane];
    __syncthreads();
  }

  // phase 1: forward count of column x, mask_back of column x, list of rows with any band pixel
  int cnt = 0;
  for (int y = warp; y < H; y += kUfWarps) {
    float mback = 0.f;
    if (xin) {
      float z;
      float m = band_mask(c, y, x, &z);
      if (a.prm.quash) {
        float cl = z * m;
        if (cl == 0.f) cl = 1000.f;
        m = m * ((cl > cmin_x[lane] - 1.0f && cl < cmin_x[lane] + 1.0f) ? 1.f : 0.f);
      }
      cnt += (m != 0.f);
      // img_utils.py:541-545: mask_back = grid_sample(mask, inverse shift, nearest)
      int ry = y;
      if (c.pshift != 0) ry = nearest_src(y, H, -(float)c.pshift);
      if (ry >= 0 && xc >= 0) {
        float zz;
        float m2 = band_mask(c, ry, xc, &zz);
        if (a.prm.quash) {
          float cl = zz * m2;
          if (cl == 0.f) cl = 1000.f;
          m2 = m2 * ((cl > cmin_c[lane] - 1.0f && cl < cmin_c[lane] + 1.0f) ? 1.f : 0.f);
        }
        mback = m2;
      }
      if (a.depth_zero) a.depth_zero[((size_t)b * H + y) * W + x] = c.mean[(size_t)y * W + x] * mback;
    }
    // mask values are 0/1 unless a fractional user mask is given; keep the float for that case
    mb[y * kUfCols + lane] = (mback != 0.f) ? 1 : 0;
    unsigned any = __ballot_sync(0xffffffffu, mback != 0.f);
    if (lane == 0 && any) rows[atomicAdd(&nrows, 1)] = y;
  }
  if (cnt) atomicAdd(&count[lane], cnt);
  __syncthreads();
  // deterministic row order: sort the (short) active list ascending — simple odd-even by one warp
  const int n = nrows;
  if (warp == 0) {
    for (int pass = 0; pass < n; pass++) {
      for (int i = 2 * lane + (pass & 1); i + 1 < n; i += 64) {
        int r0 = rows[i], r1 = rows[i + 1];
        if (r0 > r1) { rows[i] = r1; rows[i + 1] = r0; }
      }
      __syncwarp();
    }
  }
  __syncthreads();

  // phase 2: UF[d,x] = sum_y p[d,y,x] * mask_back[y,x] / count[x]   (img_utils.py:551-557)
  const float denom = (float)count[lane];
  const size_t HW = (size_t)H * W;
  for (int d = warp; d < D; d += kUfWarps) {
    const float* plane = a.dpv + ((size_t)b * D + d) * HW + x;
    float acc = 0.f;
    int i = 0;
    for (; i + 4 <= n; i += 4) {
      int y0 = rows[i], y1 = rows[i + 1], y2 = rows[i + 2], y3 = rows[i + 3];
      bool m0 = xin && mb[y0 * kUfCols + lane], m1 = xin && mb[y1 * kUfCols + lane];
      bool m2 = xin && mb[y2 * kUfCols + lane], m3 = xin && mb[y3 * kUfCols + lane];
      float v0 = m0 ? __ldg(plane + (size_t)y0 * W) : 0.f;
      float v1 = m1 ? __ldg(plane + (size_t)y1 * W) : 0.f;
      float v2 = m2 ? __ldg(plane + (size_t)y2 * W) : 0.f;
      float v3 = m3 ? __ldg(plane + (size_t)y3 * W) : 0.f;
      if (a.prm.bv_log) { v0 = expf(v0); v1 = expf(v1); v2 = expf(v2); v3 = expf(v3); }
      if (m0) acc += v0;
      if (m1) acc += v1;
      if (m2) acc += v2;
      if (m3) acc += v3;
    }
    for (; i < n; i++) {
      int y0 = rows[i];
      if (xin && mb[y0 * kUfCols + lane]) {
        float v0 = __ldg(plane + (size_t)y0 * W);
        acc += a.prm.bv_log ? expf(v0) : v0;
      }
    }
    if (xin) a.uf[((size_t)b * D + d) * W + x] = acc / denom;
  }
}

}  // namespace dpv

using namespace dpv;

extern "C" int dpv_ufield(const float* dpv_, const float* mean, const float* mask, int B, int D, int H, int W,
                          const float* d_candi, const float* K_up, const DpvUfieldParams* prm, float* uf,
                          float* depth_zero, void* stream) {
  DPV_REQUIRE(dpv_ && mean && d_candi && K_up && prm && uf, "dpv_ufield: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 1 && W > 1, "dpv_ufield: bad shape");
  DPV_REQUIRE(B <= 65535, "dpv_ufield: B too large");
  UfArgs a;
  a.dpv = dpv_; a.mean = mean; a.mask = mask; a.d_candi = d_candi; a.K_up = K_up; a.uf = uf;
  a.depth_zero = depth_zero; a.D = D; a.H = H; a.W = W; a.prm = *prm;
  size_t smem = (size_t)H * kUfCols + (size_t)H * sizeof(int);
  DPV_REQUIRE(smem <= 200 * 1024, "dpv_ufield: H=%d too large", H);
  if (smem > 48 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(ufield_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((W + kUfCols - 1) / kUfCols, B);
  ufield_kernel<<<grid, kUfCols * kUfWarps, smem, (cudaStream_t)stream>>>(a);
  return check_launch("dpv_ufield");
}
