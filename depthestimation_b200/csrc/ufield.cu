// (d1) DPV -> top-down uncertainty field (UF).  utils/img_utils.py:475-565 gen_ufield.
//
// The reference makes ~6 full passes over the 50 MB volume (2 nearest grid_samples, 2 expectations,
// exp * mask, sum).  Here the row-shifted grid_samples collapse into integer index functions
// (mask_back(y,x) is a pure function of mean[y',x'] at one source pixel), the
// expectation comes from the kernel that last wrote the DPV, and only the pixels inside the road band are
// read from the volume.  Three small grids instead of one long-running block per 32 columns (which was bound
// by serial DRAM latency: 113 us for 19 MB of traffic):
//   colmin (quash modes only)  per-column minimum of the masked depth
//   mask                       forward count per column, inverse-shifted band mask as one 32-bit word per
//                              (row, 32-column group), depth_zero
//   sum                        one warp per (depth plane, 32 columns): walks the non-empty rows in ascending
//                              order, eight loads in flight, and divides by the column count.
#include <mutex>

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
// (`step` = grid_step(n): a double division, taken once per thread, not once per pixel)
__device__ __forceinline__ float grid_step(int n) { return (float)(2.0 / (double)(n - 1)); }
__device__ __forceinline__ int nearest_src(int i, int n, float flow, float step) {
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
  float step_h, step_w;  // grid_step(H), grid_step(W)
};

// the column a shifted sample of column `col` reads (a function of the column only: callers hoist it out of their row loop)
__device__ __forceinline__ int shifted_col(const UfCtx& c, int col) {
  return c.pshift != 0 ? nearest_src(col, c.W, 0.f, c.step_w) : col;
}

// depth_s at (r,c): expectation of the row-shifted DPV (img_utils.py:509-514); cs = shifted_col(c, col)
__device__ __forceinline__ float shifted_depth(const UfCtx& c, int r, int cs, float* ms) {
  int rs = r;
  if (c.pshift != 0) rs = nearest_src(r, c.H, (float)c.pshift, c.step_h);
  if (rs < 0 || cs < 0) {
    *ms = 0.f;
    return c.fill;
  }
  *ms = c.mask ? c.mask[(size_t)rs * c.W + cs] : 1.f;
  return c.mean[(size_t)rs * c.W + cs];
}

// forward band mask at (r,c) before quash (img_utils.py:516-530); cs = shifted_col(c, c_)
__device__ __forceinline__ float band_mask(const UfCtx& c, int r, int cs, float* depth_s) {
  float ms;
  float z = shifted_depth(c, r, cs, &ms);
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
constexpr int kUfRows = 64;          // rows per block in the per-pixel kernels

__device__ __forceinline__ UfCtx make_ctx(const UfArgs& a, int b, float fill) {
  UfCtx c;
  c.mean = a.mean + (size_t)b * a.H * a.W;
  c.mask = a.mask ? a.mask + (size_t)b * a.H * a.W : nullptr;
  c.H = a.H; c.W = a.W; c.pshift = a.prm.pshift;
  const float* K = a.K_up + b * 9;
  c.fy = K[4]; c.cy = K[5];
  c.ystart = a.prm.ystart; c.yend = a.prm.yend;
  c.maxd1 = a.prm.maxd - 1.f; c.mind = a.prm.mind;
  c.fill = fill;
  c.step_h = grid_step(a.H);
  c.step_w = grid_step(a.W);
  return c;
}

// expectation of a zero-filled log-DPV (p = 1 in every bin): what an out-of-bounds shifted pixel sees.
// A sequential sum over D (the reference's order); one thread of the CTA takes it, the others read it after the
// CTA's first barrier (it was 2 x D instructions in every thread, for eight pixels each).
__device__ __forceinline__ float zero_fill_depth(const UfArgs& a) {
  float s = 0.f;
  if (a.prm.bv_log)
    for (int d = 0; d < a.D; d++) s += a.d_candi[d];
  return s;
}

// masked depth with empty pixels pushed to 1000 (img_utils.py:533-535)
__device__ __forceinline__ float quash_key(float z, float m) {
  const float cl = z * m;
  return cl == 0.f ? 1000.f : cl;
}

// grid (column groups, row blocks, B): cmin[b,x] = min_y quash_key   (img_utils.py:532-536)
__global__ void __launch_bounds__(kUfCols* kUfWarps)
ufield_colmin_kernel(UfArgs a, float* __restrict__ cmin) {
  __shared__ float s_min[kUfCols];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, b = blockIdx.z;
  const int x = blockIdx.x * kUfCols + lane;
  __shared__ float s_fill;
  if (threadIdx.x < kUfCols) s_min[threadIdx.x] = 3.0e38f;
  if (threadIdx.x == 0) s_fill = zero_fill_depth(a);
  __syncthreads();
  const UfCtx c = make_ctx(a, b, s_fill);
  if (x < a.W) {
    float mx = 3.0e38f;
    const int cs = shifted_col(c, x);
    for (int y = blockIdx.y * kUfRows + warp; y < min(a.H, (int)(blockIdx.y + 1) * kUfRows); y += kUfWarps) {
      float z;
      const float m = band_mask(c, y, cs, &z);
      mx = fminf(mx, quash_key(z, m));
    }
    atomic_min_pos(&s_min[lane], mx);
  }
  __syncthreads();
  if (threadIdx.x < kUfCols && x < a.W) atomic_min_pos(&cmin[(size_t)b * a.W + x], s_min[lane]);
}

// grid (column groups, row blocks, B): forward count, mask_back words, depth_zero
__global__ void __launch_bounds__(kUfCols* kUfWarps)
ufield_mask_kernel(UfArgs a, const float* __restrict__ cmin, unsigned* __restrict__ words, int* __restrict__ count) {
  __shared__ int s_cnt[kUfCols];
  const int H = a.H, W = a.W;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, b = blockIdx.z;
  const int x = blockIdx.x * kUfCols + lane;
  const bool xin = x < W;
  __shared__ float s_fill;
  if (threadIdx.x < kUfCols) s_cnt[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_fill = zero_fill_depth(a);
  __syncthreads();
  const UfCtx c = make_ctx(a, b, s_fill);
  // column that mask_back(.,x) looks at (inverse shift samples the forward mask at (ri[y], cs[x]))
  const int xc = !xin ? -1 : (c.pshift != 0 ? nearest_src(x, W, 0.f, c.step_w) : x);
  // ... and the columns the forward masks at x and at xc read in turn (column-only: out of the row loop)
  const int cs_x = xin ? shifted_col(c, x) : -1, cs_xc = xc >= 0 ? shifted_col(c, xc) : -1;
  float qx = 0.f, qc = 0.f;
  if (a.prm.quash) {
    if (xin) qx = cmin[(size_t)b * W + x];
    if (xc >= 0) qc = cmin[(size_t)b * W + xc];
  }
  int cnt = 0;
  for (int y = blockIdx.y * kUfRows + warp; y < min(H, (int)(blockIdx.y + 1) * kUfRows); y += kUfWarps) {
    float mback = 0.f;
    if (xin) {
      float z;
      float m = band_mask(c, y, cs_x, &z);
      if (a.prm.quash) {                                          // img_utils.py:537-539
        const float cl = quash_key(z, m);
        m = m * ((cl > qx - 1.0f && cl < qx + 1.0f) ? 1.f : 0.f);
      }
      cnt += (m != 0.f);
      // img_utils.py:541-545: mask_back = grid_sample(mask, inverse shift, nearest)
      int ry = y;
      if (c.pshift != 0) ry = nearest_src(y, H, -(float)c.pshift, c.step_h);
      if (ry >= 0 && xc >= 0) {
        float zz;
        float m2 = band_mask(c, ry, cs_xc, &zz);
        if (a.prm.quash) {
          const float cl = quash_key(zz, m2);
          m2 = m2 * ((cl > qc - 1.0f && cl < qc + 1.0f) ? 1.f : 0.f);
        }
        mback = m2;
      }
      if (a.depth_zero) a.depth_zero[((size_t)b * H + y) * W + x] = c.mean[(size_t)y * W + x] * mback;
    }
    // mask values are 0/1 (a fractional user mask only switches pixels on or off, as before)
    const unsigned any = __ballot_sync(0xffffffffu, mback != 0.f);
    if (lane == 0) words[((size_t)b * gridDim.x + blockIdx.x) * H + y] = any;
  }
  if (cnt) atomicAdd(&s_cnt[lane], cnt);
  __syncthreads();
  if (threadIdx.x < kUfCols && xin && s_cnt[lane]) atomicAdd(&count[(size_t)b * W + x], s_cnt[lane]);
}

// grid (column groups, ceil(D / warps), B): UF[d,x] = sum_y p[d,y,x] * mask_back[y,x] / count[x]  (:551-557)
__global__ void __launch_bounds__(kUfCols* kUfWarps)
ufield_sum_kernel(UfArgs a, const unsigned* __restrict__ words, const int* __restrict__ count) {
  const int H = a.H, W = a.W, D = a.D;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, b = blockIdx.z;
  const int d = blockIdx.y * kUfWarps + warp;
  const int x = blockIdx.x * kUfCols + lane;
  if (d >= D) return;
  const unsigned* wrow = words + ((size_t)b * gridDim.x + blockIdx.x) * H;
  const float* plane = a.dpv + ((size_t)b * D + d) * (size_t)H * W + x;
  const unsigned full = 0xffffffffu;
  float acc = 0.f;
  for (int y0 = 0; y0 < H; y0 += 32) {
    const unsigned myw = (y0 + lane < H) ? wrow[y0 + lane] : 0u;
    unsigned act = __ballot_sync(full, myw != 0u);
    while (act) {                                                // rows with a band pixel, ascending, 8 at a time
      float v[8];
      bool on[8];
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const int r = act ? (__ffs(act) - 1) : 0;
        const unsigned w = __shfl_sync(full, myw, r);
        on[k] = act != 0u && ((w >> lane) & 1u);
        act &= act - 1u;                                          // 0 stays 0
        v[k] = on[k] ? __ldg(plane + (size_t)(y0 + r) * W) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < 8; k++)
        if (on[k]) acc += a.prm.bv_log ? expf(v[k]) : v[k];
    }
  }
  if (x < W) a.uf[((size_t)b * D + d) * W + x] = acc / (float)count[(size_t)b * W + x];
}

}  // namespace dpv

using namespace dpv;

extern "C" size_t dpv_ufield_scratch_bytes(int B, int H, int W) {
  const size_t G = (size_t)(W + kUfCols - 1) / kUfCols;
  return (size_t)B * G * H * sizeof(unsigned) + 2 * (size_t)B * W * sizeof(float);
}

extern "C" int dpv_ufield(const float* dpv_, const float* mean, const float* mask, int B, int D, int H, int W,
                          const float* d_candi, const float* K_up, const DpvUfieldParams* prm, float* uf,
                          float* depth_zero, void* scratch, size_t scratch_bytes, void* stream) {
  DPV_REQUIRE(dpv_ && mean && d_candi && K_up && prm && uf && scratch, "dpv_ufield: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 1 && W > 1, "dpv_ufield: bad shape");
  DPV_REQUIRE(B <= 65535 && (D + kUfWarps - 1) / kUfWarps <= 65535, "dpv_ufield: B or D too large");
  DPV_REQUIRE(scratch_bytes >= dpv_ufield_scratch_bytes(B, H, W), "dpv_ufield: scratch too small (%zu < %zu bytes)",
              scratch_bytes, dpv_ufield_scratch_bytes(B, H, W));
  DPV_REQUIRE(((uintptr_t)scratch & 3) == 0, "dpv_ufield: scratch must be 4-byte aligned");
  UfArgs a;
  a.dpv = dpv_; a.mean = mean; a.mask = mask; a.d_candi = d_candi; a.K_up = K_up; a.uf = uf;
  a.depth_zero = depth_zero; a.D = D; a.H = H; a.W = W; a.prm = *prm;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = (W + kUfCols - 1) / kUfCols;
  unsigned* words = reinterpret_cast<unsigned*>(scratch);
  int* count = reinterpret_cast<int*>(words + (size_t)B * G * H);
  float* cmin = reinterpret_cast<float*>(count + (size_t)B * W);
  DPV_CUDA(cudaMemsetAsync(count, 0, (size_t)B * W * sizeof(int), st));
  const dim3 gpix(G, (H + kUfRows - 1) / kUfRows, B);
  int launches = 0;
  if (a.prm.quash) {
    DPV_CUDA(cudaMemsetAsync(cmin, 0x7f, (size_t)B * W * sizeof(float), st));     // 0x7f7f7f7f = 3.39e38
    launch_high_priority(ufield_colmin_kernel, gpix, dim3(kUfCols * kUfWarps), 0, st, a, cmin);
    launches++;
  }
  launch_high_priority(ufield_mask_kernel, gpix, dim3(kUfCols * kUfWarps), 0, st, a, cmin, words, count);
  launch_high_priority(ufield_sum_kernel, dim3(G, (D + kUfWarps - 1) / kUfWarps, B), dim3(kUfCols * kUfWarps), 0, st, a,
                       words, count);
  launches++;
  const int rc = check_launch("dpv_ufield");
  if (rc == DPV_OK) count_launch(launches);
  return rc;
}

// ---------------------------------------------------------------------------------------------------------------
// dpv_ufield_static: dpv_ufield for callers that may not allocate (no counterpart in the reference).
// The scratch that dpv_ufield takes from its caller lives in a file-scope device
// array instead, so neither the caller nor this function calls cudaMalloc / cudaFree or any allocating routine.
// Every GPU that loads the library holds the array until the process exits.
//
// In host code the array's NAME is a host-side stand-in, not a device address: handing it to a kernel would fault
// (or, where the GPU can reach host memory, silently use the wrong memory).  Its device address is resolved with
// cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels get.
namespace dpv {

constexpr size_t kUfStaticBytes = DPV_UFIELD_STATIC_MAX_BYTES;      // 1 MiB: C2 (8 x 256 x 768) needs 240 KiB, C5 (2 x 512 x 1536) 216 KiB
static __device__ __align__(256) unsigned char g_uf_static_scratch[kUfStaticBytes];

constexpr int kUfMaxDevices = 64;

// Device address of g_uf_static_scratch on the CURRENT device (cached per device ordinal).
static int uf_static_scratch_ptr(void** out) {
  static void* ptrs[kUfMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  DPV_CUDA(cudaGetDevice(&dev));
  DPV_REQUIRE(dev >= 0 && dev < kUfMaxDevices, "dpv_ufield_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    DPV_CUDA(cudaGetSymbolAddress(&p, g_uf_static_scratch));
    DPV_REQUIRE(p != nullptr && ((uintptr_t)p & 255) == 0, "dpv_ufield_static: bad scratch address %p", p);
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return DPV_OK;
}

}  // namespace dpv

extern "C" int dpv_ufield_static(const float* dpv_, const float* mean, const float* mask, int B, int D, int H, int W,
                                 const float* d_candi, const float* K_up, const DpvUfieldParams* prm, float* uf,
                                 float* depth_zero, void* stream) {
  DPV_REQUIRE(B > 0 && D > 0 && H > 1 && W > 1, "dpv_ufield_static: bad shape");
  const size_t need = dpv_ufield_scratch_bytes(B, H, W);
  DPV_REQUIRE(need <= kUfStaticBytes,
              "dpv_ufield_static: B=%d H=%d W=%d needs %zu bytes of scratch, the static maximum is %zu bytes "
              "(use dpv_ufield with caller scratch for larger batches)", B, H, W, need, (size_t)kUfStaticBytes);
  void* scratch = nullptr;
  const int rc = uf_static_scratch_ptr(&scratch);
  if (rc != DPV_OK) return rc;
  // dpv_ufield resets what must start at zero (per-column counts, and the column minima of the quash branch) with
  // cudaMemsetAsync on `stream` and its mask kernel rewrites every band-mask word, so nothing of the previous call
  // survives into this one.
  return dpv_ufield(dpv_, mean, mask, B, D, H, W, d_candi, K_up, prm, uf, depth_zero, scratch, kUfStaticBytes, stream);
}
