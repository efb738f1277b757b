// (b) DPV normalisation / statistics kernels and the stand-alone DPV utilities:
//   dpv_logsoftmax_stats, dpv_expectation, dpv_spread, dpv_resample_depth, dpv_bayes_fuse.
// All are HBM-streaming: one thread owns one pixel and keeps its whole depth column in registers,
// so the volume is read once and written once (warp = 32 consecutive x => each depth plane access
// is one fully-used 128 B line).
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "dpv_pixel.cuh"

namespace dpv {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

constexpr int kThreads = 128;

// dpv_softmax_tma.cu: DPV_OK / DPV_E* when it took the call, -1 when the shape is not its case
int logsoftmax_stats_tma(const float* x, int B, int D, int H, int W, const float* d_candi, float* logp, float* mean,
                         float* var, float* entropy, float* carry, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// log_softmax over D (+ optional addend) with fused expectation / variance / entropy / x0.25 carry.
// models/models.py:351,671,777 ; utils/img_utils.py:197-215 ; trainer/default_trainer.py:284,446-450
// ---------------------------------------------------------------------------------------------
// The depth column of a pixel is split over T adjacent lanes (DPT bins each, lane t owns bins
// [t*DPT, (t+1)*DPT)): 2 x 32 registers per lane instead of 2 x 64 keeps ~20 warps resident per SM, and the
// column reductions (max, sum, mean, entropy, variance) finish with log2(T) shuffles.
template <int T>
__device__ __forceinline__ float lane_group_sum(float v) {
#pragma unroll
  for (int off = T / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

template <int DPT, int T, bool FULL>
__global__ void __launch_bounds__(kThreads, DPT <= 16 ? 8 : 4)
lsm_stats_kernel(const float* __restrict__ x, const float* __restrict__ add, int D, int H, int W,
                 const float* __restrict__ d_candi, float* __restrict__ logp, float* __restrict__ mean_o,
                 float* __restrict__ var_o, float* __restrict__ ent_o, float* __restrict__ carry) {
  __shared__ float dc[DPT * T];
  for (int i = threadIdx.x; i < DPT * T; i += kThreads) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  // grid.y = sample, grid.x covers one image: the per-pixel index arithmetic stays in 32 bits.
  // Every lane of a warp stays alive for the shuffles: out-of-range pixels are clamped and not stored.
  const unsigned HW = (unsigned)H * (unsigned)W;
  const unsigned hw_raw = (blockIdx.x * kThreads + threadIdx.x) / T;
  const bool live = hw_raw < HW;
  const unsigned hw = live ? hw_raw : HW - 1;
  const int tl = threadIdx.x & (T - 1);
  const int d0 = tl * DPT;
  const float* dcl = dc + d0;
  const int b = blockIdx.y;
  const size_t pix = (size_t)b * HW + hw;
  const size_t base = ((size_t)b * D + d0) * HW + hw;

  // FULL (D == DPT*T): no per-bin bound checks.  Plane pointers are bumped by HW (one 64-bit add each).
  auto valid = [&](int d) -> bool { return FULL || d0 + d < D; };
  float t[DPT];
  {
    const float* xp = x + base;
#pragma unroll
    for (int d = 0; d < DPT; d++, xp = bump(xp, HW)) t[d] = valid(d) ? ld_stream(xp) : -INFINITY;
  }
  if (add != nullptr) {
    const float* ap = add + base;
#pragma unroll
    for (int d = 0; d < DPT; d++, ap = bump(ap, HW))
      if (valid(d)) t[d] += ld_stream(ap);
  }
  float m0 = t[0], m1 = t[1];
#pragma unroll
  for (int d = 2; d < DPT; d += 2) { m0 = fmaxf(m0, t[d]); m1 = fmaxf(m1, t[d + 1]); }
  float m = fmaxf(m0, m1);
#pragma unroll
  for (int off = T / 2; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  float e[DPT];
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d += 2) {
    t[d] -= m;
    t[d + 1] -= m;
    e[d] = exp2f_approx(t[d] * 1.4426950408889634f);       // exp(-inf) = 0 for the padding bins
    e[d + 1] = exp2f_approx(t[d + 1] * 1.4426950408889634f);
    s0 += e[d];
    s1 += e[d + 1];
  }
  const float ssum = lane_group_sum<T>(s0 + s1);
  const float lse = log2f_approx(ssum) * 0.6931471805599453f;   // ssum in [1, D]
  const float inv = 1.f / ssum;
  float mean0 = 0.f, mean1 = 0.f, ent0 = 0.f, ent1 = 0.f;
  float* lpp = logp + base;
#pragma unroll
  for (int d = 0; d < DPT; d += 2, lpp = bump(bump(lpp, HW), HW)) {
    const float lp0 = t[d] - lse, lp1 = t[d + 1] - lse;
    e[d] *= inv;
    e[d + 1] *= inv;
    if (valid(d)) {
      if (live) st_stream(lpp, lp0);
      mean0 = fmaf(dcl[d], e[d], mean0);
      ent0 = fmaf(-e[d], lp0, ent0);
    }
    if (valid(d + 1)) {
      if (live) st_stream(lpp + HW, lp1);
      mean1 = fmaf(dcl[d + 1], e[d + 1], mean1);
      ent1 = fmaf(-e[d + 1], lp1, ent1);
    }
    t[d] = lp0;
    t[d + 1] = lp1;
  }
  const float mean = lane_group_sum<T>(mean0 + mean1);
  const float ent = lane_group_sum<T>(ent0 + ent1);
  const bool writer = live && tl == 0;
  if (mean_o && writer) mean_o[pix] = mean;
  if (ent_o && writer) ent_o[pix] = ent;
  if (var_o) {
    float v0 = 0.f, v1 = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d += 2) {
      const float a0 = dcl[d] - mean, a1 = dcl[d + 1] - mean;
      v0 = fmaf(a0 * a0, e[d], v0);
      v1 = fmaf(a1 * a1, e[d + 1], v1);
    }
    const float var = lane_group_sum<T>(v0 + v1);
    if (writer) var_o[pix] = var;
  }
  if (carry && live) {
    // F.interpolate(scale_factor=0.25, 'nearest'): dst (j,i) <- src (4j, 4i)
    const int y = (int)(hw / (unsigned)W), xx = (int)(hw - (unsigned)y * (unsigned)W);
    const int h4 = H / 4, w4 = W / 4;
    if ((y & 3) == 0 && (xx & 3) == 0 && (y >> 2) < h4 && (xx >> 2) < w4) {
      float* c = carry + ((size_t)b * D + d0) * h4 * w4 + (size_t)(y >> 2) * w4 + (xx >> 2);
#pragma unroll
      for (int d = 0; d < DPT; d++)
        if (valid(d)) c[(size_t)d * h4 * w4] = t[d];
    }
  }
}

// Fallback for D > 128: three streaming passes, no register column.
__global__ void lsm_stats_generic_kernel(const float* __restrict__ x, const float* __restrict__ add, int D, int H,
                                         int W, long long npix, const float* __restrict__ d_candi,
                                         float* __restrict__ logp, float* __restrict__ mean_o,
                                         float* __restrict__ var_o, float* __restrict__ ent_o,
                                         float* __restrict__ carry) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const size_t HW = (size_t)H * W;
  const int b = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)b * HW);
  const size_t base = (size_t)b * D * HW + hw;
  float m = -INFINITY;
  for (int d = 0; d < D; d++) {
    float v = x[base + (size_t)d * HW] + (add ? add[base + (size_t)d * HW] : 0.f);
    m = fmaxf(m, v);
  }
  float s = 0.f;
  for (int d = 0; d < D; d++) {
    float v = x[base + (size_t)d * HW] + (add ? add[base + (size_t)d * HW] : 0.f);
    s += expf(v - m);
  }
  const float lse = logf(s);
  float mean = 0.f, ent = 0.f;
  const int y = (int)(hw / W), xx = (int)(hw % W);
  const int h4 = H / 4, w4 = W / 4;
  const bool do_carry = carry && (y & 3) == 0 && (xx & 3) == 0 && (y >> 2) < h4 && (xx >> 2) < w4;
  for (int d = 0; d < D; d++) {
    float v = x[base + (size_t)d * HW] + (add ? add[base + (size_t)d * HW] : 0.f);
    float lp = (v - m) - lse;
    logp[base + (size_t)d * HW] = lp;
    float p = expf(lp);
    mean += d_candi[d] * p;
    ent -= p * lp;
    if (do_carry) carry[(size_t)b * D * h4 * w4 + (size_t)d * h4 * w4 + (size_t)(y >> 2) * w4 + (xx >> 2)] = lp;
  }
  if (mean_o) mean_o[pix] = mean;
  if (ent_o) ent_o[pix] = ent;
  if (var_o) {
    float var = 0.f;
    for (int d = 0; d < D; d++) {
      float dd = d_candi[d] - mean;
      var += dd * dd * expf(logp[base + (size_t)d * HW]);
    }
    var_o[pix] = var;
  }
}

// ---------------------------------------------------------------------------------------------
template <int DPT>
__global__ void __launch_bounds__(kThreads)
expectation_kernel(const float* __restrict__ dpv, int D, size_t HW, long long npix,
                   const float* __restrict__ d_candi, int bv_log, float* __restrict__ mean_o) {
  __shared__ float dc[DPT];
  for (int i = threadIdx.x; i < DPT; i += kThreads) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const int b = (int)(pix / (long long)HW);
  const size_t base = (size_t)b * D * HW + (size_t)(pix - (long long)b * HW);
  float v[DPT];
#pragma unroll
  for (int d = 0; d < DPT; d++) v[d] = d < D ? __ldg(dpv + base + (size_t)d * HW) : 0.f;
  float mean = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) mean += dc[d] * (bv_log ? expf(v[d]) : v[d]);
  mean_o[pix] = mean;
}

__global__ void expectation_generic_kernel(const float* __restrict__ dpv, int D, size_t HW, long long npix,
                                           const float* __restrict__ d_candi, int bv_log,
                                           float* __restrict__ mean_o) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const int b = (int)(pix / (long long)HW);
  const size_t base = (size_t)b * D * HW + (size_t)(pix - (long long)b * HW);
  float mean = 0.f;
  for (int d = 0; d < D; d++) {
    float v = dpv[base + (size_t)d * HW];
    mean += d_candi[d] * (bv_log ? expf(v) : v);
  }
  mean_o[pix] = mean;
}

// ---------------------------------------------------------------------------------------------
template <int DPT>
__global__ void __launch_bounds__(kThreads)
spread_kernel(const float* __restrict__ p_in, int D, size_t HW, long long npix, int N, float* __restrict__ out) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const int b = (int)(pix / (long long)HW);
  const size_t base = (size_t)b * D * HW + (size_t)(pix - (long long)b * HW);
  float p[DPT];
#pragma unroll
  for (int d = 0; d < DPT; d++) p[d] = d < D ? ld_stream(p_in + base + (size_t)d * HW) : 0.f;
  spread_dispatch<DPT>(p, D, N);
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) st_stream(out + base + (size_t)d * HW, p[d]);
}

// ---------------------------------------------------------------------------------------------
// upsample_dpv: F.interpolate(bilinear, align_corners=False) along D only (the W axis keeps its
// size => identity weights), renormalise over the new bins, optional exp in / log out.
// ATen area_pixel_compute_source_index: src = scale*(dst+0.5)-0.5 clamped at 0, scale = D/Dout.
__global__ void resample_depth_kernel(const float* __restrict__ dpv, int D, size_t HW, long long npix, int Dout,
                                      int bv_log, float* __restrict__ out) {
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const int b = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)b * HW);
  const float* src = dpv + (size_t)b * D * HW + hw;
  float* dst = out + (size_t)b * Dout * HW + hw;
  const float scale = (float)D / (float)Dout;
  float total = 0.f;
  for (int pass = 0; pass < 2; pass++) {
    for (int o = 0; o < Dout; o++) {
      float s = fmaxf(scale * ((float)o + 0.5f) - 0.5f, 0.f);
      int i0 = (int)s;
      int i1 = i0 + (i0 < D - 1 ? 1 : 0);
      float l1 = s - (float)i0, l0 = 1.f - l1;
      float a = src[(size_t)i0 * HW], c = src[(size_t)i1 * HW];
      if (bv_log) { a = expf(a); c = expf(c); }
      float v = l0 * a + l1 * c;
      if (pass == 0) total += v;
      else {
        v = v / total;
        dst[(size_t)o * HW] = bv_log ? logf(v) : v;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Bayesian multiply-renormalise with explicit likelihood volumes (models/models.py:1096-1122).
template <int DPT>
__global__ void __launch_bounds__(kThreads)
bayes_fuse_kernel(const float* __restrict__ logp_in, const float* __restrict__ liks, int nlik, int D, size_t HW,
                  long long npix, const float* __restrict__ d_candi, int spread_iter, int spread_n,
                  float* __restrict__ logp_out, float* __restrict__ mean_out) {
  __shared__ float dc[DPT];
  for (int i = threadIdx.x; i < DPT; i += kThreads) dc[i] = i < D ? d_candi[i] : 0.f;
  __syncthreads();
  const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const int b = (int)(pix / (long long)HW);
  const size_t hw = (size_t)(pix - (long long)b * HW);
  const size_t base = (size_t)b * D * HW + hw;
  float p[DPT];
  load_prob_from_log<DPT>(p, logp_in + base, HW, D);
  for (int l = 0; l < nlik; l++) {
    const float* lk = liks + ((size_t)b * nlik + l) * D * HW + hw;
    float s = 0.f;
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) {
        float v = ld_stream(lk + (size_t)d * HW);
        if (v != v) v = 0.f;
        v = fminf(fmaxf(v, eps_floor()), 1.f);
        p[d] *= v;
        s += p[d];
      }
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) p[d] = p[d] / s;
  }
  for (int i = 0; i < spread_iter; i++) spread_dispatch<DPT>(p, D, spread_n);
  float mean = store_log_and_mean<DPT>(p, logp_out + base, HW, D, dc);
  if (mean_out) mean_out[pix] = mean;
}

static inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

}  // namespace dpv

using namespace dpv;

extern "C" {

int dpv_version(void) { return 100; }
const char* dpv_last_error(void) { return g_err; }
long long dpv_launch_count(void) { return g_launches.load(); }

int dpv_logsoftmax_stats(const float* x, const float* add, int B, int D, int H, int W, const float* d_candi,
                         float* logp, float* mean, float* var, float* entropy, float* carry, void* stream) {
  DPV_REQUIRE(x && logp && d_candi, "dpv_logsoftmax_stats: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0, "dpv_logsoftmax_stats: bad shape B=%d D=%d H=%d W=%d", B, D, H, W);
  cudaStream_t st = (cudaStream_t)stream;
  const long long npix = (long long)B * H * W, hw_img = (long long)H * W;
  DPV_REQUIRE(B <= 65535 && hw_img * 32 < (1ll << 31), "dpv_logsoftmax_stats: B=%d or image too large", B);
  // full-resolution volumes (D = 64 / 128, no addend, rows of at least 32 pixels): the TMA-tiled kernel
  // (dpv_softmax_tma.cu); anything else, or DPV_LSM_TMA=0, takes the register-column kernels below
  if (add == nullptr) {
    const int rc = logsoftmax_stats_tma(x, B, D, H, W, d_candi, logp, mean, var, entropy, carry, st);
    if (rc >= 0) return rc;
  }
  const unsigned g = blocks_for(npix, kThreads);
#define LAUNCH(DPT, T)                                                                                       \
  do {                                                                                                          \
    if (D == DPT * T)                                                                                           \
      lsm_stats_kernel<DPT, T, true><<<dim3(blocks_for(hw_img * T, kThreads), (unsigned)B), kThreads, 0, st>>>( \
          x, add, D, H, W, d_candi, logp, mean, var, entropy, carry);                                           \
    else                                                                                                        \
      lsm_stats_kernel<DPT, T, false><<<dim3(blocks_for(hw_img * T, kThreads), (unsigned)B), kThreads, 0, st>>>( \
          x, add, D, H, W, d_candi, logp, mean, var, entropy, carry);                                           \
  } while (0)
  if (D <= 16) LAUNCH(16, 1);
  else if (D <= 32) LAUNCH(32, 1);
  else if (D <= 64) {
    static const int lanes = getenv("DPV_LSM_LANES") ? atoi(getenv("DPV_LSM_LANES")) : 4;
    if (lanes == 2) LAUNCH(32, 2);                             // 4 lanes: 64 registers, twice the resident warps
    else LAUNCH(16, 4);
  } else if (D <= 128) LAUNCH(32, 4);
  else
    lsm_stats_generic_kernel<<<g, kThreads, 0, st>>>(x, add, D, H, W, npix, d_candi, logp, mean, var, entropy, carry);
#undef LAUNCH
  return check_launch("dpv_logsoftmax_stats");
}

int dpv_expectation(const float* dpv_, int B, int D, int H, int W, const float* d_candi, int bv_log, float* mean,
                    void* stream) {
  DPV_REQUIRE(dpv_ && mean && d_candi, "dpv_expectation: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0, "dpv_expectation: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t HW = (size_t)H * W;
  const long long npix = (long long)B * HW;
  const unsigned g = blocks_for(npix, kThreads);
  if (D <= 16) expectation_kernel<16><<<g, kThreads, 0, st>>>(dpv_, D, HW, npix, d_candi, bv_log, mean);
  else if (D <= 32) expectation_kernel<32><<<g, kThreads, 0, st>>>(dpv_, D, HW, npix, d_candi, bv_log, mean);
  else if (D <= 64) expectation_kernel<64><<<g, kThreads, 0, st>>>(dpv_, D, HW, npix, d_candi, bv_log, mean);
  else if (D <= 128) expectation_kernel<128><<<g, kThreads, 0, st>>>(dpv_, D, HW, npix, d_candi, bv_log, mean);
  else expectation_generic_kernel<<<g, kThreads, 0, st>>>(dpv_, D, HW, npix, d_candi, bv_log, mean);
  return check_launch("dpv_expectation");
}

int dpv_spread(const float* p, int B, int D, int H, int W, int N, float* out, void* stream) {
  DPV_REQUIRE(p && out, "dpv_spread: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0, "dpv_spread: bad shape");
  DPV_REQUIRE(N == 1 || N == 3 || N == 5 || N == 7, "dpv_spread: kernel size N=%d not in {1,3,5,7}", N);
  DPV_REQUIRE(D <= 128, "dpv_spread: D=%d > 128 unsupported", D);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t HW = (size_t)H * W;
  const long long npix = (long long)B * HW;
  const unsigned g = blocks_for(npix, kThreads);
  if (D <= 16) spread_kernel<16><<<g, kThreads, 0, st>>>(p, D, HW, npix, N, out);
  else if (D <= 32) spread_kernel<32><<<g, kThreads, 0, st>>>(p, D, HW, npix, N, out);
  else if (D <= 64) spread_kernel<64><<<g, kThreads, 0, st>>>(p, D, HW, npix, N, out);
  else spread_kernel<128><<<g, kThreads, 0, st>>>(p, D, HW, npix, N, out);
  return check_launch("dpv_spread");
}

int dpv_resample_depth(const float* dpv_, int B, int D, int H, int W, int Dout, int bv_log, float* out,
                       void* stream) {
  DPV_REQUIRE(dpv_ && out, "dpv_resample_depth: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0 && Dout > 0, "dpv_resample_depth: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t HW = (size_t)H * W;
  const long long npix = (long long)B * HW;
  resample_depth_kernel<<<blocks_for(npix, 256), 256, 0, st>>>(dpv_, D, HW, npix, Dout, bv_log, out);
  return check_launch("dpv_resample_depth");
}

int dpv_bayes_fuse(const float* logp_in, const float* liks, int nlik, int B, int D, int H, int W,
                   const float* d_candi, int spread_iter, int spread_n, float* logp_out, float* mean_out,
                   void* stream) {
  DPV_REQUIRE(logp_in && logp_out && d_candi && (liks || nlik == 0), "dpv_bayes_fuse: null pointer");
  DPV_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0 && nlik >= 0 && spread_iter >= 0, "dpv_bayes_fuse: bad shape");
  DPV_REQUIRE(spread_iter == 0 || spread_n == 1 || spread_n == 3 || spread_n == 5 || spread_n == 7,
              "dpv_bayes_fuse: spread_n=%d not in {1,3,5,7}", spread_n);
  DPV_REQUIRE(D <= 128, "dpv_bayes_fuse: D=%d > 128 unsupported", D);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t HW = (size_t)H * W;
  const long long npix = (long long)B * HW;
  const unsigned g = blocks_for(npix, kThreads);
#define LAUNCH(DPT)                                                                                          \
  bayes_fuse_kernel<DPT><<<g, kThreads, 0, st>>>(logp_in, liks, nlik, D, HW, npix, d_candi, spread_iter, \
                                                 spread_n, logp_out, mean_out)
  if (D <= 16) LAUNCH(16);
  else if (D <= 32) LAUNCH(32);
  else if (D <= 64) LAUNCH(64);
  else LAUNCH(128);
#undef LAUNCH
  return check_launch("dpv_bayes_fuse");
}

}  // extern "C"
