// (d3) DP planner solve kernel.  external/lcsim/lcsim/src/planning.cpp:75-126 (optimizedDesignPts),
// :162-217 (Trajectory compare).
//
// dp(i,k) = umap(k,i) + best over the edge window [lo,hi] of ray i+1, rays processed right to left.
// The reference's total order over a node's edges is
//   (unc = fl(u + unc[e]) larger) > (las = dth^2 + las[e] smaller) > (first in edge order), "no edge" first,
// and max-propagation through zero regions of the map makes exact ties in the first key COMMON, so the
// second key decides real paths and must be evaluated exactly.  Here LANES threads share one node and
// scan its window in two cheap vectorised passes over shared memory (LDS.128):
//   pass 1: m = max unc[e]  (fl(u + x) is monotone in x, so fl(u + m) is the winning first key);
//   pass 2: among the e with fl(u + unc[e]) == fl(u + m), the lexicographic minimum of (las, e).
// Both passes are plain associative reductions, so the selected index is exactly the one the reference's
// sequential strict-'>' scan selects: results are bit-exact.  One CTA = one umap; back-pointers for all
// W x R nodes stay in shared memory (W*R bytes) for the final path unroll.
#include <cstdlib>

#include "lc_device.cuh"

namespace dpv {

struct RayIn {
  float u, th;
  int lo, hi;
};

template <int LANES>
__global__ void __launch_bounds__(1024)
planner_solve_kernel(const float* __restrict__ umaps, int W, int R, int RP, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned char smraw[];
  float* unc = reinterpret_cast<float*>(smraw);            // [2][RP]   RP = R rounded up to 4 (pad = -inf)
  float* las = unc + 2 * RP;                               // [2][RP]
  float* th = las + 2 * RP;                                // [2][RP]
  int16_t* path = reinterpret_cast<int16_t*>(th + 2 * RP); // [W rounded to even]
  uint8_t* sub = reinterpret_cast<uint8_t*>(path + ((W + 1) & ~1));     // [W][R]
  __shared__ int s_cnt;

  const int k = threadIdx.x / LANES, s = threadIdx.x % LANES;
  const bool active = k < R;
  const float* umap = umaps + (size_t)blockIdx.x * R * W;
  const unsigned full = 0xffffffffu;
  // shuffles inside per-node branches name only the node's own lane group (branches differ between nodes of a warp)
  const unsigned gmask = (LANES == 32 ? 0xffffffffu : ((1u << LANES) - 1u)) << ((threadIdx.x & 31) & ~(LANES - 1));
  for (int i = threadIdx.x; i < 6 * RP; i += blockDim.x) unc[i] = -INFINITY;   // padding lanes never win a max
  __syncthreads();

  // lane s of a node prefetches (u, theta, window) of ray i0 - s one group of LANES rays ahead
  RayIn pre;
  auto fetch = [&](int i0) {
    const int i = i0 - s;
    pre.u = 0.f; pre.th = 0.f; pre.lo = 1; pre.hi = 0;
    if (active && i >= 0) {
      pre.u = __ldg(umap + (size_t)k * W + i);
      pre.th = __ldg(theta + (size_t)i * R + k);
      if (i < W - 1) {
        const uchar2 w2 = *reinterpret_cast<const uchar2*>(win + ((size_t)i * R + k) * 2);
        pre.lo = w2.x; pre.hi = w2.y;
      }
    }
  };
  fetch(W - 1);
  int buf = 0;
  for (int i0 = W - 1; i0 >= 0; i0 -= LANES) {
    const RayIn grp = pre;
    fetch(i0 - LANES);
    for (int j = 0; j < LANES && i0 - j >= 0; j++) {
      const int i = i0 - j;
      const int src = (threadIdx.x & 31 & ~(LANES - 1)) + j;
      const float u = __shfl_sync(full, grp.u, src);
      const float tn = __shfl_sync(full, grp.th, src);
      const int lo = __shfl_sync(full, grp.lo, src);
      const int hi = __shfl_sync(full, grp.hi, src);
      const float4* unc4 = reinterpret_cast<const float4*>(unc + (buf ^ 1) * RP);
      const float4* las4 = reinterpret_cast<const float4*>(las + (buf ^ 1) * RP);
      const float4* th4 = reinterpret_cast<const float4*>(th + (buf ^ 1) * RP);
      float bu = u, bl = 0.f;                                   // Trajectory(node): planning.cpp:169-184
      int be = -1;
      if (lo <= hi) {                                           // uniform within the node's lane group
        const int q0 = lo >> 2, q1 = hi >> 2;
        // ---- pass 1: window maximum of unc[] (entries outside [lo,hi] masked) ----
        float m = -INFINITY;
        for (int q = q0 + s; q <= q1; q += LANES) {
          const float4 v = unc4[q];
          const int e = q << 2;
          float a = (e >= lo && e <= hi) ? v.x : -INFINITY;
          float b = (e + 1 >= lo && e + 1 <= hi) ? v.y : -INFINITY;
          float c = (e + 2 >= lo && e + 2 <= hi) ? v.z : -INFINITY;
          float d = (e + 3 >= lo && e + 3 <= hi) ? v.w : -INFINITY;
          m = fmaxf(fmaxf(m, fmaxf(a, b)), fmaxf(c, d));
        }
#pragma unroll
        for (int off = LANES / 2; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(gmask, m, off));
        const float tmax = __fadd_rn(u, m);                     // :194
        if (tmax > u) {                                         // strict '>' against the bare node
          // ---- pass 2: among first-key ties, min (las, e) ----
          float cl = INFINITY;
          int ce = 0x7fffffff;
          for (int q = q0 + s; q <= q1; q += LANES) {
            const float4 v = unc4[q];
            const int e = q << 2;
            const bool ha = (e >= lo && e <= hi) && __fadd_rn(u, v.x) == tmax;
            const bool hb = (e + 1 >= lo && e + 1 <= hi) && __fadd_rn(u, v.y) == tmax;
            const bool hc = (e + 2 >= lo && e + 2 <= hi) && __fadd_rn(u, v.z) == tmax;
            const bool hd = (e + 3 >= lo && e + 3 <= hi) && __fadd_rn(u, v.w) == tmax;
            if (ha | hb | hc | hd) {
              const float4 t4 = th4[q], l4 = las4[q];
              float dth, tl;
              if (ha) { dth = __fsub_rn(t4.x, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.x); if (tl < cl) { cl = tl; ce = e; } }
              if (hb) { dth = __fsub_rn(t4.y, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.y); if (tl < cl) { cl = tl; ce = e + 1; } }
              if (hc) { dth = __fsub_rn(t4.z, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.z); if (tl < cl) { cl = tl; ce = e + 2; } }
              if (hd) { dth = __fsub_rn(t4.w, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.w); if (tl < cl) { cl = tl; ce = e + 3; } }
            }
          }
#pragma unroll
          for (int off = LANES / 2; off > 0; off >>= 1) {
            const float ol = __shfl_xor_sync(gmask, cl, off);
            const int oe = __shfl_xor_sync(gmask, ce, off);
            if (ol < cl || (ol == cl && oe < ce)) { cl = ol; ce = oe; }
          }
          bu = tmax; bl = cl; be = ce;                          // :197-198, :210-217
        }
      }
      if (active && s == 0) {
        unc[buf * RP + k] = bu;
        las[buf * RP + k] = bl;
        th[buf * RP + k] = tn;
        sub[(size_t)i * R + k] = be < 0 ? 255 : (uint8_t)be;
      }
      __syncthreads();
      buf ^= 1;
    }
  }
  // ray 0 results are in buffer buf^1.  Best start against a default (0,0,null) trajectory (:99-104)
  if (threadIdx.x == 0) {
    const float* u0 = unc + (buf ^ 1) * RP;
    const float* l0 = las + (buf ^ 1) * RP;
    float bu = 0.f, bl = 0.f;
    int bk = -1;
    for (int kk = 0; kk < R; kk++) {
      float tu = u0[kk], tl = l0[kk];
      bool gt = (tu > bu) ? true : ((tu == bu) ? (tl < bl) : false);
      if (gt) { bu = tu; bl = tl; bk = kk; }
    }
    int cnt = 0;
    if (bk >= 0) {                       // the reference dereferences a null node otherwise (:116)
      int i = 0, kk = bk;
      while (true) {
        path[cnt++] = (int16_t)kk;
        int nx = sub[(size_t)i * R + kk];
        if (nx == 255 || i == W - 1) break;
        kk = nx;
        i++;
      }
    }
    s_cnt = cnt;
    if (counts) counts[blockIdx.x] = cnt;
  }
  __syncthreads();
  const int cnt = s_cnt;
  for (int i = threadIdx.x; i < W; i += blockDim.x) {
    int kk = i < cnt ? path[i] : -1;
    if (ks) ks[(size_t)blockIdx.x * W + i] = kk;
    if (xz) {
      float x = 0.f, z = 0.f;
      if (kk >= 0) {
        x = node_xz[((size_t)i * R + kk) * 2];
        z = node_xz[((size_t)i * R + kk) * 2 + 1];
      }
      xz[((size_t)blockIdx.x * W + i) * 2] = x;
      xz[((size_t)blockIdx.x * W + i) * 2 + 1] = z;
    }
  }
}

}  // namespace dpv

using namespace dpv;

extern "C" int lc_planner_solve(LcPlanner* p, const float* umaps, int n, int32_t* ks, float* xz, int32_t* counts,
                                void* stream) {
  DPV_REQUIRE(p && umaps, "lc_planner_solve: null pointer");
  DPV_REQUIRE(n > 0, "lc_planner_solve: n=%d", n);
  const int W = p->W, R = p->R;
  const int RP = ((R + 3) / 4) * 4;
  int lanes = 8;
  if (const char* e = getenv("DPV_PLANNER_LANES")) lanes = atoi(e) > 0 ? atoi(e) : lanes;   // tuning aid (1,2,4,8)
  while (lanes > 1 && R * lanes > 1024) lanes >>= 1;
  const int threads = ((R * lanes + 31) / 32) * 32;
  size_t smem = (size_t)6 * RP * 4 + (size_t)((W + 1) & ~1) * 2 + (size_t)W * R;
  DPV_REQUIRE(smem <= 227 * 1024, "lc_planner_solve: W*R = %d*%d back-pointers exceed shared memory", W, R);
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(L)                                                                                               \
  do {                                                                                                          \
    DPV_CUDA(cudaFuncSetAttribute(planner_solve_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                                  (int)smem));                                                                  \
    planner_solve_kernel<L><<<n, threads, smem, st>>>(umaps, W, R, RP, p->d_theta, p->d_xz, p->d_win, ks, xz,   \
                                                      counts);                                                  \
  } while (0)
  if (lanes == 8) LAUNCH(8);
  else if (lanes == 4) LAUNCH(4);
  else if (lanes == 2) LAUNCH(2);
  else LAUNCH(1);
#undef LAUNCH
  return check_launch("lc_planner_solve");
}
