// (d3) DP planner solve kernel.  external/lcsim/lcsim/src/planning.cpp:75-126 (optimizedDesignPts),
// :162-217 (Trajectory compare).
//
// dp(i,k) = umap(k,i) + best over the edge window [lo,hi] of ray i+1, rays processed right to left.
// The reference scans every edge (W*R*|window| ~ 10 M relaxations per solve).  The primary key
// fl(u + unc[e]) is monotone in unc[e], so the winner is the window's arg-max of unc[] unless another
// entry rounds to the same sum.  Per ray the CTA builds a sparse table (range top-2 query: max, its
// first index, second max) over the R values of the ray just finished; each node then needs ONE O(1)
// query: if fl(u + second) < fl(u + max) the arg-max is the unique winner, else (ties after rounding:
// rare) it falls back to the reference's sequential scan with its exact tie rules
//   (unc larger) > (las smaller) > (first in edge order), "no edge" before all.
// Either way the selected index is the one the reference selects: results are bit-exact.
// One CTA = one umap, one thread = one range node; back-pointers for all W x R nodes stay in
// shared memory (W*R bytes) for the final path unroll.
#include "lc_device.cuh"

namespace dpv {

struct Top2 {
  float v1, v2;
  int p1;
};

__device__ __forceinline__ Top2 top2_merge(const Top2& a, const Top2& b) {   // a covers the lower indices
  Top2 r;
  if (a.p1 == b.p1) { r.v1 = a.v1; r.p1 = a.p1; r.v2 = fmaxf(a.v2, b.v2); }
  else if (a.v1 > b.v1) { r.v1 = a.v1; r.p1 = a.p1; r.v2 = fmaxf(a.v2, b.v1); }
  else if (a.v1 < b.v1) { r.v1 = b.v1; r.p1 = b.p1; r.v2 = fmaxf(a.v1, b.v2); }
  else { r.v1 = a.v1; r.p1 = a.p1; r.v2 = a.v1; }
  return r;
}

struct RayIn {
  float u, th;
  int lo, hi;
};

__global__ void __launch_bounds__(256)
planner_solve_kernel(const float* __restrict__ umaps, int W, int R, int RP, int LV, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts) {
  extern __shared__ unsigned char smraw[];
  float* unc = reinterpret_cast<float*>(smraw);            // [2][RP]
  float* las = unc + 2 * RP;                               // [2][RP]
  float* th = las + 2 * RP;                                // [2][RP]
  float* tv1 = th + 2 * RP;                                // [LV][RP]  (level 0 unused: it is unc itself)
  float* tv2 = tv1 + LV * RP;                              // [LV][RP]
  int16_t* path = reinterpret_cast<int16_t*>(tv2 + LV * RP);            // [W rounded to even]
  uint8_t* tp1 = reinterpret_cast<uint8_t*>(path + ((W + 1) & ~1));     // [LV][RP]
  uint8_t* sub = tp1 + LV * RP;                            // [W][R]
  __shared__ int s_cnt;

  const int k = threadIdx.x;
  const bool active = k < R;
  const float* umap = umaps + (size_t)blockIdx.x * R * W;

  auto fetch = [&](int i) {
    RayIn r;
    r.u = 0.f; r.th = 0.f; r.lo = 1; r.hi = 0;
    if (active && i >= 0) {
      r.u = __ldg(umap + (size_t)k * W + i);
      r.th = __ldg(theta + (size_t)i * R + k);
      if (i < W - 1) {
        const uchar2 w2 = *reinterpret_cast<const uchar2*>(win + ((size_t)i * R + k) * 2);
        r.lo = w2.x; r.hi = w2.y;
      }
    }
    return r;
  };
  RayIn cur = fetch(W - 1), n1 = fetch(W - 2), n2 = fetch(W - 3);
  int buf = 0;
  for (int i = W - 1; i >= 0; i--) {
    const RayIn in = cur;
    cur = n1; n1 = n2; n2 = fetch(i - 3);
    const float* uncn = unc + (buf ^ 1) * RP;
    const float* lasn = las + (buf ^ 1) * RP;
    const float* thn = th + (buf ^ 1) * RP;
    if (active) {
      float bu = in.u, bl = 0.f;                              // Trajectory(node): planning.cpp:169-184
      int be = -1;
      if (in.lo <= in.hi) {
        const int len = in.hi - in.lo + 1;
        const int j = 31 - __clz(len);
        const int ib = in.hi - (1 << j) + 1;
        Top2 a, b;
        if (j == 0) {
          a.v1 = uncn[in.lo]; a.p1 = in.lo; a.v2 = -INFINITY;
          b = a;
        } else {
          a.v1 = tv1[j * RP + in.lo]; a.v2 = tv2[j * RP + in.lo]; a.p1 = tp1[j * RP + in.lo];
          b.v1 = tv1[j * RP + ib]; b.v2 = tv2[j * RP + ib]; b.p1 = tp1[j * RP + ib];
        }
        const Top2 m = top2_merge(a, b);
        const float tmax = __fadd_rn(in.u, m.v1);             // :194
        if (tmax > in.u) {                                    // otherwise no edge beats the bare node (strict >)
          if (__fadd_rn(in.u, m.v2) < tmax) {                 // unique maximiser of the primary key
            const float dth = __fsub_rn(thn[m.p1], in.th);
            bu = tmax;
            bl = __fadd_rn(__fmul_rn(dth, dth), lasn[m.p1]);  // :197-198
            be = m.p1;
          } else {                                            // rounding ties: the reference's own scan
            for (int e = in.lo; e <= in.hi; e++) {
              const float tu = __fadd_rn(in.u, uncn[e]);
              const float dth = __fsub_rn(thn[e], in.th);
              const float tl = __fadd_rn(__fmul_rn(dth, dth), lasn[e]);
              const bool gt = (tu > bu) ? true : ((tu == bu) ? (tl < bl) : false);     // :210-217
              if (gt) { bu = tu; bl = tl; be = e; }
            }
          }
        }
      }
      unc[buf * RP + k] = bu;
      las[buf * RP + k] = bl;
      th[buf * RP + k] = in.th;
      sub[(size_t)i * R + k] = be < 0 ? 255 : (uint8_t)be;
    }
    __syncthreads();
    // sparse table over this ray's unc[] for the next ray's queries
    if (i > 0) {
      const float* src = unc + buf * RP;
      for (int j = 1; j < LV; j++) {
        const int half = 1 << (j - 1);
        if (k + 2 * half <= R) {
          Top2 a, b;
          if (j == 1) {
            a.v1 = src[k]; a.p1 = k; a.v2 = -INFINITY;
            b.v1 = src[k + 1]; b.p1 = k + 1; b.v2 = -INFINITY;
          } else {
            const int o = (j - 1) * RP;
            a.v1 = tv1[o + k]; a.v2 = tv2[o + k]; a.p1 = tp1[o + k];
            b.v1 = tv1[o + k + half]; b.v2 = tv2[o + k + half]; b.p1 = tp1[o + k + half];
          }
          const Top2 m = top2_merge(a, b);
          tv1[j * RP + k] = m.v1; tv2[j * RP + k] = m.v2; tp1[j * RP + k] = (uint8_t)m.p1;
        }
        __syncthreads();
      }
    }
    buf ^= 1;
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
  const int RP = ((R + 31) / 32) * 32;
  int LV = 1;
  while ((1 << LV) <= R) LV++;                      // levels 0 .. floor(log2 R)
  size_t smem = (size_t)6 * RP * 4 + (size_t)2 * LV * RP * 4 + (size_t)((W + 1) & ~1) * 2 + (size_t)LV * RP +
                (size_t)W * R;
  DPV_REQUIRE(smem <= 227 * 1024, "lc_planner_solve: W*R = %d*%d back-pointers exceed shared memory", W, R);
  DPV_CUDA(cudaFuncSetAttribute(planner_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  planner_solve_kernel<<<n, RP, smem, (cudaStream_t)stream>>>(umaps, W, R, RP, LV, p->d_theta, p->d_xz, p->d_win, ks,
                                                              xz, counts);
  return check_launch("lc_planner_solve");
}
