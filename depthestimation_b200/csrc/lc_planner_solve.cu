// (d3) DP planner solve kernel.  external/lcsim/lcsim/src/planning.cpp:75-126 (optimizedDesignPts),
// :162-217 (Trajectory compare).
//
// dp(i,k) = umap(k,i) + best over the edge window [lo,hi] of ray i+1, rays processed right to left.
// The reference's total order over a node's edges is
//   (unc = fl(u + unc[e]) larger) > (las = dth^2 + las[e] smaller) > (first in edge order), "no edge" first,
// and max-propagation through zero regions of the map makes exact ties in the first key COMMON (5-20 % of the
// nodes), so the second key decides real paths and must be evaluated exactly.  The kernel is bound by
// instruction issue on the ~W*R*|window| edge visits, so the scan is organised to touch most edges at a
// quarter instruction each:
//   * the ray just finished keeps, next to unc[], the maxima of its aligned groups of four (qm[]);
//   * pass 1: window max m from qm[] (interior) + the two boundary groups entry by entry;
//     fl(u + x) is monotone in x, so fl(u + m) is the winning first key;
//   * pass 2: a group can hold a first-key tie only if fl(u + qm) == fl(u + m); only those groups are
//     opened and their entries compared exactly on (las, e).
// Both passes are associative reductions under the reference's order, so the selected index is exactly the
// one its sequential strict-'>' scan selects: results are bit-exact.  One CTA = one umap, LANES threads per
// node; back-pointers for all W x R nodes stay in shared memory (W*R bytes) for the final path unroll.
#include <cstdlib>

#include "lc_device.cuh"

namespace dpv {

struct RayIn {
  float u, th;
  int lo, hi;
};

struct Best {          // running minimum of (las, e) among first-key ties
  float l;
  int e;
};

template <int LANES>
__global__ void __launch_bounds__(1024)
planner_solve_kernel(const float* __restrict__ umaps, int W, int R, int RP, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned char smraw[];
  const int NQ = RP / 4;
  float* unc = reinterpret_cast<float*>(smraw);            // [2][RP]   RP = R rounded up to 4 (pad = -inf)
  float* las = unc + 2 * RP;                               // [2][RP]
  float* th = las + 2 * RP;                                // [2][RP]
  float* qm = th + 2 * RP;                                 // [2][NQ]   max of unc[4q..4q+3]
  int16_t* path = reinterpret_cast<int16_t*>(qm + 2 * NQ + (NQ & 1) * 2);   // [W rounded to even]
  uint8_t* sub = reinterpret_cast<uint8_t*>(path + ((W + 1) & ~1));         // [W][R]
  __shared__ int s_cnt;

  const int k = threadIdx.x / LANES, s = threadIdx.x % LANES;
  const bool active = k < R;
  const float* umap = umaps + (size_t)blockIdx.x * R * W;
  const unsigned full = 0xffffffffu;
  // shuffles inside per-node branches name only the node's own lanes (branches differ between nodes of a warp)
  const unsigned gmask = ((1u << LANES) - 1u) << ((threadIdx.x & 31) & ~(LANES - 1));
  for (int i = threadIdx.x; i < 6 * RP + 2 * NQ; i += blockDim.x) unc[i] = -INFINITY;   // pads never win a max
  __syncthreads();

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
  // software pipeline: inputs of ray i-3 are requested while ray i is processed
  RayIn cur = fetch(W - 1), n1 = fetch(W - 2), n2 = fetch(W - 3);
  int buf = 0;
  for (int i = W - 1; i >= 0; i--) {
    const RayIn in = cur;
    cur = n1; n1 = n2; n2 = fetch(i - 3);
    const float* uncn = unc + (buf ^ 1) * RP;
    const float* lasn = las + (buf ^ 1) * RP;
    const float* thn = th + (buf ^ 1) * RP;
    const float* qmn = qm + (buf ^ 1) * NQ;
    const float u = in.u, tn = in.th;
    const int lo = in.lo, hi = in.hi;
    float bu = u, bl = 0.f;                                     // Trajectory(node): planning.cpp:169-184
    int be = -1;
    if (lo <= hi) {                                             // same for the LANES threads of a node
      const int qlo = lo >> 2, qhi = hi >> 2;
      // ---- pass 1: window maximum --------------------------------------------------------
      float m = -INFINITY, mb = -INFINITY;                      // mb: this lane's boundary-group maximum
      if (s == 0) {                                             // boundary groups, entry by entry
        const float4 a = reinterpret_cast<const float4*>(uncn)[qlo];
        const int e = qlo << 2;
        if (e >= lo && e <= hi) m = fmaxf(m, a.x);
        if (e + 1 >= lo && e + 1 <= hi) m = fmaxf(m, a.y);
        if (e + 2 >= lo && e + 2 <= hi) m = fmaxf(m, a.z);
        if (e + 3 >= lo && e + 3 <= hi) m = fmaxf(m, a.w);
        mb = m;
      }
      float mb2 = -INFINITY;
      if (s == LANES - 1 && qhi != qlo) {
        const float4 a = reinterpret_cast<const float4*>(uncn)[qhi];
        const int e = qhi << 2;
        if (e <= hi) mb2 = fmaxf(mb2, a.x);
        if (e + 1 <= hi) mb2 = fmaxf(mb2, a.y);
        if (e + 2 <= hi) mb2 = fmaxf(mb2, a.z);
        if (e + 3 <= hi) mb2 = fmaxf(mb2, a.w);
        m = fmaxf(m, mb2);
      }
      {
        float m1 = -INFINITY;
        int q = qlo + 1 + s;
        for (; q + LANES < qhi; q += 2 * LANES) {               // interior groups: one value per four edges
          m = fmaxf(m, qmn[q]);
          m1 = fmaxf(m1, qmn[q + LANES]);
        }
        if (q < qhi) m = fmaxf(m, qmn[q]);
        m = fmaxf(m, m1);
      }
#pragma unroll
      for (int off = LANES / 2; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(gmask, m, off));
      const float tmax = __fadd_rn(u, m);                       // :194
      if (tmax > u) {                                           // strict '>' against the bare node
        // ---- pass 2: among first-key ties, min (las, e) ---------------------------------
        Best bb;
        bb.l = INFINITY; bb.e = 0x7fffffff;
        auto open_group = [&](int q) {
          const float4 v = reinterpret_cast<const float4*>(uncn)[q];
          const float4 t4 = reinterpret_cast<const float4*>(thn)[q];
          const float4 l4 = reinterpret_cast<const float4*>(lasn)[q];
          const int e = q << 2;
          float dth, tl;
          if (e >= lo && e <= hi && __fadd_rn(u, v.x) == tmax) {
            dth = __fsub_rn(t4.x, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.x);          // :197-198
            if (tl < bb.l || (tl == bb.l && e < bb.e)) { bb.l = tl; bb.e = e; }
          }
          if (e + 1 >= lo && e + 1 <= hi && __fadd_rn(u, v.y) == tmax) {
            dth = __fsub_rn(t4.y, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.y);
            if (tl < bb.l || (tl == bb.l && e + 1 < bb.e)) { bb.l = tl; bb.e = e + 1; }
          }
          if (e + 2 >= lo && e + 2 <= hi && __fadd_rn(u, v.z) == tmax) {
            dth = __fsub_rn(t4.z, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.z);
            if (tl < bb.l || (tl == bb.l && e + 2 < bb.e)) { bb.l = tl; bb.e = e + 2; }
          }
          if (e + 3 >= lo && e + 3 <= hi && __fadd_rn(u, v.w) == tmax) {
            dth = __fsub_rn(t4.w, tn); tl = __fadd_rn(__fmul_rn(dth, dth), l4.w);
            if (tl < bb.l || (tl == bb.l && e + 3 < bb.e)) { bb.l = tl; bb.e = e + 3; }
          }
        };
        if (s == 0 && __fadd_rn(u, mb) == tmax) open_group(qlo);
        if (s == LANES - 1 && qhi != qlo && __fadd_rn(u, mb2) == tmax) open_group(qhi);
        for (int q = qlo + 1 + s; q < qhi; q += LANES)
          if (__fadd_rn(u, qmn[q]) == tmax) open_group(q);
#pragma unroll
        for (int off = LANES / 2; off > 0; off >>= 1) {
          const float ol = __shfl_xor_sync(gmask, bb.l, off);
          const int oe = __shfl_xor_sync(gmask, bb.e, off);
          if (ol < bb.l || (ol == bb.l && oe < bb.e)) { bb.l = ol; bb.e = oe; }
        }
        bu = tmax; bl = bb.l; be = bb.e;                        // :210-217
      }
    }
    // group-of-four maxima of the new unc[] (nodes 4j..4j+3 are 4*LANES adjacent lanes)
    float gq = active ? bu : -INFINITY;
    gq = fmaxf(gq, __shfl_xor_sync(full, gq, LANES));
    gq = fmaxf(gq, __shfl_xor_sync(full, gq, 2 * LANES));
    if (s == 0 && k < RP) {
      if (active) {
        unc[buf * RP + k] = bu;
        las[buf * RP + k] = bl;
        th[buf * RP + k] = tn;
        sub[(size_t)i * R + k] = be < 0 ? 255 : (uint8_t)be;
      }
      if ((k & 3) == 0) qm[buf * NQ + (k >> 2)] = gq;
    }
    __syncthreads();
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
  const int RP = ((R + 3) / 4) * 4;
  int lanes = 4;
  if (const char* e = getenv("DPV_PLANNER_LANES")) lanes = atoi(e);                // tuning aid (1,2,4,8)
  if (lanes != 1 && lanes != 2 && lanes != 4 && lanes != 8) lanes = 4;
  while (lanes > 1 && RP * lanes > 1024) lanes >>= 1;
  const int threads = ((RP * lanes + 31) / 32) * 32;
  const int NQ = RP / 4;
  size_t smem = (size_t)6 * RP * 4 + (size_t)(2 * NQ + (NQ & 1) * 2) * 4 + (size_t)((W + 1) & ~1) * 2 + (size_t)W * R;
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
