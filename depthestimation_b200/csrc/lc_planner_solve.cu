// (d3) DP planner solve kernel.  external/lcsim/lcsim/src/planning.cpp:75-126 (optimizedDesignPts),
// :162-217 (Trajectory compare).
//
// dp(i,k) = umap(k,i) + best over the edge window [lo,hi] of ray i+1, rays processed right to left.
// The reference's total order over a node's edges is
//   (unc = fl(u + unc[e]) larger) > (las = dth^2 + las[e] smaller) > (first in edge order), "no edge" first,
// and max-propagation through zero regions of the map makes exact ties in the first key COMMON (about a fifth
// of the nodes), so the second key decides real paths and must be evaluated exactly.
//
// The rays are a serial chain (W steps) and one CTA works on one umap on an otherwise idle SM, so what counts
// is the length of the dependent instruction chain per ray, not throughput.  A ray step has two parts, written
// branch-free with all shared-memory loads issued up front, and run by different warps one ray apart:
//   A. (warps 0..NA-1, ray i) LANES threads per node.  The finished ray keeps, next to unc[], the maxima of
//      its aligned groups of four (qm[]).  A lane holds up to NQL interior groups of the node's window in
//      registers; lanes 0 and LANES-1 also hold the two boundary groups entry by entry.  Window maximum m;
//      fl(u + x) is monotone in x, so fl(u + m) is the winning first key and the node's new unc.  The groups /
//      boundary entries that can hold a first-key tie (fl(u + .) == fl(u + m)) are counted.  If exactly one
//      group qualifies, its lane opens it; if exactly one of its entries ties, the successor is known.
//      Everything else (several groups or several entries tie) is queued.
//   B. (warps NA.., ray i+1) one WARP per queued node: the previous ray sits in the warp's registers, four
//      entries per lane; tie flags, the second key for the tied entries, and two REDUX.min give the exact
//      (las, e) minimum with no divergence, however many entries tie.
// unc/th/qm of a ray are complete after A, its las only after B, and A never needs las: the las of a node A
// finished is filled in one step later by the same thread (its successor's las is complete by then).  So A(i)
// and B(i+1) overlap, with one barrier per ray and three rays of state live (ring of three buffers).
// Both parts are associative reductions under the reference's order, so the selected edge is exactly the one
// its sequential strict-'>' scan selects: results are bit-exact.  Back-pointers for all W x R nodes stay in
// shared memory (W*R bytes) for the final path unroll.
#include <cstdlib>

#include "lc_device.cuh"

namespace dpv {

struct RayIn {
  float u, th;
  unsigned w;         // lo | hi << 8, unpacked only when used (an early unpack would wait for the load)
};

struct __align__(16) TieRec {       // a node queued for phase B
  int k, lo, hi;
  float u;
  float tn, tmax, pad0, pad1;
};

constexpr unsigned kNoWindow = 1u;  // lo = 1 > hi = 0

// LANES threads per node, NQL interior groups per lane: covers R <= 4 * LANES * NQL range entries
template <int LANES, int NQL>
__global__ void __launch_bounds__(1024)
planner_solve_kernel(const float* __restrict__ umaps, int W, int R, int NA, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts) {
  constexpr int NQ = LANES * NQL;                                 // groups of four per ray
  constexpr int RP = 4 * NQ;                                      // padded ray length (multiple of 128)
  constexpr int NCH = RP / 128;
  constexpr int SLOT = 3 * RP + 2 * NQ;                           // floats per ray: unc (pad -inf), las, th, qm (2*NQ)
  static_assert(RP % 128 == 0, "ray length must fill whole warps in phase B");
  extern __shared__ __align__(16) unsigned char smraw[];
  float* state = reinterpret_cast<float*>(smraw);                 // [3][SLOT] ring over rays
  TieRec* rec = reinterpret_cast<TieRec*>(state + 3 * SLOT);      // [2][R] queues, by ray parity
  int16_t* path = reinterpret_cast<int16_t*>(rec + 2 * R);        // [W rounded to 8]
  uint8_t* sub = reinterpret_cast<uint8_t*>(path + ((W + 7) & ~7));   // [W][R]
  __shared__ int s_ties[3];                                       // queue lengths, ring like state
  __shared__ int s_cnt;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int k = threadIdx.x / LANES, s = threadIdx.x % LANES;
  const bool active = k < R;                                      // phase-A thread with a node
  const bool in_a = warp < NA;                                    // NA = warps covering R*LANES threads
  const unsigned full = 0xffffffffu;

  for (int i = threadIdx.x; i < 3 * SLOT; i += blockDim.x) {
    const int o = i % SLOT;
    state[i] = (o < RP || o >= 3 * RP) ? -INFINITY : 0.f;         // unc and qm lose every maximum; las = th = 0
  }
  if (threadIdx.x < 3) s_ties[threadIdx.x] = 0;
  __syncthreads();

  // input pointers of this thread's node, walked right to left
  const int kc = active ? k : 0;
  const float* pu = umaps + (size_t)blockIdx.x * R * W + (size_t)kc * W + (W - 1);
  const float* pth = theta + (size_t)(W - 1) * R + kc;
  const uint16_t* pw = reinterpret_cast<const uint16_t*>(win) + (size_t)(W - 1) * R + kc;
  auto fetch = [&](int i) {                                       // i = ray the pointers stand on
    RayIn r;
    r.u = 0.f; r.th = 0.f; r.w = kNoWindow;
    if (in_a && active && i >= 0) {
      r.u = __ldg(pu);
      r.th = __ldg(pth);
      if (i < W - 1) r.w = __ldg(pw);
    }
    pu -= 1; pth -= R; pw -= R;
    return r;
  };
  // software pipeline: inputs of ray i-3 are requested while ray i is processed
  RayIn cur = fetch(W - 1), n1 = fetch(W - 2), n2 = fetch(W - 3);
  // a node this thread finished on the previous ray still waits for its las
  int pend_e = -1;
  float pend_tn = 0.f;

  // ring slots: s0 = ray i (written by A), s1 = ray i+1 (read by A, las completed by B), s2 = ray i+2 (read by B)
  int s0 = (W - 1) % 3, s1 = W % 3, s2 = (W + 1) % 3;
  for (int i = W - 1; i >= -1; i--) {
    float* st0 = state + s0 * SLOT;
    float* st1 = state + s1 * SLOT;
    const float* st2 = state + s2 * SLOT;
    if (in_a) {
      // ---- las of the node finished one ray ago (ray i+1): its successor on ray i+2 is complete now  :197-198
      if (pend_e >= 0) {
        const float dth = __fsub_rn(st2[2 * RP + pend_e], pend_tn);
        st1[RP + k] = __fadd_rn(__fmul_rn(dth, dth), st2[RP + pend_e]);
      }
      pend_e = -1;
      if (i >= 0) {
        // ================= part A on ray i =================
        const float* uncn = st1;
        const float* qmn = st1 + 3 * RP;
        const RayIn in = cur;
        cur = n1; n1 = n2; n2 = fetch(i - 3);
        const float u = in.u, tn = in.th;
        const int lo = (int)(in.w & 0xffu), hi = (int)(in.w >> 8);
        const int qlo = lo >> 2, qhi = hi >> 2;                   // no window: lo = 1 > hi = 0 masks everything below
        // ---- loads: this lane's boundary group (lane 0: group of lo, lane LANES-1: group of hi) and its
        //      interior groups qlo+1+s, +LANES, ... < qhi  (qm[] is padded, no index clamp)
        const bool is_b = s == 0 || (s == LANES - 1 && qhi != qlo);
        const int qb = s == 0 ? qlo : qhi;
        const float4 araw = reinterpret_cast<const float4*>(uncn)[qb];
        const float* qp = qmn + qlo + 1 + s;
        const int rem = qhi - qlo - 1 - s;                        // interior group j exists iff j*LANES < rem
        float qv[NQL];
#pragma unroll
        for (int j = 0; j < NQL; j++) qv[j] = qp[j * LANES];
        const int eb = qb << 2;
        float4 a;                                                 // boundary entries inside the window, else -inf
        a.x = (is_b && eb >= lo && eb <= hi) ? araw.x : -INFINITY;
        a.y = (is_b && eb + 1 >= lo && eb + 1 <= hi) ? araw.y : -INFINITY;
        a.z = (is_b && eb + 2 >= lo && eb + 2 <= hi) ? araw.z : -INFINITY;
        a.w = (is_b && eb + 3 >= lo && eb + 3 <= hi) ? araw.w : -INFINITY;
        float m = fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)), m1 = -INFINITY;
#pragma unroll
        for (int j = 0; j < NQL; j++) {
          if (!(j * LANES < rem)) qv[j] = -INFINITY;
          if (j & 1) m1 = fmaxf(m1, qv[j]);
          else m = fmaxf(m, qv[j]);
        }
        m = fmaxf(m, m1);
#pragma unroll
        for (int off = LANES / 2; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(full, m, off));
        const float tmax = __fadd_rn(u, m);                       // :194
        const bool better = tmax > u;                             // strict '>' against the bare node
        // ---- which groups / boundary entries can hold a first-key tie (-inf never does once tmax > u)
        unsigned bb = 0, qbits = 0;
        if (__fadd_rn(u, a.x) == tmax) bb |= 1u;
        if (__fadd_rn(u, a.y) == tmax) bb |= 2u;
        if (__fadd_rn(u, a.z) == tmax) bb |= 4u;
        if (__fadd_rn(u, a.w) == tmax) bb |= 8u;
#pragma unroll
        for (int j = 0; j < NQL; j++)
          if (__fadd_rn(u, qv[j]) == tmax) qbits |= 1u << j;
        if (!better) { bb = 0; qbits = 0; }
        const int mine = __popc(bb) + __popc(qbits);
        int tot = mine;
#pragma unroll
        for (int off = LANES / 2; off > 0; off >>= 1) tot += __shfl_xor_sync(full, tot, off);
        // ---- exactly one candidate: its lane names the successor, or finds several ties inside an interior group
        const bool owner = tot == 1 && mine == 1;                 // implies better
        bool solved = false;
        int e1 = -1;
        if (owner) {
          e1 = eb + __ffs(bb) - 1;                                // the tied boundary entry
          solved = bb != 0;
          if (!solved) {                                          // an interior group: open it
            const int q = qlo + 1 + s + (__ffs(qbits) - 1) * LANES;
            const float4 v = reinterpret_cast<const float4*>(uncn)[q];
            unsigned eq = 0;
            if (__fadd_rn(u, v.x) == tmax) eq |= 1u;
            if (__fadd_rn(u, v.y) == tmax) eq |= 2u;
            if (__fadd_rn(u, v.z) == tmax) eq |= 4u;
            if (__fadd_rn(u, v.w) == tmax) eq |= 8u;
            e1 = (q << 2) + __ffs(eq) - 1;
            solved = __popc(eq) == 1;
          }
        }
        // ---- queue what is left for part B (one pushing lane per node)
        const bool push = active && ((owner && !solved) || (s == 0 && tot >= 2));
        const unsigned pb = __ballot_sync(full, push);
        if (pb) {
          int base = 0;
          if (lane == 0) base = atomicAdd(&s_ties[s0], __popc(pb));
          base = __shfl_sync(full, base, 0);
          if (push) {
            TieRec r;
            r.k = k; r.lo = lo; r.hi = hi; r.u = u;
            r.tn = tn; r.tmax = tmax; r.pad0 = 0.f; r.pad1 = 0.f;
            rec[(i & 1) * R + base + __popc(pb & ((1u << lane) - 1u))] = r;
          }
        }
        // ---- results; group-of-four maxima of the new unc[] (4*LANES adjacent lanes)
        const float bu = better ? tmax : u;                       // :210-217
        float gq = active ? bu : -INFINITY;
        gq = fmaxf(gq, __shfl_xor_sync(full, gq, LANES));
        gq = fmaxf(gq, __shfl_xor_sync(full, gq, 2 * LANES));
        if (active) {
          uint8_t* subi = sub + (size_t)i * R;
          if (s == 0) {
            st0[k] = bu;
            st0[2 * RP + k] = tn;
            if ((k & 3) == 0) st0[3 * RP + (k >> 2)] = gq;
            if (!better) {                                        // the bare node: las 0, no successor
              st0[RP + k] = 0.f;
              subi[k] = 255;
            }
          }
          if (solved) {                                           // successor known, las follows one ray later
            subi[k] = (uint8_t)e1;
            pend_e = e1;
            pend_tn = tn;
          }
        }
      }
    } else if (i < W - 1) {
      // ================= part B on ray i+1: queued nodes, one warp each =================
      const int nt = s_ties[s1];
      const TieRec* rq = rec + ((i + 1) & 1) * R;
      uint8_t* subi = sub + (size_t)(i + 1) * R;
      for (int j = warp - NA; j < nt; j += nwarps - NA) {
        const float4 r0 = reinterpret_cast<const float4*>(rq + j)[0], r1 = reinterpret_cast<const float4*>(rq + j)[1];
        const int rk = __float_as_int(r0.x), rlo = __float_as_int(r0.y), rhi = __float_as_int(r0.z);
        const float ru = r0.w, rtn = r1.x, rtmax = r1.y;
        float tl_best = INFINITY;
        int e_best = 0x7fffffff;
#pragma unroll
        for (int c = 0; c < NCH; c++) {                           // chunks of 128 entries, four per lane
          const int e = c * 128 + lane * 4;
          const float4 v = reinterpret_cast<const float4*>(st2)[c * 32 + lane];
          const float4 l4 = reinterpret_cast<const float4*>(st2 + RP)[c * 32 + lane];
          const float4 t4 = reinterpret_cast<const float4*>(st2 + 2 * RP)[c * 32 + lane];
          const float vv[4] = {v.x, v.y, v.z, v.w}, tt[4] = {t4.x, t4.y, t4.z, t4.w}, ll[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const float dth = __fsub_rn(tt[q], rtn);              // :197-198
            const float tl = __fadd_rn(__fmul_rn(dth, dth), ll[q]);
            const bool tie = e + q >= rlo && e + q <= rhi && __fadd_rn(ru, vv[q]) == rtmax;
            if (tie && tl < tl_best) { tl_best = tl; e_best = e + q; }    // increasing e: first minimum stays
          }
        }
        const unsigned tmin = __reduce_min_sync(full, __float_as_uint(tl_best));   // las >= +0: bit order = value order
        const int be = __reduce_min_sync(full, __float_as_uint(tl_best) == tmin ? e_best : 0x7fffffff);
        if (lane == 0) {
          st1[RP + rk] = __uint_as_float(tmin);
          subi[rk] = (uint8_t)be;
        }
      }
    }
    if (threadIdx.x == blockDim.x - 1) s_ties[s2] = 0;            // read last step (as s1), pushed next step (as s0)
    __syncthreads();
    const int t = s2; s2 = s1; s1 = s0; s0 = t;                   // ray i-1 takes the slot ray i+2 leaves
  }
  // ray 0 is now in slot s2 (after the last rotation).  Best start against a default (0,0,null) trajectory (:99-104)
  if (threadIdx.x == 0) {
    const float* u0 = state + s2 * SLOT;
    const float* l0 = u0 + RP;
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
        const float2 p = *reinterpret_cast<const float2*>(node_xz + ((size_t)i * R + kk) * 2);
        x = p.x; z = p.y;
      }
      *reinterpret_cast<float2*>(xz + ((size_t)blockIdx.x * W + i) * 2) = make_float2(x, z);
    }
  }
}

}  // namespace dpv

using namespace dpv;

namespace {

template <int LANES, int NQL>
int launch_solve(const LcPlanner* p, const float* umaps, int n, int warps, int32_t* ks, float* xz, int32_t* counts,
                 cudaStream_t st) {
  static_assert(LANES >= 2, "the two boundary groups of a window are held by different lanes");
  const int W = p->W, R = p->R;
  constexpr int RP = 4 * LANES * NQL;
  DPV_REQUIRE(R <= RP, "lc_planner_solve: R=%d does not fit the %d-entry kernel", R, RP);
  const size_t smem = (size_t)3 * (3 * RP + 2 * (RP / 4)) * 4 + (size_t)2 * R * sizeof(TieRec) +
                      (size_t)((W + 7) & ~7) * 2 + (size_t)W * R;
  DPV_REQUIRE(smem <= 227 * 1024, "lc_planner_solve: W*R = %d*%d back-pointers exceed shared memory", W, R);
  auto kern = planner_solve_kernel<LANES, NQL>;
  DPV_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int na = (R * LANES + 31) / 32;                           // part A: LANES threads per node
  if (warps < 1) warps = 1;
  if (na + warps > 32) warps = 32 - na;                           // part B: the remaining warps, one queued node each
  DPV_REQUIRE(warps >= 1, "lc_planner_solve: R=%d leaves no warp for tie resolution", R);
  kern<<<n, 32 * (na + warps), smem, st>>>(umaps, W, R, na, p->d_theta, p->d_xz, p->d_win, ks, xz, counts);
  return DPV_OK;
}

}  // namespace

extern "C" int lc_planner_solve(LcPlanner* p, const float* umaps, int n, int32_t* ks, float* xz, int32_t* counts,
                                void* stream) {
  DPV_REQUIRE(p && umaps, "lc_planner_solve: null pointer");
  DPV_REQUIRE(n > 0, "lc_planner_solve: n=%d", n);
  int lanes = 2, warps = 16;                                      // warps = part-B warps
  if (const char* e = getenv("DPV_PLANNER_LANES")) lanes = atoi(e);                // tuning aids
  if (const char* e = getenv("DPV_PLANNER_WARPS")) warps = atoi(e);
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (p->R > 128) rc = launch_solve<2, 32>(p, umaps, n, warps, ks, xz, counts, st);     // up to 256 entries
  else if (lanes == 4) rc = launch_solve<4, 8>(p, umaps, n, warps, ks, xz, counts, st);
  else rc = launch_solve<2, 16>(p, umaps, n, warps, ks, xz, counts, st);
  if (rc != DPV_OK) return rc;
  return check_launch("lc_planner_solve");
}
