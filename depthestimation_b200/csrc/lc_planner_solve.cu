// (d3) DP planner solve kernel.  external/lcsim/lcsim/src/planning.cpp:75-126 (optimizedDesignPts),
// :162-217 (Trajectory compare).
//
// dp(i,k) = umap(k,i) + best over the edge window [lo,hi] of ray i+1, rays processed right to left.
// The reference's total order over a node's edges is
//   (unc = fl(u + unc[e]) larger) > (las = dth^2 + las[e] smaller) > (first in edge order), "no edge" first,
// and max-propagation through zero regions of the map makes exact ties in the first key COMMON (about a fifth
// of the nodes), so the second key decides real paths and must be evaluated exactly.
//
// The rays are a serial chain (W steps), so what counts is the length of the dependent instruction chain per ray
// and how many issue slots serve it.  One umap is solved by a pair of CTAs on two SMs (a cluster of two, so both
// are always co-scheduled), connected by a one-way log in global memory:
//   A. (first CTA) LANES threads per node.  The finished ray keeps, next to unc[], the maxima of its aligned
//      groups of four (qm[]).  A lane holds up to NQL interior groups of the node's window in registers; lanes 0
//      and LANES-1 also hold the two boundary groups entry by entry.  Window maximum m; fl(u + x) is monotone
//      in x, so fl(u + m) is the winning first key and the node's new unc.  The groups / boundary entries that
//      can hold a first-key tie (fl(u + .) == fl(u + m)) are counted.  If exactly one group qualifies, its lane
//      opens it; if exactly one of its entries ties, the successor is known.  A needs only unc/qm of the
//      previous ray, never the second key, so it runs ahead at its own pace and appends, per ray, unc[] and a
//      code per node (successor | bare | tie) to the log as self-describing 8-byte entries (no fence, no flag).
//   B. (second CTA) follows the log (entries prefetched, re-read until tagged).  One step and one barrier per
//      ray: it fills in the second key las of the nodes A finished, stages the next ray (independent work), and
//      resolves the tie nodes, one WARP per node: the previous ray
//      sits in the warp's registers, four entries per lane; tie flags, the second key for the tied entries,
//      and two REDUX.min give the exact (las, e) minimum with no divergence, however many entries tie.  B owns
//      the back-pointers (W*R bytes of shared memory) and unrolls the path at the end.
// Both parts are associative reductions under the reference's order, so the selected edge is exactly the one
// its sequential strict-'>' scan selects: results are bit-exact.
#include <cooperative_groups.h>

#include <cstdlib>

#include "lc_device.cuh"

namespace cg = cooperative_groups;

namespace dpv {

constexpr unsigned kNoWindow = 1u;   // lo = 1 > hi = 0
constexpr int kCodeBare = -1;        // node keeps its own value: las 0, no successor
constexpr int kCodeTie = -2;         // several edges tie on the first key: B decides

// The log A writes and B follows (global memory, LcPlanner::d_log): one 8-byte entry per (ray, node),
//   bits 0-31 unc, 32-47 code (successor | bare | tie), 48-63 tag of the solve that wrote it.
// An aligned 8-byte store is single-copy atomic and the entry is self-describing, so no flag, fence or ordering
// between entries is needed: B re-reads an entry until its tag is the current one.
__device__ __forceinline__ unsigned long long log_entry(float unc, int code, unsigned tag) {
  return (unsigned long long)__float_as_uint(unc) | ((unsigned long long)(unsigned)(code & 0xffff) << 32) |
         ((unsigned long long)tag << 48);
}
// Both sides of the log are strong (relaxed, gpu scope) operations, so the single-copy atomicity of an aligned
// 8-byte access is what the PTX memory model guarantees, not an artefact of how a weak store happens to be issued.
__device__ __forceinline__ void st_relaxed(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v));   // no clobber: the entry depends on no other memory
}
__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// A "team" is the set of threads that solves one map inside a CTA.  PACK = 1: the whole CTA (threadIdx.x, blockDim.x,
// __syncthreads(): the code is what it was).  PACK = 2 (throughput mode, lc_planner_set_pack): a CTA of 1024 threads
// holds two teams of 512, each with its own half of the dynamic shared memory and its own named barrier, so a cluster
// of two CTAs solves two maps on two SMs instead of one: the solve is a dependent chain that leaves most issue slots
// idle (part A works with 8 of its 16 warps), and two independent chains share an SM's schedulers.
constexpr int kTeamThreads = 512;
template <int PACK> __device__ __forceinline__ int team_tid() { return PACK == 1 ? (int)threadIdx.x : (int)(threadIdx.x & (kTeamThreads - 1)); }
template <int PACK> __device__ __forceinline__ int team_size() { return PACK == 1 ? (int)blockDim.x : kTeamThreads; }
template <int PACK> __device__ __forceinline__ void team_sync(int half, int count) {
  if (PACK == 1) __syncthreads();
  else asm volatile("bar.sync %0, %1;" ::"r"(half + 1), "r"(count) : "memory");
}

// ------------------------------------------------------------------------------------------------
// part A: first keys (unc) of every ray, successor where a single edge wins
// ------------------------------------------------------------------------------------------------
template <int LANES, int NQL, int PACK>
__device__ __forceinline__ void planner_part_a(unsigned char* smraw, const float* __restrict__ umap, int W, int R,
                                               const uint8_t* __restrict__ win, unsigned long long* __restrict__ lg, unsigned tag,
                                               int half) {
  constexpr int NQ = LANES * NQL;                                 // groups of four per ray
  constexpr int RP = 4 * NQ;                                      // padded ray length
  constexpr int SLOT = RP + 2 * NQ;                               // unc (pad -inf) + qm (padded: no index clamp)
  float* state = reinterpret_cast<float*>(smraw);                 // [2][SLOT]
  const int tid = team_tid<PACK>();
  const int k = tid / LANES, s = tid % LANES;
  const bool active = k < R;
  const unsigned full = 0xffffffffu;
  const int nact = (R * LANES + 31) & ~31;                        // threads of the warps that hold nodes
  if (PACK > 1 && tid >= nact) return;                            // a named barrier counts only those
  const int nth = PACK == 1 ? team_size<PACK>() : nact;
  for (int i = tid; i < 2 * SLOT; i += nth) state[i] = -INFINITY;
  team_sync<PACK>(half, nact);
  if (PACK == 1 && tid >= nact) {                                 // warps without nodes only keep the barriers company
    for (int i = W - 1; i >= 0; i--) __syncthreads();
    return;
  }
  // input pointers of this thread's node, walked right to left
  const int kc = active ? k : 0;
  const float* pu = umap + (size_t)kc * W + (W - 1);
  const uint16_t* pw = reinterpret_cast<const uint16_t*>(win) + (size_t)(W - 1) * R + kc;
  auto fetch = [&](int i, float& u, unsigned& w) {                // i = ray the pointers stand on
    u = 0.f; w = kNoWindow;
    if (active && i >= 0) {
      u = __ldg(pu);
      if (i < W - 1) w = __ldg(pw);                               // unpacked only when used
    }
    pu -= 1; pw -= R;
  };
  float u0, u1, u2;
  unsigned w0, w1, w2;
  fetch(W - 1, u0, w0); fetch(W - 2, u1, w1); fetch(W - 3, u2, w2);   // inputs of ray i-3 are requested during ray i

  for (int i = W - 1; i >= 0; i--) {
    const int par = i & 1;
    const float* uncn = state + (par ^ 1) * SLOT;                 // ray i+1
    const float* qmn = uncn + RP;
    float* unco = state + par * SLOT;                             // ray i
    const float u = u0;
    const unsigned wbits = w0;
    u0 = u1; w0 = w1; u1 = u2; w1 = w2;
    fetch(i - 3, u2, w2);
    const int lo = (int)(wbits & 0xffu), hi = (int)(wbits >> 8);
    const int qlo = lo >> 2, qhi = hi >> 2;                       // no window: lo = 1 > hi = 0 masks everything below
    // ---- loads: this lane's boundary group (lane 0: group of lo, lane LANES-1: group of hi) and its
    //      interior groups qlo+1+s, +LANES, ... < qhi
    const bool is_b = s == 0 || (s == LANES - 1 && qhi != qlo);
    const int qb = s == 0 ? qlo : qhi;
    const float4 araw = reinterpret_cast<const float4*>(uncn)[qb];
    const float* qp = qmn + qlo + 1 + s;
    const int rem = qhi - qlo - 1 - s;                            // interior group j exists iff j*LANES < rem
    float qv[NQL];
#pragma unroll
    for (int j = 0; j < NQL; j++) qv[j] = qp[j * LANES];
    const int eb = qb << 2;
    float4 a;                                                     // boundary entries inside the window, else -inf
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
    const float tmax = __fadd_rn(u, m);                           // planning.cpp:194
    const bool better = tmax > u;                                 // strict '>' against the bare node
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
    const bool owner = tot == 1 && mine == 1;                     // implies better
    int code = kCodeTie;
    if (owner) {
      int e1 = eb + __ffs(bb) - 1;                                // the tied boundary entry
      bool solved = bb != 0;
      if (!solved) {                                              // an interior group: open it
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
      if (solved) code = e1;
    }
    // ---- results; group-of-four maxima of the new unc[] (4*LANES adjacent lanes)
    const float bu = better ? tmax : u;                           // :210-217
    float gq = active ? bu : -INFINITY;
    gq = fmaxf(gq, __shfl_xor_sync(full, gq, LANES));
    gq = fmaxf(gq, __shfl_xor_sync(full, gq, 2 * LANES));
    if (active) {
      if (s == 0) {
        unco[k] = bu;
        if ((k & 3) == 0) unco[RP + (k >> 2)] = gq;
      }
      // one lane per node appends the log entry: the owner when a single candidate exists, lane 0 otherwise
      if (!better) {
        if (s == 0) st_relaxed(lg + (size_t)i * RP + k, log_entry(bu, kCodeBare, tag));
      } else if (owner || (s == 0 && tot >= 2)) {
        st_relaxed(lg + (size_t)i * RP + k, log_entry(bu, code, tag));
      }
    }
    team_sync<PACK>(half, nact);                                  // ray i complete in shared memory
  }
}

// ------------------------------------------------------------------------------------------------
// part B: second keys (las), tie nodes, back-pointers, path
// ------------------------------------------------------------------------------------------------
struct LogRay {          // what a node's thread needs of one ray
  unsigned long long entry;
  float u, th;
  unsigned w;
};

template <int NCH, int PACK>
__device__ __forceinline__ void planner_part_b(unsigned char* smraw, const float* __restrict__ umap, int W, int R,
                                               const float* __restrict__ theta, const float* __restrict__ node_xz,
                                               const uint8_t* __restrict__ win, const unsigned long long* lg,
                                               unsigned tag, int map, int32_t* __restrict__ ks,
                                               float* __restrict__ xz, int32_t* __restrict__ counts, int half) {
  constexpr int RP = 128 * NCH;
  constexpr int SLOT = 3 * RP;                                    // unc (pad -inf), las, th of one ray
  float* state = reinterpret_cast<float*>(smraw);                 // [3][SLOT] ring over rays
  float4* nin = reinterpret_cast<float4*>(state + 3 * SLOT);      // [2][RP]: u, theta, lo, hi of a ray's nodes
  int* list = reinterpret_cast<int*>(nin + 2 * RP);               // [2][RP] tie nodes of a ray
  int16_t* path = reinterpret_cast<int16_t*>(list + 2 * RP);      // [W rounded to 8]
  uint8_t* sub = reinterpret_cast<uint8_t*>(path + ((W + 7) & ~7));   // [W][R]
  __shared__ int s_nt_all[PACK][3];                               // tie counts, ring like state
  __shared__ int s_cnt_all[PACK];
  __shared__ volatile int s_gaveup_all[PACK];                     // some wait ran out of polls: the result is void
  int* s_nt = s_nt_all[half];
  int& s_cnt = s_cnt_all[half];
  volatile int& s_gaveup = s_gaveup_all[half];

  const int tid = team_tid<PACK>(), nth = team_size<PACK>();
  const int lane = tid & 31, warp = tid >> 5, nwarps = nth >> 5;
  const int k = tid;                                              // node whose log entries this thread follows
  const bool node = k < R;
  const bool node_warp = warp * 32 < R;
  const unsigned full = 0xffffffffu;

  for (int i = tid; i < 3 * SLOT; i += nth) state[i] = ((i % SLOT) < RP) ? -INFINITY : 0.f;
  if (tid < 3) s_nt[tid] = 0;
  if (tid == 0) s_gaveup = 0;
  team_sync<PACK>(half, nth);

  auto load_ray = [&](int i) {                                    // may come back with an old tag: checked at use
    LogRay r;
    r.entry = 0ull; r.u = 0.f; r.th = 0.f; r.w = kNoWindow;
    if (node && i >= 0) {
      r.entry = ld_relaxed(lg + (size_t)i * RP + k);
      r.u = __ldg(umap + (size_t)k * W + i);
      r.th = __ldg(theta + (size_t)i * R + k);
      if (i < W - 1) r.w = __ldg(reinterpret_cast<const uint16_t*>(win) + (size_t)i * R + k);
    }
    return r;
  };
  // Stage ray i for the step that will finish it: unc and theta into its ring slot, the node inputs, the list of
  // tie nodes.  Returns the node's code (successor | bare | tie) and theta for the thread to keep.
  auto stage_ray = [&](int i, LogRay r, int slot, int& code_out, float& th_out) {
    if (node) {                                                   // A has not got here yet: wait for this node's entry
      // Part B waits here on most rays (A, the producer, sets the pace), so this loop is as tight as it gets: sleep,
      // re-read, and a poll counter in a register.  Bounded per wait: A is co-scheduled (same cluster) and produces a
      // ray in about a microsecond; 2^15 polls (~15 ms) without the entry mean that A faulted or was never
      // launched.  Instead of spinning the GPU forever the thread raises a flag that voids the map's result (thread 0
      // turns the count into -1 after the last ray), takes a "bare node" entry in place of the missing one and goes
      // on; a dead producer therefore costs one such wait per ray (W x ~15 ms), then an error -- not a hung device.
      // Nothing about it sits in the per-ray path, and it is written the one way that keeps this kernel at 62
      // registers: an abort check per ray cost 6 %, the same wait out of line 8 %, a separate block after the loop
      // 10 %, filling in the remaining entries from here 11 % (each of them 64 registers); as written the bound is
      // free (0.743 ms with it and the strong log stores, 0.758 without, 0.762 before).
      unsigned polls = 0;
      while ((unsigned)(r.entry >> 48) != tag) {
        if (++polls > (1u << 15)) {
          s_gaveup = 1;
          r.entry = log_entry(0.f, kCodeBare, tag);
          break;
        }
        __nanosleep(32);
        r.entry = ld_relaxed(lg + (size_t)i * RP + k);
      }
    }
    const int code = (int)(short)(unsigned short)(r.entry >> 32);
    const int lo = (int)(r.w & 0xffu), hi = (int)(r.w >> 8);
    if (node) {
      float* s = state + slot * SLOT;
      s[k] = __uint_as_float((unsigned)r.entry);
      s[2 * RP + k] = r.th;
      nin[(i & 1) * RP + k] = make_float4(r.u, r.th, __int_as_float(lo), __int_as_float(hi));
    }
    const bool tie = node && code == kCodeTie;
    const unsigned pb = __ballot_sync(full, tie);
    if (pb) {
      int base = 0;
      if (lane == 0) base = atomicAdd(&s_nt[slot], __popc(pb));
      base = __shfl_sync(full, base, 0);
      if (tie) list[(i & 1) * RP + base + __popc(pb & ((1u << lane) - 1u))] = k;
    }
    code_out = code;
    th_out = r.th;
  };

  // ring slots: sa = ray i+1 (complete), sb = ray i (staged, finished in this step), sc_ = ray i-1 (staged in this step)
  int sa = W % 3, sb = (W - 1) % 3, sc_ = (W + 1) % 3;
  LogRay nxt = load_ray(W - 2), nx2 = load_ray(W - 3);            // entries are requested two rays before their staging
  int code = kCodeBare;
  float tn = 0.f;
  if (node_warp) stage_ray(W - 1, load_ray(W - 1), sb, code, tn);
  team_sync<PACK>(half, nth);

  for (int i = W - 1; i >= 0; i--) {
    const float* sp = state + sa * SLOT;                          // ray i+1: unc, las, th
    float* sc = state + sb * SLOT;                                // ray i
    uint8_t* subi = sub + (size_t)i * R;
    if (tid == 0) s_nt[sa] = 0;                                   // its ties were read a step ago; pushed again next step
    if (node_warp) {
      // ---- nodes A finished: second key from the successor (planning.cpp:197-198); bare nodes: las 0
      if (node) {
        if (code >= 0) {
          const float dth = __fsub_rn(sp[2 * RP + code], tn);
          sc[RP + k] = __fadd_rn(__fmul_rn(dth, dth), sp[RP + code]);
          subi[k] = (uint8_t)code;
        } else if (code == kCodeBare) {
          sc[RP + k] = 0.f;
          subi[k] = 255;
        }
      }
      // ---- stage ray i-1 for the next step (independent of everything this step computes)
      const LogRay r = nxt;
      nxt = nx2;
      nx2 = load_ray(i - 3);
      if (i > 0) stage_ray(i - 1, r, sc_, code, tn);
    }
    // ---- tie nodes of ray i, one warp each, handed out from the last warp down (the first warps hold the nodes)
    const int nt = s_nt[sb];
    const float4* ninb = nin + (i & 1) * RP;
    const int* listb = list + (i & 1) * RP;
    for (int j = nwarps - 1 - warp; j < nt; j += nwarps) {
      const int rk = listb[j];
      const float4 in = ninb[rk];
      const float ru = in.x, rtn = in.y, rtmax = sc[rk];
      const int rlo = __float_as_int(in.z), rhi = __float_as_int(in.w);
      float tl_best = INFINITY;
      int e_best = 0x7fffffff;
#pragma unroll
      for (int c = 0; c < NCH; c++) {                             // chunks of 128 entries, four per lane
        const int e = c * 128 + lane * 4;
        const float4 v = reinterpret_cast<const float4*>(sp)[c * 32 + lane];
        const float4 l4 = reinterpret_cast<const float4*>(sp + RP)[c * 32 + lane];
        const float4 t4 = reinterpret_cast<const float4*>(sp + 2 * RP)[c * 32 + lane];
        const float vv[4] = {v.x, v.y, v.z, v.w}, tt[4] = {t4.x, t4.y, t4.z, t4.w}, ll[4] = {l4.x, l4.y, l4.z, l4.w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const float dth = __fsub_rn(tt[q], rtn);                // :197-198
          const float tl = __fadd_rn(__fmul_rn(dth, dth), ll[q]);
          const bool tie = e + q >= rlo && e + q <= rhi && __fadd_rn(ru, vv[q]) == rtmax;
          if (tie && tl < tl_best) { tl_best = tl; e_best = e + q; }      // increasing e: first minimum stays
        }
      }
      const unsigned tmin = __reduce_min_sync(full, __float_as_uint(tl_best));   // las >= +0: bit order = value order
      const int be = __reduce_min_sync(full, __float_as_uint(tl_best) == tmin ? e_best : 0x7fffffff);
      if (lane == 0) {
        sc[RP + rk] = __uint_as_float(tmin);
        subi[rk] = (uint8_t)be;
      }
    }
    team_sync<PACK>(half, nth);
    const int t = sa; sa = sb; sb = sc_; sc_ = t;                 // ray i+1's slot is free for ray i-2
  }
  const int pc = sa;                                              // ray 0 after the last rotation
  // ray 0 is in buffer pc.  Best start against a default (0,0,null) trajectory (planning.cpp:99-104)
  // (the thread index and team size are read again here instead of being kept in registers across the ray loop, and
  // for PACK = 1 with the built-in variables' own types: either difference puts the kernel at 64 registers and into
  // its slow schedule)
  if ((PACK == 1 ? threadIdx.x : (unsigned)team_tid<PACK>()) == 0) {
    const float* u0 = state + pc * 3 * RP;
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
    if (s_gaveup) cnt = -1;              // a wait ran out of polls: the path is void (ks -1, points 0 below)
    s_cnt = cnt;
    if (counts) counts[map] = cnt;
  }
  team_sync<PACK>(half, kTeamThreads);
  const int cnt = s_cnt;
  const unsigned first = PACK == 1 ? threadIdx.x : (unsigned)team_tid<PACK>();
  const unsigned stride = PACK == 1 ? blockDim.x : (unsigned)kTeamThreads;
  for (int i = first; i < W; i += stride) {
    int kk = i < cnt ? path[i] : -1;
    if (ks) ks[(size_t)map * W + i] = kk;
    if (xz) {
      float x = 0.f, z = 0.f;
      if (kk >= 0) {
        const float2 p = *reinterpret_cast<const float2*>(node_xz + ((size_t)i * R + kk) * 2);
        x = p.x; z = p.y;
      }
      *reinterpret_cast<float2*>(xz + ((size_t)map * W + i) * 2) = make_float2(x, z);
    }
  }
}

// grid = 2 CTAs per umap, launched as clusters of two so that the producer and its consumer are co-resident
template <int LANES, int NQL, bool DEVTAG, int PACK>
__device__ __forceinline__ void planner_solve_body(const float* __restrict__ umaps, int W, int R, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, unsigned long long* log_base,
                     size_t log_per_map, unsigned host_tag, unsigned* __restrict__ epoch, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts, int nmaps, unsigned team_smem) {
  constexpr int RP = 4 * LANES * NQL;
  static_assert(RP % 128 == 0, "ray length must fill whole warps in part B");
  static_assert(PACK == 1 || !DEVTAG, "the device-counter tag variant is one map per CTA pair");
  extern __shared__ __align__(16) unsigned char smraw_all[];
  const int half = PACK == 1 ? 0 : (int)(threadIdx.x / kTeamThreads);
  unsigned char* smraw = PACK == 1 ? smraw_all : smraw_all + (size_t)half * team_smem;
  const int map = PACK == 1 ? (int)(blockIdx.x >> 1) : (int)(blockIdx.x >> 1) * PACK + half;
  if (PACK > 1 && map >= nmaps) return;                           // an odd number of maps: the last pair's second team
  const float* umap = umaps + (size_t)map * R * W;
  unsigned long long* lg = log_base + (size_t)map * log_per_map;
  // The tag of this solve's log entries.  Launched directly, it is a kernel parameter the host counts up (1..32767;
  // as a constant-bank operand it costs the per-entry code nothing).  Captured in a CUDA graph the parameter would be
  // frozen, so that variant takes it from the handle's device-side solve counter instead (32768..65535: disjoint
  // from the host's range, and the last CTA of the grid to finish advances the counter after every CTA has read it).
  // Either way it is never 0 = "written by no solve" and differs from the previous solve's tag, whose entries are
  // the only stale ones a reader can meet (every solve rewrites all entries of its maps).
  const unsigned tag = DEVTAG ? 32768u + (*reinterpret_cast<volatile unsigned*>(epoch) & 32767u) : (host_tag & 0xffffu);
  if ((blockIdx.x & 1) == 0) {
    if (host_tag >> 31) return;                                   // test hook (DPV_PLANNER_TEST_NO_PRODUCER): A never writes
    planner_part_a<LANES, NQL, PACK>(smraw, umap, W, R, win, lg, tag, half);
  }
  else planner_part_b<RP / 128, PACK>(smraw, umap, W, R, theta, node_xz, win, lg, tag, map, ks, xz, counts, half);
  if (DEVTAG) {
    __syncthreads();
    if (threadIdx.x == 0) {
      if (atomicAdd(epoch + 1, 1u) == gridDim.x - 1) {
        epoch[1] = 0;
        epoch[0] = epoch[0] + 1;
      }
    }
  }
}

template <int LANES, int NQL, bool DEVTAG>
__global__ void __launch_bounds__(1024)
planner_solve_kernel(const float* __restrict__ umaps, int W, int R, const float* __restrict__ theta,
                     const float* __restrict__ node_xz, const uint8_t* __restrict__ win, unsigned long long* log_base,
                     size_t log_per_map, unsigned host_tag, unsigned* __restrict__ epoch, int32_t* __restrict__ ks,
                     float* __restrict__ xz, int32_t* __restrict__ counts) {
  planner_solve_body<LANES, NQL, DEVTAG, 1>(umaps, W, R, theta, node_xz, win, log_base, log_per_map, host_tag, epoch, ks, xz,
                                            counts, 0, 0u);
}

// two maps per CTA pair (lc_planner_set_pack(2)): teams of 512 threads, see team_sync
template <int LANES, int NQL>
__global__ void __launch_bounds__(1024)
planner_solve_pack2_kernel(const float* __restrict__ umaps, int W, int R, const float* __restrict__ theta,
                           const float* __restrict__ node_xz, const uint8_t* __restrict__ win, unsigned long long* log_base,
                           size_t log_per_map, unsigned host_tag, unsigned* __restrict__ epoch, int32_t* __restrict__ ks,
                           float* __restrict__ xz, int32_t* __restrict__ counts, int nmaps, unsigned team_smem) {
  planner_solve_body<LANES, NQL, false, 2>(umaps, W, R, theta, node_xz, win, log_base, log_per_map, host_tag, epoch, ks, xz,
                                           counts, nmaps, team_smem);
}

}  // namespace dpv

using namespace dpv;

namespace {

template <int LANES, int NQL>
int launch_solve(LcPlanner* p, const float* umaps, int n, int warps, int32_t* ks, float* xz, int32_t* counts,
                 cudaStream_t st) {
  static_assert(LANES >= 2, "the two boundary groups of a window are held by different lanes");
  const int W = p->W, R = p->R;
  constexpr int RP = 4 * LANES * NQL;
  DPV_REQUIRE(R <= RP, "lc_planner_solve: R=%d does not fit the %d-entry kernel", R, RP);
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) {
    (void)cudaGetLastError();
    cap = cudaStreamCaptureStatusNone;
  }
  const bool capturing = cap != cudaStreamCaptureStatusNone;
  // the log A writes and B follows: one per stream (solves on different streams may overlap on the GPU), grown on
  // demand, kept in the handle.  The entry in the map is only touched under the lock; its address is stable.
  const size_t per_map = (size_t)W * RP;                          // entries per umap
  LcPlanner::Log* lgp;
  {
    std::lock_guard<std::mutex> lock(p->mu);
    lgp = &p->logs[(void*)st];
    if (lgp->log_bytes < per_map * n * 8 || !lgp->d_epoch) {
      // cudaMalloc / cudaFree are not allowed while the stream is being captured: size the log with one solve of
      // this many maps before the capture (pipeline.py does: an eager step precedes every capture)
      DPV_REQUIRE(!capturing, "lc_planner_solve: the log of this stream must grow (%d maps) but the stream is being "
                  "captured -- run one solve of this size on it before the capture", n);
      if (lgp->d_log) DPV_CUDA(cudaFree(lgp->d_log));
      lgp->d_log = nullptr;
      lgp->log_bytes = 0;
      DPV_CUDA(cudaMalloc(&lgp->d_log, per_map * n * 8));
      lgp->log_bytes = per_map * n * 8;
      lgp->last_n = 0;
      if (!lgp->d_epoch) {
        DPV_CUDA(cudaMalloc(&lgp->d_epoch, 2 * sizeof(unsigned)));
        DPV_CUDA(cudaMemsetAsync(lgp->d_epoch, 0, 2 * sizeof(unsigned), st));
      }
    }
  }
  LcPlanner::Log& L = *lgp;
  if (n != L.last_n) {
    // a different number of maps than the previous solve: the regions of maps that sat out keep older tags, which a
    // wrapped-around 16-bit tag could meet again -- back to tag 0 = "written by no solve" everywhere
    DPV_CUDA(cudaMemsetAsync(L.d_log, 0, L.log_bytes, st));
    L.last_n = n;
  }
  const size_t smem_b = (size_t)9 * RP * 4 + (size_t)2 * RP * 16 + (size_t)2 * RP * 4 + (size_t)((W + 7) & ~7) * 2 +
                        (size_t)W * R;
  const size_t smem_a = (size_t)2 * (RP + 2 * (RP / 4)) * 4;
  const size_t smem = smem_a > smem_b ? smem_a : smem_b;
  DPV_REQUIRE(smem + 64 <= 227 * 1024,   /* + the kernel's static shared memory */ "lc_planner_solve: W*R = %d*%d back-pointers exceed shared memory", W, R);
  L.host_epoch = L.host_epoch % 32767 + 1;                        // 1..32767
  unsigned host_tag = (unsigned)L.host_epoch;
  // Captured in a CUDA graph the tag parameter is frozen, so every replay would meet its own previous entries as
  // valid ones.  Two ways out: clear this solve's log region inside the graph (a memset node of W*RP*8 bytes per map,
  // ~6 us at C2) and keep the parameter-tag kernel, or the device-counter variant (no memset; measured 63 us slower
  // per solve: its two extra registers).  The memset is the default; DPV_PLANNER_DEVTAG=1 selects the counter.
  // Frozen tags come from 32768..65535, so a directly launched solve (1..32767) never takes a replay's entries for
  // its own.
  static const bool want_devtag = getenv("DPV_PLANNER_DEVTAG") && atoi(getenv("DPV_PLANNER_DEVTAG")) != 0;
  const bool devtag = capturing && want_devtag;
  if (capturing && !devtag) {
    DPV_CUDA(cudaMemsetAsync(L.d_log, 0, per_map * n * 8, st));
    host_tag = 32768u + ((unsigned)L.host_epoch & 32767u);
  }
  if (const char* e = getenv("DPV_PLANNER_TEST_NO_PRODUCER"))    // exercises part B's bounded wait (tests only)
    if (atoi(e) != 0 && !devtag) host_tag |= 0x80000000u;
  int threads = ((R * LANES + 31) / 32) * 32;                     // part A: LANES threads per node
  if (warps < 1) warps = 1;                                       // part B: one warp per tie node
  if (threads < 32 * warps) threads = 32 * warps;
  if (threads < ((R + 31) / 32) * 32) threads = ((R + 31) / 32) * 32;            // B follows one node per thread
  DPV_REQUIRE(threads <= 1024, "lc_planner_solve: R=%d needs %d threads", R, threads);
  // two maps per CTA pair (lc_planner_set_pack) when both teams' shared memory fits and a team is 512 threads
  const size_t team_smem = (smem + 127) & ~(size_t)127;
  const bool pack2 = p->pack == 2 && !devtag && n >= 2 && threads <= kTeamThreads && 2 * team_smem + 128 <= 227 * 1024;
  auto kern = devtag ? planner_solve_kernel<LANES, NQL, true> : planner_solve_kernel<LANES, NQL, false>;
  auto kern2 = planner_solve_pack2_kernel<LANES, NQL>;
  if (pack2) DPV_CUDA(cudaFuncSetAttribute(kern2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * team_smem)));
  else DPV_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(pack2 ? 2 * ((n + 1) / 2) : 2 * n));
  cfg.blockDim = dim3((unsigned)(pack2 ? 2 * kTeamThreads : threads));
  cfg.dynamicSmemBytes = pack2 ? 2 * team_smem : smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributePriority;                       // see high_priority(): 48 CTAs, 0.76 ms of serial work
  attr[1].val.priority = high_priority();
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  if (pack2)
    DPV_CUDA(cudaLaunchKernelEx(&cfg, kern2, umaps, W, R, (const float*)p->d_theta, (const float*)p->d_xz,
                                (const uint8_t*)p->d_win, (unsigned long long*)L.d_log, per_map, host_tag, L.d_epoch, ks,
                                xz, counts, n, (unsigned)team_smem));
  else
    DPV_CUDA(cudaLaunchKernelEx(&cfg, kern, umaps, W, R, (const float*)p->d_theta, (const float*)p->d_xz,
                                (const uint8_t*)p->d_win, (unsigned long long*)L.d_log, per_map, host_tag, L.d_epoch, ks, xz,
                                counts));
  return DPV_OK;
}

}  // namespace

extern "C" int lc_planner_set_pack(LcPlanner* p, int pack) {
  DPV_REQUIRE(p, "lc_planner_set_pack: null pointer");
  DPV_REQUIRE(pack == 1 || pack == 2, "lc_planner_set_pack: pack must be 1 or 2, got %d", pack);
  p->pack = pack;
  return DPV_OK;
}

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
  else if (lanes == 2) rc = launch_solve<2, 16>(p, umaps, n, warps, ks, xz, counts, st);
  else rc = launch_solve<4, 8>(p, umaps, n, warps, ks, xz, counts, st);
  if (rc != DPV_OK) return rc;
  return check_launch("lc_planner_solve");
}
