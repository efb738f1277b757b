// (d4) One LC iteration's sensing + Bayesian update in a single pass over the DPV.
//   models/models.py:1064-1122 (sense each curtain, clamp, multiply, renormalise, spread, clamp, log)
//   with lc/light_curtain.py:847-888 + :228-252 (likelihood) and dprocessor.cpp:806-993 (image formation)
//   evaluated per pixel in registers.  Compiled with -fmad=false for the geometry (see lc_sense.cuh).
//
// T adjacent lanes own one pixel: its DPV column p[D] and the current curtain's likelihood column g[D] live in
// registers as f32x2 pairs (straight-line code over the bins; at the simulated baseline of 0.2 m the curtain thickness
// grows to metres, so the Gaussian really is dense over the bins and is evaluated for every bin with one
// MUFU ex2 each).  Multiply, box-spread and renormalise are all homogeneous in the DPV, so the
// per-curtain and per-spread renormalisations of the reference collapse into one scale at the end
// (identical up to fp32 rounding; an extra rescale is inserted only if the running sum leaves fp32 range).
// The volume may be kept as log-probability (the reference's format) or as clamped probability between
// the iterations of one lc_process call (flags), which removes exp/log from the inner iterations.
#include <cstdlib>

#include "lc_sense.cuh"

namespace dpv {

#ifndef DPV_FUSE_THREADS
#define DPV_FUSE_THREADS 128
#endif
constexpr int kFuseThreads = DPV_FUSE_THREADS;
constexpr int kMaxCurtains = 8;
#ifndef DPV_FUSE_BLOCKS16
#define DPV_FUSE_BLOCKS16 8
#endif
constexpr int kFuseBlocks16 = DPV_FUSE_BLOCKS16 * 128 / DPV_FUSE_THREADS;   // resident CTAs per SM asked for when a lane holds 16 bins

__device__ __forceinline__ float2 splat2(float a) { return make_float2(a, a); }

// sum over the T lanes that share one pixel (T is a power of two; lanes are adjacent)
template <int T>
__device__ __forceinline__ float group_sum(float v, unsigned gmask) {
#pragma unroll
  for (int off = T / 2; off > 0; off >>= 1) v += __shfl_xor_sync(gmask, v, off);
  return v;
}

// A lane's DPT bins as DPT/2 register pairs (packed f32x2 arithmetic: one FMUL2/FADD2/FFMA2 per two bins).
template <int NP>
struct Col {
  float2 v[NP];
  __device__ __forceinline__ float get(int d) const { return (d & 1) ? v[d >> 1].y : v[d >> 1].x; }
  __device__ __forceinline__ void set(int d, float x) {
    if (d & 1) v[d >> 1].y = x;
    else v[d >> 1].x = x;
  }
};

// One box-sum pass along depth over a column that is split over T lanes (DPT bins each, lane t owns bins
// [t*DPT, (t+1)*DPT)).  q[d] = sum_{|j|<=HALF} p[d+j], zero outside [0,D).  img_utils.py:408-436.
// Bins at and beyond D hold 0 on entry and on exit, so the neighbour lane's halo needs no bound check.
template <int DPT, int T, int HALF, bool FULL>
__device__ __forceinline__ void spread_pass(Col<DPT / 2>& p, int d0, int D, int t, int lane, unsigned gmask) {
  constexpr int NH = HALF > 0 ? HALF : 1;
  float lo_h[NH], hi_h[NH];                                    // lo_h[j] = p[d0-1-j], hi_h[j] = p[d0+DPT+j]
#pragma unroll
  for (int j = 0; j < HALF; j++) {
    float from_prev = 0.f, from_next = 0.f;
    if (T > 1) {
      from_prev = __shfl_sync(gmask, p.get(DPT - 1 - j), (lane - 1) & 31);
      from_next = __shfl_sync(gmask, p.get(j), (lane + 1) & 31);
    }
    lo_h[j] = (T > 1 && t > 0) ? from_prev : 0.f;
    hi_h[j] = (T > 1 && t < T - 1) ? from_next : 0.f;
  }
  auto at = [&](int i) -> float {                              // i is a compile-time constant after unrolling
    return i < 0 ? lo_h[(-1 - i) < NH ? (-1 - i) : 0] : (i >= DPT ? hi_h[(i - DPT) < NH ? (i - DPT) : 0] : p.get(i));
  };
  Col<DPT / 2> q;
  if constexpr (HALF == 2) {
    // pair sums S_k = p[2k] + p[2k+1]:  q[2k] = S_{k-1} + S_k + p[2k+2],  q[2k+1] = p[2k-1] + S_k + S_{k+1}
    float S[DPT / 2 + 2];
#pragma unroll
    for (int k = -1; k <= DPT / 2; k++) S[k + 1] = at(2 * k) + at(2 * k + 1);
#pragma unroll
    for (int k = 0; k < DPT / 2; k++) {
      q.v[k].x = (S[k] + S[k + 1]) + at(2 * k + 2);
      q.v[k].y = (S[k + 1] + S[k + 2]) + at(2 * k - 1);
    }
  } else {
#pragma unroll
    for (int d = 0; d < DPT; d++) {
      float acc = at(d);
#pragma unroll
      for (int j = 1; j <= HALF; j++) acc += at(d - j) + at(d + j);
      q.set(d, acc);
    }
  }
#pragma unroll
  for (int d = 0; d < DPT; d++) p.set(d, (FULL || d0 + d < D) ? q.get(d) : 0.f);
}

template <int DPT, int T, int HALF, bool FULL>
__global__ void __launch_bounds__(kFuseThreads, DPT <= 16 ? kFuseBlocks16 : 4 * 128 / kFuseThreads)
sense_fuse_kernel(LcDeviceView dv, const float* __restrict__ vol_in, const float* __restrict__ depth,
                  const float* __restrict__ planes, const float* __restrict__ noise, int nc,
                  const float* __restrict__ d_candi, int D, int spread_iter, int flags,
                  float* __restrict__ vol_out, float* __restrict__ mean_out, int rows_per_cta) {
  constexpr int NP = DPT / 2;
  constexpr int STR = DPT + 4;      // lane slices of d_candi, padded: 16-byte aligned and 4 banks apart
  __shared__ __align__(16) float dc[STR * T];
  __shared__ float a_lut[256];      // mapping(i / 255) for the 256 intensities the simulator emits (lc_sense.cuh)
  for (int d = threadIdx.x; d < DPT * T; d += kFuseThreads)           // DPT is a power of two: shifts, no division
    dc[(d / DPT) * STR + (d % DPT)] = d < D ? d_candi[d] : 0.f;
  for (int i = threadIdx.x; i < 4 * T; i += kFuseThreads) dc[(i / 4) * STR + DPT + (i % 4)] = 0.f;   // the pads
  for (int i = threadIdx.x; i < 256; i += kFuseThreads) a_lut[i] = mapping_of((float)i);
  __syncthreads();
  // grid = (column chunks, rows, samples): no per-thread integer division, all pixel indices in 32 bits.
  // Every lane of a warp stays alive (full-mask shuffles, no divergence bookkeeping): columns past the end of
  // the row are clamped to the last one and not stored
  const unsigned HW = (unsigned)dv.H * (unsigned)dv.W;
  const int u_raw = (int)(blockIdx.x * (kFuseThreads / T) + threadIdx.x / T);
  const bool live = u_raw < dv.W;
  const int u = live ? u_raw : dv.W - 1;
  const int lane = threadIdx.x & 31;
  const int t = lane & (T - 1);
  const int gbase = lane & ~(T - 1);
  constexpr unsigned gmask = 0xffffffffu;
  const int d0 = t * DPT;
  // this lane's slice of d_candi as a 32-bit shared-memory address held in a register (a generic pointer is
  // re-derived from the CTA's shared window inside every loop that is not unrolled)
  const unsigned dc_addr = (unsigned)__cvta_generic_to_shared(dc + t * STR);
  auto dc4_at = [&](int i) -> float4 {                       // i is a compile-time constant after unrolling
    float4 q;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w) : "r"(dc_addr + 16u * i));
    return q;
  };
  const int b = blockIdx.z;
  auto valid = [&](int d) -> bool { return FULL || d0 + d < D; };

  // ---- what depends on the column only (lc_sense.cuh): this lane's curtain (t < nc: one sensing round covers all
  // curtains) through this lane's column.  Once per CTA instead of once per pixel.
  const bool col_form = nc <= T;
  ColGeom cg;
  cg.kind = 2;
  if (col_form && t < nc) {
    const float4* P4 = reinterpret_cast<const float4*>(planes + ((size_t)(b * nc + t) * dv.W + u) * 12);
    float Pl[12];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float4 q = __ldg(P4 + k);
      Pl[4 * k] = q.x; Pl[4 * k + 1] = q.y; Pl[4 * k + 2] = q.z; Pl[4 * k + 3] = q.w;
    }
    cg = column_geometry(Pl, dv.nn_x[u]);
  }

  const int v_end = min(dv.H, ((int)blockIdx.y + 1) * rows_per_cta);
#pragma unroll 1
  for (int v = (int)blockIdx.y * rows_per_cta; v < v_end; v++) {
  const unsigned hw = (unsigned)v * (unsigned)dv.W + (unsigned)u;
  const size_t pix = (size_t)b * HW + hw;
  const size_t base = ((size_t)b * D + d0) * HW + hw;
  // ---- issue the volume loads first; the curtain geometry below overlaps their latency
  Col<NP> p;
  {
    const float* src = vol_in + base;                        // bumped by one plane per bin: one 64-bit add each
#pragma unroll
    for (int d = 0; d < DPT; d++, src = bump(src, HW)) p.set(d, valid(d) ? ld_stream(src) : 0.f);
  }
  const float gt = depth[pix];

  int c0 = 0;
  int unchecked = (flags & 4) ? 0 : 1;                       // multiplications since the column's total was last seen
#pragma unroll 1
  do {
    // ---- sensing: in rounds of T curtains, lane t of the group senses curtain c0 + t (all lanes busy, same code)
    Lik mine;
    mine.z = 0.f; mine.den = 1.f; mine.A = 0.f; mine.Bu = 0.f;      // idle lane: never broadcast
    if (c0 + t < nc) {
      const int cu = b * nc + c0 + t;
      const float* nzp = noise ? noise + (size_t)cu * HW + hw : nullptr;
      Ret r;
      if (cg.kind == 0) {
        r = sense_pixel_col(dv, cg, u, v, gt, nzp);
      } else if (cg.kind == 1) {                             // column without a design point: no return
        r.x = r.y = r.z = NAN;
        r.inten = 0.f;
        r.thick = NAN;
      } else {
        const float4* P4 = reinterpret_cast<const float4*>(planes + ((size_t)cu * dv.W + u) * 12);
        float Pl[12];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          const float4 q = __ldg(P4 + k);
          Pl[4 * k] = q.x; Pl[4 * k + 1] = q.y; Pl[4 * k + 2] = q.z; Pl[4 * k + 3] = q.w;
        }
        r = sense_pixel(dv, Pl, u, v, gt, nzp);
      }
      mine = lik_params(r, gt, D, a_lut);
    }
    if (c0 == 0) {                                           // models.py:1097 (after the loads have landed)
      if (flags & 4) {
        // the caller's own output of the previous iteration: already inside [eps, 1]
      } else if (flags & 1) {
#pragma unroll
        for (int d = 0; d < DPT; d++) p.set(d, valid(d) ? fminf(fmaxf(p.get(d), eps_floor()), 1.f) : 0.f);
      } else {
#pragma unroll
        for (int d = 0; d < DPT; d++) p.set(d, valid(d) ? fminf(fmaxf(exp2f_approx(p.get(d) * 1.4426950408889634f), eps_floor()), 1.f) : 0.f);
      }
    }
    // g[d] = exp(-(dc-z)^2 / den) as ex2((dc-z)^2 * kk), kk = -log2(e)/den  (img_utils.py:183-186), divided once by
    // the sensing lane; den = 0 -> kk = -inf: g = 0 (NaN at dc == z), as in the reference
    // (an un-sensed pixel has den == 0: answered without the division's special-case path)
    const bool den_pos = mine.den > 0.f;
    const float kk_q = -1.4426950408889634f / opaque(den_pos ? mine.den : 1.f);
    const float mine_kk = den_pos ? kk_q : (mine.den == 0.f ? -INFINITY : NAN);   // den = 2 sigma^2 is never -0
    const int jn = min(T, nc - c0);
    unsigned slow_mask = 0;                                  // curtains of this round left to the fix-up below
#pragma unroll 1
    for (int j = 0; j < jn; j++) {
      float Lz = mine.z, Lkk = mine_kk, LA = mine.A, LBu = mine.Bu;
      if (T > 1) {
        Lz = __shfl_sync(gmask, mine.z, gbase + j);
        Lkk = __shfl_sync(gmask, mine_kk, gbase + j);
        LA = __shfl_sync(gmask, mine.A, gbase + j);
        LBu = __shfl_sync(gmask, mine.Bu, gbase + j);
      }
      // ---- likelihood of this curtain over the depth bins (mixed_model, light_curtain.py:247-252)
      const float2 kk2 = splat2(Lkk), nz2 = splat2(-Lz);
      Col<NP> g;
      float2 a0 = splat2(0.f), a1 = splat2(0.f);
#pragma unroll
      for (int k = 0; k < NP; k += 2) {
        const float4 q = dc4_at(k >> 1);
        float2 e0 = __fadd2_rn(make_float2(q.x, q.y), nz2), e1 = __fadd2_rn(make_float2(q.z, q.w), nz2);
        e0 = __fmul2_rn(__fmul2_rn(e0, e0), kk2);
        e1 = __fmul2_rn(__fmul2_rn(e1, e1), kk2);
        g.v[k] = make_float2(valid(2 * k) ? exp2f_approx(e0.x) : 0.f, valid(2 * k + 1) ? exp2f_approx(e0.y) : 0.f);
        g.v[k + 1] =
            make_float2(valid(2 * k + 2) ? exp2f_approx(e1.x) : 0.f, valid(2 * k + 3) ? exp2f_approx(e1.y) : 0.f);
        a0 = __fadd2_rn(a0, g.v[k]);
        a1 = __fadd2_rn(a1, g.v[k + 1]);
      }
      const float sgl = (a0.x + a0.y) + (a1.x + a1.y);       // this lane's share of sum(g)
      const float sg = group_sum<T>(sgl, gmask);
      // sg == 0 or NaN: g/sg is NaN -> -1 in every bin -> constant mixture -> uniform likelihood -> no-op.
      // A SENSED pixel whose sum is (nearly) flushed is skipped here too and remembered: see the fix-up below.
      bool ok = sg >= 1e-26f;
#ifndef DPV_FUSE_NO_DENORMAL_PATH
      slow_mask |= (unsigned)(!ok && Lkk > -INFINITY) << j;   // kk is -inf / NaN for an un-sensed pixel, else < 0
#endif
      // m = soft * A + uni * B before its clamp at 0 (kept signed: the floor eps*sm below is positive, so
      // max(m, eps*sm) == max(max(m, 0), eps*sm)); sum(max(m, 0)) = (sum(m) + sum|m|) / 2 with sum(m) in closed form
      const float As = LA * rcp_approx(sg);
      const float2 As2 = splat2(As), Bu2 = splat2(LBu);
      float b0 = 0.f, b1 = 0.f;
      int nvalid = 0;
#pragma unroll
      for (int k = 0; k < NP; k++) {
        float2 m = __ffma2_rn(g.v[k], As2, Bu2);
        if (!valid(2 * k)) m.x = 0.f; else nvalid++;
        if (!valid(2 * k + 1)) m.y = 0.f; else nvalid++;
        g.v[k] = m;
        b0 += fabsf(m.x);
        b1 += fabsf(m.y);
      }
      const float sml = 0.5f * (fmaf(As, sgl, (float)nvalid * LBu) + (b0 + b1));
      const float sm = group_sum<T>(sml, gmask);
      ok = ok && (sm > 0.f);                                 // 0/0 in every bin -> NaN -> 0 -> eps: uniform again
      if (ok) {
        // models.py:1109-1112: p *= clamp(m / sm, eps, 1), renormalise.  Everything here is homogeneous in p,
        // so the common factor 1/sm is left to the final renormalisation: p *= max(m, eps * sm); m <= sm
        // already bounds the ratio by 1.
        const float esm = eps_floor() * sm;
#pragma unroll
        for (int k = 0; k < NP; k++)
          p.v[k] = __fmul2_rn(p.v[k], make_float2(fmaxf(g.v[k].x, esm), fmaxf(g.v[k].y, esm)));
      }
      // One multiplication shrinks a column's total by at most eps (and grows it by at most sm <= 3), so a column
      // that starts normalised (largest bin >= 1/D) stays inside fp32 range over two of them; its total is looked
      // at, and rescaled if it left [1e-12, 1e12], after every second one (after the first one too when the input
      // is not known to be normalised).
      if (++unchecked >= 2) {
        unchecked = 0;
        a0 = splat2(0.f); a1 = splat2(0.f);
#pragma unroll
        for (int k = 0; k < NP; k += 2) {
          a0 = __fadd2_rn(a0, p.v[k]);
          a1 = __fadd2_rn(a1, p.v[k + 1]);
        }
        const float s = group_sum<T>((a0.x + a0.y) + (a1.x + a1.y), gmask);
        if (s < 1e-12f || s > 1e12f) {
          const float2 inv2 = splat2(1.f / s);
#pragma unroll
          for (int k = 0; k < NP; k++) p.v[k] = __fmul2_rn(p.v[k], inv2);
        }
      }
    }
#ifndef DPV_FUSE_NO_DENORMAL_PATH
    // ---- fix-up for thin sheets (rare; warp-uniform branches, every lane takes part in the shuffles).
    // When the sensed surface lies between (or outside) the depth bins and sigma is small, the whole Gaussian falls
    // below 2^-126: ex2.approx.ftz flushes it and the loop above skipped the curtain, while the reference's exp()
    // keeps the denormals and normalises them (g / sum(g) = 1 at the nearest bin, img_utils.py:183-190) -- at
    // d_min = 3 m that is whole image columns (tests/test_gpu_fullsize.py C5, profiles/r02_fullsize_bins_diag.txt).
    // The update is a product of likelihoods, so the skipped factor can be applied now, evaluated the way the
    // reference sees it: an fp32 denormal is a whole number of units of 2^-149, so the reference's exp() is
    // rint(2^(e + 149)) units (e = log2 of the Gaussian; exact above 2^23 units, i.e. for normal results) --
    // including the few-bit quantisation that shows in g / sum(g) when the nearest bin is only a few hundred
    // units, and including a sum of exactly 0 (uniform likelihood, nothing to apply) when even the nearest bin
    // is below half a unit: surfaces many sigma outside the bins (image edges, z < d_min), the common case here.
    // ex2.approx on e + 149 is a normal number again; the scale 2^149 drops out of g / sum(g).
    // Above 2^-86 (the loop's threshold) a flushed bin is < 2^-40 of the sum: nothing the 1/D term can see.
    if (__any_sync(gmask, slow_mask != 0u)) {
#pragma unroll 1
      for (int j = 0; j < jn; j++) {
        const bool slow = (slow_mask >> j) & 1u;
        if (!__any_sync(gmask, slow)) continue;
        float Lz = mine.z, Lkk = mine_kk, LA = mine.A, LBu = mine.Bu;
        if (T > 1) {
          Lz = __shfl_sync(gmask, mine.z, gbase + j);
          Lkk = __shfl_sync(gmask, mine_kk, gbase + j);
          LA = __shfl_sync(gmask, mine.A, gbase + j);
          LBu = __shfl_sync(gmask, mine.Bu, gbase + j);
        }
        const float2 kk2 = splat2(Lkk), nz2 = splat2(-Lz), up2 = splat2(149.f);
        Col<NP> g;
        float2 a0 = splat2(0.f), a1 = splat2(0.f);
#pragma unroll
        for (int k = 0; k < NP; k += 2) {
          const float4 q = dc4_at(k >> 1);
          float2 e0 = __fadd2_rn(make_float2(q.x, q.y), nz2), e1 = __fadd2_rn(make_float2(q.z, q.w), nz2);
          e0 = __ffma2_rn(__fmul2_rn(e0, e0), kk2, up2);
          e1 = __ffma2_rn(__fmul2_rn(e1, e1), kk2, up2);
          g.v[k] = make_float2(slow && valid(2 * k) ? rintf(exp2f_approx(e0.x)) : 0.f,
                               slow && valid(2 * k + 1) ? rintf(exp2f_approx(e0.y)) : 0.f);
          g.v[k + 1] = make_float2(slow && valid(2 * k + 2) ? rintf(exp2f_approx(e1.x)) : 0.f,
                                   slow && valid(2 * k + 3) ? rintf(exp2f_approx(e1.y)) : 0.f);
          a0 = __fadd2_rn(a0, g.v[k]);
          a1 = __fadd2_rn(a1, g.v[k + 1]);
        }
        const float sgl = (a0.x + a0.y) + (a1.x + a1.y);
        const float sg = group_sum<T>(sgl, gmask);
        const float As = LA / sg;
        float sml = 0.f;
#pragma unroll
        for (int k = 0; k < NP; k++) {                         // mix_value (lc_sense.cuh) per bin, clamped at 0
          float2 m = __ffma2_rn(g.v[k], splat2(As), splat2(LBu));
          m.x = valid(2 * k) ? fmaxf(m.x, 0.f) : 0.f;
          m.y = valid(2 * k + 1) ? fmaxf(m.y, 0.f) : 0.f;
          g.v[k] = m;
          sml += m.x + m.y;
        }
        const float sm = group_sum<T>(sml, gmask);
        if (slow && sg > 0.f && sm > 0.f) {
          const float esm = eps_floor() * sm;
#pragma unroll
          for (int k = 0; k < NP; k++)
            p.v[k] = __fmul2_rn(p.v[k], make_float2(fmaxf(g.v[k].x, esm), fmaxf(g.v[k].y, esm)));
        }
        // one more multiplication than the loop's range check counted on: look at the total right away
        a0 = splat2(0.f); a1 = splat2(0.f);
#pragma unroll
        for (int k = 0; k < NP; k += 2) {
          a0 = __fadd2_rn(a0, p.v[k]);
          a1 = __fadd2_rn(a1, p.v[k + 1]);
        }
        const float s = group_sum<T>((a0.x + a0.y) + (a1.x + a1.y), gmask);
        if (s < 1e-12f || s > 1e12f) {
          const float2 inv2 = splat2(1.f / s);
#pragma unroll
          for (int k = 0; k < NP; k++) p.v[k] = __fmul2_rn(p.v[k], inv2);
        }
      }
    }
#endif
    c0 += T;
  } while (c0 < nc);

  // ---- box spread along depth x spread_iter (img_utils.py:408-436), one renormalisation at the end
#pragma unroll 1
  for (int i = 0; i < spread_iter; i++) spread_pass<DPT, T, HALF, FULL>(p, d0, D, t, lane, gmask);
  float2 a0 = splat2(0.f), a1 = splat2(0.f);
#pragma unroll
  for (int k = 0; k < NP; k += 2) {
    a0 = __fadd2_rn(a0, p.v[k]);
    a1 = __fadd2_rn(a1, p.v[k + 1]);
  }
  const float2 inv2 = splat2(1.f / group_sum<T>((a0.x + a0.y) + (a1.x + a1.y), gmask));

  // ---- clamp, (log), expectation, store (models.py:1119-1122) -------------------------------
  // LC_FUSE_OUT_RAW: more curtains follow in another call -- the reference only clamps after the last one
  const float floor_out = (flags & 8) ? 0.f : eps_floor();
  float2 m0 = splat2(0.f), m1 = splat2(0.f);
#pragma unroll
  for (int k = 0; k < NP; k += 2) {
    const float4 q = dc4_at(k >> 1);
    float2 c0v = __fmul2_rn(p.v[k], inv2), c1v = __fmul2_rn(p.v[k + 1], inv2);
    // clamp(., eps, 1): the upper bound cannot bind -- a sum of positive floats is >= each term, and
    // x * fl(1/x) <= 1 + 2^-24 rounds to 1 (ties to even)
    c0v = make_float2(fmaxf(c0v.x, floor_out), fmaxf(c0v.y, floor_out));
    c1v = make_float2(fmaxf(c1v.x, floor_out), fmaxf(c1v.y, floor_out));
    p.v[k] = c0v;
    p.v[k + 1] = c1v;
    m0 = __ffma2_rn(make_float2(q.x, q.y), c0v, m0);         // d_candi is 0 in the padding bins
    m1 = __ffma2_rn(make_float2(q.z, q.w), c1v, m1);
  }
  {
    float* dst = vol_out + base;
    if (flags & 2) {
#pragma unroll
      for (int d = 0; d < DPT; d++, dst = bump(dst, HW))
        if (valid(d) && live) st_stream(dst, p.get(d));
    } else {
#pragma unroll
      for (int d = 0; d < DPT; d++, dst = bump(dst, HW))
        if (valid(d) && live) st_stream(dst, log2f_approx(p.get(d)) * 0.6931471805599453f);
    }
  }
  const float mean = group_sum<T>((m0.x + m0.y) + (m1.x + m1.y), gmask);
  if (mean_out && t == 0 && live) mean_out[pix] = mean;
  }  // rows of this CTA
}

}  // namespace dpv

using namespace dpv;

extern "C" int lc_sense_fuse(const LcDevice* dev, const float* logp_in, const float* depth, const float* planes,
                             const float* noise, int B, int nc, const float* d_candi, int D, int spread_iter,
                             int spread_n, int flags, float* logp_out, float* mean_out, void* stream) {
  DPV_REQUIRE(dev && logp_in && depth && planes && d_candi && logp_out, "lc_sense_fuse: null pointer");
  DPV_REQUIRE(B > 0 && nc >= 0 && nc <= kMaxCurtains && D > 0 && spread_iter >= 0,
              "lc_sense_fuse: bad shape (B=%d nc=%d D=%d)", B, nc, D);
  DPV_REQUIRE(spread_iter == 0 || spread_n == 1 || spread_n == 3 || spread_n == 5 || spread_n == 7,
              "lc_sense_fuse: spread_n=%d not in {1,3,5,7}", spread_n);
  DPV_REQUIRE(D <= 128, "lc_sense_fuse: D=%d > 128 unsupported", D);
  DPV_REQUIRE(((uintptr_t)planes & 15) == 0, "lc_sense_fuse: planes must be 16-byte aligned");
  DPV_REQUIRE((flags & ~15) == 0, "lc_sense_fuse: unknown flags %d", flags);
  DPV_REQUIRE(!(flags & 4) || (flags & 1), "lc_sense_fuse: LC_FUSE_IN_CLAMPED needs LC_FUSE_IN_PROB");
  DPV_REQUIRE(!(flags & 8) || (flags & 2), "lc_sense_fuse: LC_FUSE_OUT_RAW needs LC_FUSE_OUT_PROB");
  const long long hw_img = (long long)dev->H * dev->W;
  DPV_REQUIRE(B <= 65535 && dev->H <= 65535 && hw_img * 32 < (1ll << 31), "lc_sense_fuse: B=%d or image too large", B);
  cudaStream_t st = (cudaStream_t)stream;
  const int half = spread_iter == 0 ? 0 : spread_n / 2;
  // A CTA keeps its 32 columns and walks `rows` image rows: the prologue (depth bins and mapping table into shared
  // memory) and the column-only sheet geometry are paid once per CTA.  At most 4 rows (C2: 0.301 / 0.281 / 0.272 /
  // 0.279 ms for 1 / 2 / 4 / 8 rows), fewer while the grid would have under ten waves of CTAs on 148 SMs x 8 resident
  // CTAs: with five waves the last, partly filled one already costs what the longer loop saves (tuning aid:
  // DPV_FUSE_ROWS).
  static const int rows_env = getenv("DPV_FUSE_ROWS") ? atoi(getenv("DPV_FUSE_ROWS")) : 0;
  int rows = rows_env > 0 ? rows_env : 4;
  if (rows_env <= 0) {
    const long long ctas_per_row = (long long)((dev->W * 4 + kFuseThreads - 1) / kFuseThreads) * B;
    while (rows > 1 && ctas_per_row * ((dev->H + rows - 1) / rows) < 10ll * 148 * 8) rows >>= 1;
  }
#define LAUNCH4(DPT, T, HALF, FULL)                                                                                \
  sense_fuse_kernel<DPT, T, HALF, FULL>                                                                            \
      <<<dim3((unsigned)((dev->W * T + kFuseThreads - 1) / kFuseThreads), (unsigned)((dev->H + rows - 1) / rows), \
              (unsigned)B), kFuseThreads, 0, st>>>(                                                                \
          dev->view, logp_in, depth, planes, noise, nc, d_candi, D, spread_iter, flags, logp_out, mean_out, rows)
#define LAUNCH3(DPT, T, HALF)                        \
  do {                                               \
    if (D == DPT * T) { LAUNCH4(DPT, T, HALF, true); }   \
    else { LAUNCH4(DPT, T, HALF, false); }               \
  } while (0)
#define LAUNCH(DPT, T)                       \
  do {                                       \
    if (half == 0) LAUNCH3(DPT, T, 0);       \
    else if (half == 1) LAUNCH3(DPT, T, 1);  \
    else if (half == 2) LAUNCH3(DPT, T, 2);  \
    else LAUNCH3(DPT, T, 3);                 \
  } while (0)
  // each pixel's depth column is split over T adjacent lanes, at most 32 bins per lane (register budget);
  // D == DPT*T (16, 32, 64, 128) compiles without any per-bin bound check
  if (D <= 16) LAUNCH(16, 1);
  else if (D <= 32) LAUNCH(32, 1);
  else if (D <= 64) {
    static const int lanes = getenv("DPV_FUSE_LANES") ? atoi(getenv("DPV_FUSE_LANES")) : 4;
    if (lanes == 2) LAUNCH(32, 2);                             // 4 lanes: 64 registers, twice the resident warps
    else LAUNCH(16, 4);
  }
  else LAUNCH(32, 4);
#undef LAUNCH
#undef LAUNCH3
#undef LAUNCH4
  return check_launch("lc_sense_fuse");
}
