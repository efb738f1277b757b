// Per-pixel light-curtain image formation and likelihood parameters shared by lc_sense_nofma.cu and
// lc_fuse_nofma.cu (both compiled with -fmad=false: fp32 expressions round exactly as written).
//   external/lcsim/lcsim/src/dprocessor.cpp:806-887, :942-993 ; util.cpp:47-52
//   lc/light_curtain.py:847-888, :228-252 ; utils/img_utils.py:176-195
#pragma once
#include "dpv_pixel.cuh"
#include "lc_device.cuh"

namespace dpv {

constexpr int kSenseThreads = 128;

struct Ret {          // one pixel of LCDevice.get_return
  float x, y, z, inten;   // NaN x => no return
  float thick;
};

// util.cpp:47-52
__device__ __forceinline__ void intersect(const float* P, float rx, float ry, float rz, float out[3]) {
  float t = -P[3] / (P[0] * rx + P[1] * ry + P[2]);
  out[0] = rx * t;
  out[1] = ry * t;
  out[2] = rz * t;
}

__device__ __forceinline__ Ret sense_pixel(const LcDeviceView& dv, const float* __restrict__ P, int u, int v,
                                           float zval, const float* noise) {
  Ret r;
  // A column without a design point carries a NaN centre plane and all-zero thickness planes
  // (curtain_planes_kernel): every quotient below is then NaN and the pixel has no return.  Same result without
  // sending three divisions and a double-precision square root down their NaN slow paths.
  {
    const bool nan0 = (P[0] != P[0]) || (P[1] != P[1]) || (P[2] != P[2]) || (P[3] != P[3]);
    bool zero12 = true;
#pragma unroll
    for (int k = 4; k < 12; k++) zero12 = zero12 && (P[k] == 0.f);
    if (nan0 && zero12) {
      r.x = r.y = r.z = NAN;
      r.inten = 0.f;
      r.thick = NAN;
      return r;
    }
  }
  const float rx = dv.nn_x[u], ry = dv.nn_y[v];
  float c[3], bot[3], top[3];
  intersect(P, rx, ry, 1.0f, c);
  float c3 = 0.f;
  if (c[2] < 0 || fabs((double)c[1]) > 3.0) { c[0] = c[1] = c[2] = c3 = NAN; }      // :818-823
  intersect(P + 4, rx, ry, 1.0f, bot);
  intersect(P + 8, rx, ry, 1.0f, top);
  double dx = (double)(bot[0] - top[0]), dy = (double)(bot[1] - top[1]), dz = (double)(bot[2] - top[2]);
  const float unc = (float)sqrt(dx * dx + dy * dy + dz * dz);                        // :884-887
  r.thick = unc;
  if (!(c[0] != c[0]) && !(unc != unc) && !(zval != zval)) {                         // computeDepthHits
    float srange = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    float drange = zval / dv.ztoramap[(size_t)v * dv.W + u];
    float inten = 0.f;
    if (dv.hit_mode == 0) {
      float err = fabsf(drange - srange);
      float color = 0;
      if (err >= 0) color = (float)(1. - 1. * (err / (unc / 2.)));
      if (err < unc / 2.) inten = color;
    } else {
      float err = drange - srange;
      // (float)(err / (unc / 2.)) of dprocessor.cpp:975: unc / 2 is exact, and a double-precision quotient of two
      // floats rounded to float IS the correctly rounded float quotient (53 >= 2 * 24 + 2 bits: double rounding is
      // innocuous for division), so the fp32 IEEE division gives the same bits without the fp64 pipe
      const float half_unc = unc * 0.5f;
      float scale = __fdiv_rn(err, half_unc);
      if (fabsf(err) > half_unc) scale = (float)(((0.f < err) - (err < 0.f)) * 1.);
      int idx = (int)(((scale + 1.) / 2.) * (dv.hit_N - 1));
      idx = min(max(idx, 0), dv.hit_N - 1);
      inten = dv.hit_lut[idx];
    }
    if (noise) {
      inten += *noise;
      inten = fmaxf(0.f, fminf(inten, 1.f));
    }
    c3 = (float)(int)(255 * inten);
  }
  r.x = c[0]; r.y = c[1]; r.z = c[2]; r.inten = c3;
  return r;
}

// ---- the same image formation with its column-only part factored out --------------------------------------
// When none of a column's three sheet planes has a y component (P[1] == P[5] == P[9] == 0: every laser that only
// rotates about the camera's y axis, i.e. every rig lcsim simulates on this path) the ray parameters
// t = -P3 / (P0 rx + P1 ry + P2) of util.cpp:47-52 do not depend on the image row: P1 * ry is +-0 and adding it
// changes nothing, so the quotients below are bit for bit the ones sense_pixel forms at every pixel of the column.
struct ColGeom {
  float tc, tb, tt;     // ray parameter on the centre / lower / upper sheet plane
  float c0;             // surface x = rx * tc
  float dxf, dzf;       // (bot - top).x and .z
  int kind;             // 0: column-only form valid, 1: no design point (no return anywhere), 2: general planes
};

__device__ __forceinline__ ColGeom column_geometry(const float* P, float rx) {
  ColGeom g;
  g.tc = g.tb = g.tt = g.c0 = g.dxf = g.dzf = 0.f;
  const bool nan0 = (P[0] != P[0]) || (P[1] != P[1]) || (P[2] != P[2]) || (P[3] != P[3]);
  bool zero12 = true;
#pragma unroll
  for (int k = 4; k < 12; k++) zero12 = zero12 && (P[k] == 0.f);
  if (nan0 && zero12) { g.kind = 1; return g; }
  if (!(P[1] == 0.f && P[5] == 0.f && P[9] == 0.f)) { g.kind = 2; return g; }
  g.kind = 0;
  g.tc = -P[3] / (P[0] * rx + P[2]);
  g.tb = -P[7] / (P[4] * rx + P[6]);
  g.tt = -P[11] / (P[8] * rx + P[10]);
  g.c0 = rx * g.tc;
  g.dxf = rx * g.tb - rx * g.tt;
  g.dzf = g.tb - g.tt;
  return g;
}

// sense_pixel for a column whose ColGeom.kind == 0 (hit_mode 1 and 0 alike)
__device__ __forceinline__ Ret sense_pixel_col(const LcDeviceView& dv, const ColGeom& g, int u, int v, float zval,
                                               const float* noise) {
  Ret r;
  const float ry = dv.nn_y[v];
  float c[3] = {g.c0, ry * g.tc, g.tc};
  float c3 = 0.f;
  if (c[2] < 0 || fabs((double)c[1]) > 3.0) { c[0] = c[1] = c[2] = c3 = NAN; }      // :818-823
  const float dyf = ry * g.tb - ry * g.tt;
  const double dx = (double)g.dxf, dy = (double)dyf, dz = (double)g.dzf;
  const float unc = (float)sqrt(dx * dx + dy * dy + dz * dz);                        // :884-887
  r.thick = unc;
  if (!(c[0] != c[0]) && !(unc != unc) && !(zval != zval)) {                         // computeDepthHits
    float srange = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    float drange = zval / dv.ztoramap[(size_t)v * dv.W + u];
    float inten = 0.f;
    if (dv.hit_mode == 0) {
      float err = fabsf(drange - srange);
      float color = 0;
      if (err >= 0) color = (float)(1. - 1. * (err / (unc / 2.)));
      if (err < unc / 2.) inten = color;
    } else {
      float err = drange - srange;
      const float half_unc = unc * 0.5f;
      float scale = __fdiv_rn(err, half_unc);
      if (fabsf(err) > half_unc) scale = (float)(((0.f < err) - (err < 0.f)) * 1.);
      int idx = (int)(((scale + 1.) / 2.) * (dv.hit_N - 1));
      idx = min(max(idx, 0), dv.hit_N - 1);
      inten = dv.hit_lut[idx];
    }
    if (noise) {
      inten += *noise;
      inten = fmaxf(0.f, fminf(inten, 1.f));
    }
    c3 = (float)(int)(255 * inten);
  }
  r.x = c[0]; r.y = c[1]; r.z = c[2]; r.inten = c3;
  return r;
}

// Per-pixel, per-curtain parameters of the likelihood (sense_high tail + mixed_model)
struct Lik {
  float z;      // sensed z * mask
  float den;    // 2 * sigma^2
  float A, Bu;  // mixture weight, (1-A) * (1/D)
};

// a / b for a finite b > 0, rounded like the IEEE division, with a zero numerator answered directly (+-0 / b = +-0):
// 0 / b goes through the division's special-case path (FCHK), and un-sensed pixels (most of the image) have exactly that
__device__ __forceinline__ float div_zero_fast(float a, float b) {
  const bool z = a == 0.f;
  const float q = opaque(z ? b : a) / b;      // opaque: the compiler otherwise divides a itself and selects after
  return z ? a : q;
}

// it / 255 for the integer intensities 0..255 the simulator emits (dprocessor.cpp:990): reciprocal, one Newton
// step, one residual correction -- the division's own fast path, correctly rounded for all 256 values (checked
// exhaustively on the host), without its special-case path for it == 0
__device__ __forceinline__ float div255(float it) {
  if (!(it >= 0.f && it <= 255.f)) return it / 255.0f;
  const float rc0 = 1.0f / 255.0f;
  const float rc = fmaf(fmaf(-255.f, rc0, 1.f), rc0, rc0);
  const float q = it * rc;
  return fmaf(fmaf(-255.f, q, it), rc, q);
}

// mapping(intensity / 255), light_curtain.py:228-244 (m = 20, f = 0.5) with :875's division
__device__ __forceinline__ float mapping_of(float it) {
  const float x = div255(it);
  if (x > 0.5f) {
    const float cc = (float)(20.0 / ((20.0 * 0.5 + 0.5) * (20.0 * 0.5 + 0.5)));
    const float c1 = (float)(1.0 - 20.0 / ((20.0 * 0.5 + 0.5) * (20.0 * 0.5 + 0.5)));
    return cc * x + c1;
  }
  return (1.0f / (20.f * (0.025f + x))) * -1.0f + 1.0f;
}

// a_lut: optional table of mapping_of(0..255) (the simulator's intensities are (int)(255 * i), dprocessor.cpp:990),
// filled with mapping_of itself, so reading it is the same value without the two divisions per pixel and curtain
__device__ __forceinline__ Lik lik_params(const Ret& r, float gt, int D, const float* a_lut = nullptr) {
  float z = r.z, th = r.thick, it = r.inten;
  if (r.x != r.x) { z = 0.f; it = 0.f; }            // light_curtain.py:847 (whole pixel -> 0)
  if (th != th) th = 0.f;                           // :848 (only where the thickness itself is NaN)
  const float mask = gt > 0.f ? 1.f : 0.f;          // :870
  z = z * mask;
  th = th * mask;
  it = it * mask;
  float A;
  const int ii = (int)it;
  if (a_lut && it >= 0.f && it <= 255.f && (float)ii == it) A = a_lut[ii];
  else A = mapping_of(it);
  Lik L;
  L.z = z;
  const float t10 = div_zero_fast(th, 10.0f);
  const float var = t10 * t10;
  const float sg = var == 0.f ? var : sqrtf(opaque(var == 0.f ? 1.f : var));     // sqrt(+-0) = +-0, off the slow path
  L.den = 2.f * (sg * sg);
  L.A = A;
  L.Bu = (1.0f / (float)D) * (1.f - A);
  return L;
}

// exp(-|dc - z|^2 / (2 sigma^2)); exact 0 once the argument is below fp32 exp underflow
__device__ __forceinline__ float soft_g(float dc, const Lik& L) {
  float a = fabsf(dc - L.z);
  a = a * a;
  if (a > 105.f * L.den) return 0.f;      // (-a)/den < -104.9: expf rounds to +0
  return expf((-a) / L.den);
}

__device__ __forceinline__ float clamp0_keepnan(float v) { return (v != v) ? v : fmaxf(v, 0.f); }

// mix[d] before the final normalisation: clamp(g/sg * A + (1/D)*(1-A), 0, inf) with NaN -> -1 first
__device__ __forceinline__ float mix_value(float g, float sg, const Lik& L) {
  float n = g / sg;
  if (n != n) n = -1.f;                   // zero_invalid (img_utils.py:189-190)
  return clamp0_keepnan(n * L.A + L.Bu);
}

template <int DPT>
__device__ __forceinline__ void lik_sums(const float* dc, int D, const Lik& L, float& sg, float& sm) {
  sg = 0.f;
#pragma unroll
  for (int d = 0; d < DPT; d++)
    if (d < D) sg += soft_g(dc[d], L);
  sm = 0.f;
  if (sg > 0.f) {
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) sm += mix_value(soft_g(dc[d], L), sg, L);
  } else {
    // every bin is 0/0 -> -1 (or NaN sum): the mixture is the same constant in every bin
    float m = mix_value(0.f, sg, L);
#pragma unroll
    for (int d = 0; d < DPT; d++)
      if (d < D) sm += m;
  }
}

}  // namespace dpv
