// lcsim device model: host-side sensor set-up and the per-curtain geometry kernel
// (checkPoints -> findCameraIntersectionsOpt -> per-column laser planes).
//   external/lcsim/lcsim/src/dprocessor.cpp:41-235, 238-280, 514-634, 735-802, 841-863, 931-936
//   external/lcsim/lcsim/src/util.cpp:5-52 ; algo.cpp:377-385, 676-682, 696-741
// Compiled with -fmad=false: fp32 expressions round exactly as written (left to right), so the two
// intensity quantisers downstream see the same bits on the GPU as on a CPU run of the same algebra.
#include <cfloat>
#include <cmath>
#include <cstring>

#include "lc_device.cuh"

namespace dpv {

__host__ __device__ inline float sin_cr(float x) { return (float)sin((double)x); }
__host__ __device__ inline float cos_cr(float x) { return (float)cos((double)x); }
__host__ __device__ inline float atan2_cr(float y, float x) { return (float)atan2((double)y, (double)x); }

__host__ __device__ inline LcMat4 mat4_identity() {
  LcMat4 r;
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) r.m[i][j] = (i == j) ? 1.f : 0.f;
  return r;
}

__host__ __device__ inline LcMat4 mat4_mul(const LcMat4& a, const LcMat4& b) {
  LcMat4 r;
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      float s = a.m[i][0] * b.m[0][j];
      for (int k = 1; k < 4; k++) s = s + a.m[i][k] * b.m[k][j];
      r.m[i][j] = s;
    }
  return r;
}

// util.cpp:5-45 getTransformMatrix(yaw, pitch, roll, x, y, z); the reference feeds (roll,pitch,yaw)
// into parameters named (eulerZ, eulerY, eulerX).
__host__ __device__ inline LcMat4 transform_matrix(float yaw, float pitch, float roll, float x, float y, float z) {
  float eulerZ = (float)(roll * M_PI / 180.);
  float eulerY = (float)(pitch * M_PI / 180.);
  float eulerX = (float)(yaw * M_PI / 180.);
  float ci = cos_cr(eulerX), cj = cos_cr(eulerY), ch = cos_cr(eulerZ);
  float si = sin_cr(eulerX), sj = sin_cr(eulerY), sh = sin_cr(eulerZ);
  float cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
  LcMat4 r = mat4_identity();
  r.m[0][0] = cj * ch;  r.m[0][1] = sj * sc - cs;  r.m[0][2] = sj * cc + ss;
  r.m[1][0] = cj * sh;  r.m[1][1] = sj * ss + cc;  r.m[1][2] = sj * cs - sc;
  r.m[2][0] = -sj;      r.m[2][1] = cj * si;       r.m[2][2] = cj * ci;
  r.m[0][3] = x; r.m[1][3] = y; r.m[2][3] = z;
  return r;
}

// transform_matrix(0, pitch, 0, 0, 0, 0) with the zero angles' cos = 1, sin = 0 written in (same expressions, so the
// same values and zero signs; two trigonometric calls instead of six)
__host__ __device__ inline LcMat4 transform_matrix_pitch(float pitch) {
  float eulerY = (float)(pitch * M_PI / 180.);
  const float ci = 1.f, ch = 1.f, si = 0.f, sh = 0.f;
  float cj = cos_cr(eulerY), sj = sin_cr(eulerY);
  float cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
  LcMat4 r = mat4_identity();
  r.m[0][0] = cj * ch;  r.m[0][1] = sj * sc - cs;  r.m[0][2] = sj * cc + ss;
  r.m[1][0] = cj * sh;  r.m[1][1] = sj * ss + cc;  r.m[1][2] = sj * cs - sc;
  r.m[2][0] = -sj;      r.m[2][1] = cj * si;       r.m[2][2] = cj * ci;
  r.m[0][3] = 0.f; r.m[1][3] = 0.f; r.m[2][3] = 0.f;
  return r;
}

// (M^-1)^T (1,0,0,0) for a rigid M: the laser-frame plane (dprocessor.cpp:781-797, 845-857)
__host__ __device__ inline void rigid_inverse_row0(const LcMat4& M, float out[4]) {
  out[0] = M.m[0][0];
  out[1] = M.m[1][0];
  out[2] = M.m[2][0];
  float s = M.m[0][0] * M.m[0][3];
  s = s + M.m[1][0] * M.m[1][3];
  s = s + M.m[2][0] * M.m[2][3];
  out[3] = -s;
}

// planes_cframe = Q^T plane, Q = laser_to_cam.inverse()  (dprocessor.cpp:802)
__host__ __device__ inline void plane_to_cam(const LcMat4& Q, const float pl[4], float out[4]) {
  for (int r = 0; r < 4; r++) {
    float s = Q.m[0][r] * pl[0];
    for (int k = 1; k < 4; k++) s = s + Q.m[k][r] * pl[k];
    out[r] = s;
  }
}

// Eigen MatrixXf::inverse() stand-in: Gauss-Jordan with partial pivoting (exact for pure translations)
static LcMat4 mat4_inverse_general(const LcMat4& a) {
  float w[4][8];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      w[i][j] = a.m[i][j];
      w[i][j + 4] = (i == j) ? 1.f : 0.f;
    }
  for (int c = 0; c < 4; c++) {
    int p = c;
    for (int r = c + 1; r < 4; r++)
      if (std::fabs(w[r][c]) > std::fabs(w[p][c])) p = r;
    if (p != c)
      for (int j = 0; j < 8; j++) { float t = w[p][j]; w[p][j] = w[c][j]; w[c][j] = t; }
    float d = w[c][c];
    for (int j = 0; j < 8; j++) w[c][j] = w[c][j] / d;
    for (int r = 0; r < 4; r++) {
      if (r == c) continue;
      float f = w[r][c];
      if (f == 0.f) continue;
      for (int j = 0; j < 8; j++) w[r][j] = w[r][j] - f * w[c][j];
    }
  }
  LcMat4 r;
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) r.m[i][j] = w[i][j + 4];
  return r;
}

// createNormalMap (dprocessor.cpp:41-137) + hit LUT (:931-936)
static void build_normal_map(LcDevice& d) {
  const int W = d.W, H = d.H;
  double fx = d.K[0], fy = d.K[4], cx = d.K[2], cy = d.K[5];
  double ifx = 1. / fx, ify = 1. / fy;                 // cv::undistortPoints, zero distortion
  d.nn_x.resize(W);
  d.nn_y.resize(H);
  for (int i = 0; i < W; i++) d.nn_x[i] = (float)(((double)(float)(i + 1) - cx) * ifx);   // 1-based (:17-27)
  for (int j = 0; j < H; j++) d.nn_y[j] = (float)(((double)(float)(j + 1) - cy) * ify);
  d.ztoramap.resize((size_t)H * W);
  d.mid_x.resize(W);
  d.mid_z.resize(W);
  const int mid = H / 2 - 1;                           // :82
  for (int i = 0; i < W; i++)
    for (int j = 0; j < H; j++) {
      float x = d.nn_x[i], y = d.nn_y[j], z = 1.0f;
      float mag = std::sqrt(x * x + y * y + z * z);
      float vx = x / mag, vy = y / mag, vz = z / mag;
      float theta = atan2f(vx, vz);
      float phi = asinf(vy / 1);
      d.ztoramap[(size_t)j * W + i] = cosf(theta) * cosf(phi);
      if (j == mid) { d.mid_x[i] = vx; d.mid_z[i] = vz; }
    }
  d.angles.resize(W);
  for (int i = 0; i < W; i++)
    d.angles[i] = (float)(-((atan2_cr(d.mid_z[i], d.mid_x[i]) * 180 / M_PI) - 90));   // :85-91
  float z_max = 1;                                      // camera wedge :93-107
  float t = (float)(1.25 * z_max / d.mid_z[0]);
  d.view.cam_origin[0] = 0.f; d.view.cam_origin[1] = 0.f;
  d.view.p_left_cam[0] = 0.f + t * d.mid_x[0];      d.view.p_left_cam[1] = 0.f + t * d.mid_z[0];
  d.view.p_right_cam[0] = 0.f + t * d.mid_x[W - 1]; d.view.p_right_cam[1] = 0.f + t * d.mid_z[W - 1];
  d.t_max = t;
  const int N = d.hit_N;                                // algo.cpp:377-385 linspace, :676-682 gaussian
  d.hit_lut.resize(N);
  float h = (1.f - (-1.f)) / static_cast<float>(N - 1);
  float val = -1.f;
  for (int k = 0; k < N; k++, val += h) {
    float a = std::pow(std::fabs(val - 0.f), d.hit_pow);
    float den = (float)(2 * std::pow(d.hit_std, d.hit_pow));
    d.hit_lut[k] = std::exp(-a / den);
  }
  float mx = d.hit_lut[0], mn = d.hit_lut[0];
  for (float v : d.hit_lut) { mx = std::max(mx, v); mn = std::min(mn, v); }
  for (float& v : d.hit_lut) v = (v - mn) / (mx - mn);
}

// computeLaserParams (dprocessor.cpp:156-211)
static void build_laser(LcDevice& d) {
  d.laser_to_cam = mat4_inverse_general(d.cam_to_laser);
  d.laser_to_cam_inv = mat4_inverse_general(d.laser_to_cam);
  d.view.laser_origin[0] = d.laser_to_cam.m[0][3];
  d.view.laser_origin[1] = d.laser_to_cam.m[2][3];
  float th[2] = {-d.fov / 2.f, d.fov / 2.f};
  for (int s = 0; s < 2; s++) {
    float sx = (float)std::sin(M_PI / 180 * th[s]), cz = (float)std::cos(M_PI / 180 * th[s]);
    float a[3] = {0, 1, 0}, b[3] = {sx, 0, cz};
    float n[4] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0], 0.f};
    float c[4];
    for (int j = 0; j < 4; j++) {
      float acc = n[0] * d.laser_to_cam_inv.m[0][j];
      for (int k = 1; k < 4; k++) acc = acc + n[k] * d.laser_to_cam_inv.m[k][j];
      c[j] = acc;
    }
    float dir[3] = {0, -1, 0};
    float v0 = dir[1] * c[2] - dir[2] * c[1];
    float v2 = dir[0] * c[1] - dir[1] * c[0];
    float* dst = s == 0 ? d.view.p_left_laser : d.view.p_right_laser;
    dst[0] = d.view.laser_origin[0] + d.t_max * v0;
    dst[1] = d.view.laser_origin[1] + d.t_max * v2;
  }
}

}  // namespace dpv

using namespace dpv;

extern "C" {

int lc_camera_angles(const float* K, int W, int H, float* out) {
  DPV_REQUIRE(K && out && W > 0 && H > 1, "lc_camera_angles: bad argument");
  double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  double ifx = 1. / fx, ify = 1. / fy;
  int j = H / 2 - 1;
  float y = (float)(((double)(float)(j + 1) - cy) * ify);
  for (int i = 0; i < W; i++) {
    float x = (float)(((double)(float)(i + 1) - cx) * ifx);
    float z = 1.0f;
    float mag = std::sqrt(x * x + y * y + z * z);
    float vx = x / mag, vz = z / mag;
    out[i] = (float)(-((atan2f(vz, vx) * 180 / M_PI) - 90));     // algo.cpp:733-737
  }
  return DPV_OK;
}

int lc_device_create(int W, int H, const float* K, const float* lTc, float fov, float laser_timestep,
                     float laser_limit, float thickness, float divergence, int hit_mode, float limit,
                     LcDevice** out) {
  DPV_REQUIRE(K && lTc && out, "lc_device_create: null pointer");
  DPV_REQUIRE(W >= 2 && H >= 2, "lc_device_create: bad image size %dx%d", W, H);
  DPV_REQUIRE(hit_mode == 0 || hit_mode == 1, "lc_device_create: hit_mode %d", hit_mode);
  LcDevice* d = new LcDevice();
  d->W = W; d->H = H;
  std::memcpy(d->K, K, sizeof(float) * 9);
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) d->cam_to_laser.m[i][j] = lTc[i * 4 + j];
  d->fov = fov; d->laser_timestep = laser_timestep; d->laser_limit = laser_limit;
  d->thickness = thickness; d->divergence = divergence; d->hit_mode = hit_mode; d->limit = limit;
  build_normal_map(*d);
  build_laser(*d);
  // one device allocation: nn_x | nn_y | ztoramap | mid_x | mid_z | angles | lut
  size_t n = (size_t)W + H + (size_t)H * W + 3 * (size_t)W + d->hit_N;
  std::vector<float> host;
  host.reserve(n);
  auto push = [&](const std::vector<float>& v) { size_t o = host.size(); host.insert(host.end(), v.begin(), v.end()); return o; };
  size_t o_nx = push(d->nn_x), o_ny = push(d->nn_y), o_zt = push(d->ztoramap), o_mx = push(d->mid_x),
         o_mz = push(d->mid_z), o_an = push(d->angles), o_lut = push(d->hit_lut);
  cudaError_t e;
  if ((e = cudaMalloc(&d->d_tables, host.size() * sizeof(float))) != cudaSuccess ||
      (e = cudaMemcpy(d->d_tables, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess) {
    dpv::set_error("lc_device_create: %s", cudaGetErrorString(e));
    lc_device_destroy(d);
    return DPV_ECUDA;
  }
  LcDeviceView& v = d->view;
  v.W = W; v.H = H;
  v.nn_x = d->d_tables + o_nx; v.nn_y = d->d_tables + o_ny; v.ztoramap = d->d_tables + o_zt;
  v.mid_x = d->d_tables + o_mx; v.mid_z = d->d_tables + o_mz; v.angles = d->d_tables + o_an;
  v.hit_lut = d->d_tables + o_lut;
  v.hit_N = d->hit_N; v.hit_mode = hit_mode; v.limit = limit; v.thickness = thickness; v.divergence = divergence;
  d->plane_consts.off_l = transform_matrix(0, 0, 0, (thickness / 2), 0, 0);        // dprocessor.cpp:845-857
  d->plane_consts.div_l = transform_matrix(0, -divergence, 0, 0, 0, 0);
  d->plane_consts.off_r = transform_matrix(0, 0, 0, -(thickness / 2), 0, 0);
  d->plane_consts.div_r = transform_matrix(0, divergence, 0, 0, 0, 0);
  v.cam_to_laser = d->cam_to_laser;
  v.laser_to_cam_inv = d->laser_to_cam_inv;
  *out = d;
  return DPV_OK;
}

void lc_device_destroy(LcDevice* d) {
  if (!d) return;
  cudaFree(d->d_tables);
  delete d;
}

int lc_device_angles(const LcDevice* d, float* out) {
  DPV_REQUIRE(d && out, "lc_device_angles: null pointer");
  std::memcpy(out, d->angles.data(), sizeof(float) * d->W);
  return DPV_OK;
}

int lc_device_hit_lut(const LcDevice* d, float* out) {
  DPV_REQUIRE(d && out, "lc_device_hit_lut: null pointer");
  std::memcpy(out, d->hit_lut.data(), sizeof(float) * d->hit_N);
  return DPV_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Per-curtain geometry: one CTA per curtain.
// ---------------------------------------------------------------------------------------------
namespace dpv {

constexpr int kPlaneThreads = 512;
constexpr int kPlaneCols = 64;      // columns per CTA (grid.y splits the image width): 4 per warp
constexpr int kPlaneMaxChunks = 128;  // chunk summaries of the first-hit search (curtains of up to 4097 points; more: no skipping)

__device__ __forceinline__ float point_angle(float x, float z) {
  // -((atan2f(z, x) * 180 / M_PI) - 90) + 0 + 0   (dprocessor.cpp:528-531, 762)
  return (float)(-((atan2_cr(z, x) * 180 / M_PI) - 90) + 0 + 0);
}

__global__ void __launch_bounds__(kPlaneThreads)
curtain_planes_kernel(LcDeviceView dv, LcPlaneConsts pc, const float* __restrict__ xz,
                      const int32_t* __restrict__ counts, int npts, float* __restrict__ planes,
                      float* __restrict__ design) {
  extern __shared__ float sm[];
  float* px = sm;                       // [npts]
  float* pz = px + npts;                // [npts]
  float* ang = pz + npts;               // [npts]
  int* good = reinterpret_cast<int*>(ang + npts);   // [npts]
  __shared__ int s_ngood;
  __shared__ int s_warp_tot[kPlaneThreads / 32];
  const int c = blockIdx.x;
  int N = counts ? counts[c] : npts;
  N = min(max(N, 0), npts);
  const float* p = xz + (size_t)c * npts * 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_ngood = 0;
  __syncthreads();
  // checkPoints (dprocessor.cpp:238-280), order-preserving compaction
  for (int base = 0; base < N; base += kPlaneThreads) {
    const int i = base + threadIdx.x;
    bool ok = false;
    if (i < N) {
      float x = p[2 * i], z = p[2 * i + 1];
      px[i] = x; pz[i] = z;
      ang[i] = point_angle(x, z);
      float x1 = dv.cam_origin[0], z1 = dv.cam_origin[1];
      float x2 = dv.p_left_cam[0], z2 = dv.p_left_cam[1];
      float d_cam_left = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
      x2 = dv.p_right_cam[0]; z2 = dv.p_right_cam[1];
      float d_cam_right = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
      x1 = dv.laser_origin[0]; z1 = dv.laser_origin[1];
      x2 = dv.p_left_laser[0]; z2 = dv.p_left_laser[1];
      float d_las_left = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
      x2 = dv.p_right_laser[0]; z2 = dv.p_right_laser[1];
      float d_las_right = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
      ok = d_cam_left > 0 && d_las_left > 0 && d_cam_right < 0 && d_las_right < 0;
    }
    unsigned bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) s_warp_tot[warp] = __popc(bal);
    __syncthreads();
    int off = s_ngood;
    for (int w2 = 0; w2 < warp; w2++) off += s_warp_tot[w2];
    if (ok) good[off + __popc(bal & ((1u << lane) - 1u))] = i;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w2 = 0; w2 < kPlaneThreads / 32; w2++) t += s_warp_tot[w2];
      s_ngood += t;
    }
    __syncthreads();
  }
  int ngood = s_ngood;
  const bool all_pts = (ngood == 0);            // findCameraIntersectionsOpt :519-523
  if (all_pts) ngood = N;

  // Per chunk of 32 consecutive segments (33 points of the compacted sequence): the range of the points' angles and
  // whether the angle gate (:557-561) can be decided for the chunk as a whole.  The gate drops a segment when BOTH its
  // end points are more than `lim` degrees from the column's ray -- unless the two angles are equal.  If every point
  // of a chunk lies above ray + lim, or every point below ray - lim (fp32 subtraction is monotone, so the chunk's
  // minimum / maximum decide it for all of them), and no segment has equal or NaN end angles, no segment of the chunk
  // passes the gate and the chunk cannot hold the first hit: it is skipped without looking at its segments
  // (points are in front of the camera, z > 0, so the ray's backward half-line meets none of them).
  // The planner's curtains are ordered by column, so a column's ray meets one or two chunks instead of all of them.
  __shared__ float s_cmin[kPlaneMaxChunks], s_cmax[kPlaneMaxChunks];
  __shared__ int s_cflag[kPlaneMaxChunks];
  const int nchunks = ngood > 1 ? (ngood - 1 + 31) / 32 : 0;
  const bool chunked = nchunks <= kPlaneMaxChunks;
  if (chunked) {
    for (int q = warp; q < nchunks; q += kPlaneThreads / 32) {
      const int j = q * 32 + lane;
      float lo = INFINITY, hi = -INFINITY;
      bool odd = false;                         // a segment whose gate does not depend on the ray: equal or NaN angles
      if (j + 1 < ngood) {
        const float pa = ang[all_pts ? j : good[j]], pb = ang[all_pts ? j + 1 : good[j + 1]];
        odd = !(pa != pb) || pa != pa || pb != pb;
        lo = fminf(pa, pb);
        hi = fmaxf(pa, pb);
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, off));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, off));
      }
      const unsigned any_odd = __ballot_sync(0xffffffffu, odd);
      if (lane == 0) { s_cmin[q] = lo; s_cmax[q] = hi; s_cflag[q] = any_odd != 0u; }
    }
    __syncthreads();
  }

  // Columns of this CTA: [col0, col1).  Each WARP searches one column at a time: its 32 lanes test 32
  // consecutive segments j, and the lowest hitting lane is exactly the reference's "first j that hits".
  const int col0 = blockIdx.y * kPlaneCols, col1 = min(col0 + kPlaneCols, dv.W);
  __shared__ float s_dsg[kPlaneCols][4];
  for (int i = col0 + warp; i < col1; i += kPlaneThreads / 32) {
    // ---- findCameraIntersectionsOpt (dprocessor.cpp:514-634) for column i
    float dx = dv.mid_x[i], dz = dv.mid_z[i];
    float nrm = sqrtf(dx * dx + dz * dz);
    dx = dx / nrm; dz = dz / nrm;
    const float cam_angle = dv.angles[i];
    bool found = false;
    float ix = 0.f, iz = 0.f;
    for (int li = 0; li < 2 && !found; li++) {
      const float lim = li == 0 ? 1.f : 100.f;
      // Chunks in ascending order, 32 chunk summaries per ballot; only the candidates are opened.
      // pass 0 (lim = 1): a chunk whose points all lie more than a degree to one side of the ray fails the angle
      // gate as a whole.  Pass 1 (lim = 100, which passes nearly every segment and exists for segments that SPAN the
      // ray with both ends more than a degree away): such a chunk holds no segment the ray's line can cross -- the
      // crossing point lies outside the segment by sin(1 deg) of its distance, six orders of magnitude above the
      // rounding of the test below -- so the same bound skips it.  Without it a column that no segment covers (image
      // edges outside the curtain) evaluates every segment twice and the CTA waits for it.
      for (int qb = 0; qb < nchunks && !found; qb += 32) {
       const int q = qb + lane;
       bool keep = q < nchunks;
       if (keep && chunked) keep = s_cflag[q] || !(s_cmin[q] - cam_angle > 1.f || cam_angle - s_cmax[q] > 1.f);
       unsigned cand = __ballot_sync(0xffffffffu, keep);
       while (cand && !found) {
        const int j0 = (qb + __ffs(cand) - 1) * 32;
        cand &= cand - 1;
        const int j = j0 + lane;
        bool hit = false;
        float qx = 0.f, qz = 0.f;
        if (j + 1 < ngood) {
          const int a = all_pts ? j : good[j], b = all_pts ? j + 1 : good[j + 1];
          const float pa = ang[a], pb = ang[b];
          bool gate = true;
          if (pa != pb)
            if (fabsf(pa - cam_angle) > lim && fabsf(pb - cam_angle) > lim) gate = false;
          const float p1x = px[a], p1z = pz[a], p2x = px[b], p2z = pz[b];
          if (p1x == p2x && p1z == p2z) gate = false;
          if (gate) {
            float ex = p2x - p1x, ez = p2z - p1z;
            float nx = -ez, nz = ex;
            float n2 = nx * nx + nz * nz;
            if (n2 > 0.f) { float s = sqrtf(n2); nx = nx / s; nz = nz / s; }
            float off = -(p1x * nx + p1z * nz);
            float t = -(off + (nx * 0.f + nz * 0.f)) / (nx * dx + nz * dz);
            qx = 0.f + dx * t; qz = 0.f + dz * t;
            float kk = ex * ex + ez * ez;
            float kp = ex * (qx - p1x) + ez * (qz - p1z);
            if (kp < 0) hit = false;
            else if (kp > kk) hit = false;
            else if (kp == kp && kk == kk) hit = true;
            else hit = false;
          }
        }
        const unsigned mask = __ballot_sync(0xffffffffu, hit);
        if (mask) {
          const int srcl = __ffs(mask) - 1;
          ix = __shfl_sync(0xffffffffu, qx, srcl);
          iz = __shfl_sync(0xffffffffu, qz, srcl);
          found = true;
        }
       }
      }
    }
    if (lane == 0) {
      s_dsg[i - col0][0] = ix; s_dsg[i - col0][1] = iz; s_dsg[i - col0][2] = found ? 1.f : 0.f;
    }
  }
  __syncthreads();
  for (int i = col0 + threadIdx.x; i < col1; i += kPlaneThreads) {
    bool found = s_dsg[i - col0][2] != 0.f;
    const float ix = s_dsg[i - col0][0], iz = s_dsg[i - col0][1];
    float cp = (float)i / float(dv.W);
    if (dv.limit > 0) if (cp > dv.limit) found = false;
    if (dv.limit < 0) if (cp < fabsf(dv.limit)) found = false;
    float dp[4];
    if (found) { dp[0] = ix; dp[1] = 0.f; dp[2] = iz; dp[3] = 1.f; }
    else { dp[0] = 0.f; dp[1] = 0.f; dp[2] = 0.f; dp[3] = -1.f; }
    if (design) {
      float* o = design + ((size_t)c * dv.W + i) * 4;
      o[0] = dp[0]; o[1] = dp[1]; o[2] = dp[2]; o[3] = dp[3];
    }
    // ---- per-column planes (calculateSurface :751-802, :841-863)
    float pl[4];
    for (int r = 0; r < 4; r++) {
      float s = dv.cam_to_laser.m[r][0] * dp[0];
      for (int k = 1; k < 4; k++) s = s + dv.cam_to_laser.m[r][k] * dp[k];
      pl[r] = s;
    }
    float an = (pl[3] == -1) ? NAN : point_angle(pl[0], pl[2]);
    float out[12];
    for (int k = 0; k < 12; k++) out[k] = 0.f;
    float lf[4], cf[4];
    LcMat4 M = transform_matrix_pitch(an);
    rigid_inverse_row0(M, lf);
    plane_to_cam(dv.laser_to_cam_inv, lf, cf);
    for (int k = 0; k < 4; k++) out[k] = cf[k];
    if (!(an != an)) {
      LcMat4 Ml = mat4_mul(mat4_mul(M, pc.off_l), pc.div_l);
      LcMat4 Mr = mat4_mul(mat4_mul(M, pc.off_r), pc.div_r);
      rigid_inverse_row0(Ml, lf);
      plane_to_cam(dv.laser_to_cam_inv, lf, cf);
      for (int k = 0; k < 4; k++) out[4 + k] = cf[k];
      rigid_inverse_row0(Mr, lf);
      plane_to_cam(dv.laser_to_cam_inv, lf, cf);
      for (int k = 0; k < 4; k++) out[8 + k] = cf[k];
    }
    float* o = planes + ((size_t)c * dv.W + i) * 12;
    for (int k = 0; k < 12; k++) o[k] = out[k];
  }
}

}  // namespace dpv

extern "C" int lc_curtain_planes(const LcDevice* dev, const float* xz, const int32_t* counts, int n, int npts,
                                 float* planes, float* design, void* stream) {
  DPV_REQUIRE(dev && xz && planes, "lc_curtain_planes: null pointer");
  DPV_REQUIRE(n > 0 && npts > 0, "lc_curtain_planes: bad shape n=%d npts=%d", n, npts);
  size_t smem = (size_t)npts * 16;
  DPV_REQUIRE(smem <= 200 * 1024, "lc_curtain_planes: npts=%d too large", npts);
  if (smem > 48 * 1024)
    DPV_CUDA(cudaFuncSetAttribute(dpv::curtain_planes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(n, (dev->W + dpv::kPlaneCols - 1) / dpv::kPlaneCols);
  dpv::launch_high_priority(dpv::curtain_planes_kernel, grid, dim3(dpv::kPlaneThreads), smem, (cudaStream_t)stream,
                            dev->view, dev->plane_consts, xz, counts, npts, planes, design);
  return dpv::check_launch("lc_curtain_planes");
}
