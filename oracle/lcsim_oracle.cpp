// CPU oracle for the lcsim half of the light-curtain path.  TEST INFRASTRUCTURE: nothing under
// depthestimation_b200/ may link, load or call this file (only tests/, smoke() and bench.py's
// cpu_baseline / --impl reference legs do).
//
// Dependency-free C++17 restatement of the algorithm in the reference's vendored submodule
// external/lcsim (soulslicer/lcsim @ 06517e766e74), which cannot be built here: it needs Eigen3,
// OpenCV C++, PCL and Boost (external/lcsim/CMakeLists.txt:9-12), none of which are installed.
// PARITY: pinned only by the one known-answer vector the reference ships
// (external/lcsim/python/example/dpv.npy['angles'], see tests/test_oracle_golden.py); the planner
// and the curtain-image formation have no golden outputs in the reference => "parity unpinned"
// for those two, stated in DESIGN.md.
//
// Arithmetic conventions fixed by this restatement (the reference's own depend on its compiler
// flags and on Eigen/OpenCV internals that are not vendored):
//   * fp32 expressions are evaluated left to right with no FMA contraction (-ffp-contract=off);
//   * 4x4 products sum k = 0..3 in order; rigid transforms are inverted in closed form
//     (R^T, -R^T t); a general 4x4 is inverted by Gauss-Jordan with partial pivoting;
//   * cv::undistortPoints with zero distortion is (float)(((double)u - cx) * (1.0 / fx));
//   * per-column trigonometry of the *sensing* path (dprocessor.cpp) uses correctly-rounded
//     fp32 results, (float)sin((double)x) etc., where the reference calls sinf/cosf/atan2f:
//     identical to glibc in all but a few percent of arguments (<= 1 ulp there), and the same
//     value a GPU computes, so CPU and GPU runs of the path can be compared bit for bit.
//     The planner graph (planning.cpp) keeps plain libm float calls.
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace {

inline float sin_cr(float x) { return (float)std::sin((double)x); }
inline float cos_cr(float x) { return (float)std::cos((double)x); }
inline float atan2_cr(float y, float x) { return (float)std::atan2((double)y, (double)x); }

struct Mat4 {
    float m[4][4];
};

Mat4 mat4_identity() {
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) r.m[i][j] = (i == j) ? 1.f : 0.f;
    return r;
}

Mat4 mat4_mul(const Mat4& a, const Mat4& b) {
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float s = a.m[i][0] * b.m[0][j];
            for (int k = 1; k < 4; k++) s = s + a.m[i][k] * b.m[k][j];
            r.m[i][j] = s;
        }
    return r;
}

// General inverse (Gauss-Jordan, partial pivoting) — stands in for Eigen's MatrixXf::inverse()
// (dprocessor.cpp:172).  Exact for the pure-translation lTc of the KITTI-sim set-up.
Mat4 mat4_inverse_general(const Mat4& a) {
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
            for (int j = 0; j < 8; j++) std::swap(w[p][j], w[c][j]);
        float d = w[c][c];
        for (int j = 0; j < 8; j++) w[c][j] = w[c][j] / d;
        for (int r = 0; r < 4; r++) {
            if (r == c) continue;
            float f = w[r][c];
            if (f == 0.f) continue;
            for (int j = 0; j < 8; j++) w[r][j] = w[r][j] - f * w[c][j];
        }
    }
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) r.m[i][j] = w[i][j + 4];
    return r;
}

// util.cpp:5-23 setEulerYPR + util.cpp:25-45 getTransformMatrix(yaw, pitch, roll, x, y, z).
// NB the reference passes (roll, pitch, yaw) into parameters named (eulerZ, eulerY, eulerX).
Mat4 transform_matrix(float yaw, float pitch, float roll, float x, float y, float z) {
    float eulerZ = (float)(roll * M_PI / 180.);
    float eulerY = (float)(pitch * M_PI / 180.);
    float eulerX = (float)(yaw * M_PI / 180.);
    float ci = cos_cr(eulerX), cj = cos_cr(eulerY), ch = cos_cr(eulerZ);
    float si = sin_cr(eulerX), sj = sin_cr(eulerY), sh = sin_cr(eulerZ);
    float cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
    Mat4 r = mat4_identity();
    r.m[0][0] = cj * ch;  r.m[0][1] = sj * sc - cs;  r.m[0][2] = sj * cc + ss;
    r.m[1][0] = cj * sh;  r.m[1][1] = sj * ss + cc;  r.m[1][2] = sj * cs - sc;
    r.m[2][0] = -sj;      r.m[2][1] = cj * si;       r.m[2][2] = cj * ci;
    r.m[0][3] = x; r.m[1][3] = y; r.m[2][3] = z;
    return r;
}

// First row of the inverse of a rigid transform == (M^-1)^T * (1,0,0,0), the "straight plane"
// carried through lrotated_matrix.inverse().transpose() (dprocessor.cpp:781-797, 845-857).
void rigid_inverse_row0(const Mat4& M, float out[4]) {
    out[0] = M.m[0][0];
    out[1] = M.m[1][0];
    out[2] = M.m[2][0];
    float s = M.m[0][0] * M.m[0][3];
    s = s + M.m[1][0] * M.m[1][3];
    s = s + M.m[2][0] * M.m[2][3];
    out[3] = -s;
}

// planes_cframe = Q^T * plane with Q = laser_to_cam.inverse() (dprocessor.cpp:802).
void plane_to_cam(const Mat4& Q, const float pl[4], float out[4]) {
    for (int r = 0; r < 4; r++) {
        float s = Q.m[0][r] * pl[0];
        for (int k = 1; k < 4; k++) s = s + Q.m[k][r] * pl[k];
        out[r] = s;
    }
}

// util.cpp:47-52
inline void intersect(float A, float B, float C, float D, const float ray[3], float out[4]) {
    float t = -D / (A * ray[0] + B * ray[1] + C);
    out[0] = ray[0] * t;
    out[1] = ray[1] * t;
    out[2] = ray[2] * t;
    out[3] = 0.f;
}

struct Vec2 {
    float x, z;
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// Sensor model: DatumProcessor::setSensors / createNormalMap / computeLaserParams
// (dprocessor.cpp:41-235), defaults common.h:102-115 and python/sim.py:42-73.
// ---------------------------------------------------------------------------------------------
struct LcoDevice {
    int W = 0, H = 0;
    float K[9];
    Mat4 cam_to_laser, laser_to_cam, laser_to_cam_inv;
    float fov = 80.f, laser_timestep = 1.5e-5f, laser_limit = 14000.f;
    float thickness = 0.00055f, divergence = 0.11f / 2.f;
    int hit_mode = 0, hit_N = 1000;
    float hit_std = 0.3f, hit_pow = 2.f, hit_noise = 0.f, limit = 1.0f;
    std::vector<float> nn_x, nn_y;        // un-normalised pixel rays: x per column, y per row
    std::vector<float> ztoramap;          // [H*W]
    std::vector<float> mid_x, mid_z;      // normalised mid-row ray (x,z) per column
    std::vector<float> valid_angles;      // per column, degrees
    Vec2 cam_origin{0, 0}, p_left_cam, p_right_cam, laser_origin, p_left_laser, p_right_laser;
    float t_max = 0;
    std::vector<float> hit_lut;
};

static void build_normal_map(LcoDevice& d) {
    const int W = d.W, H = d.H;
    // cv::undistortPoints, zero distortion, 1-based pixel grid (dprocessor.cpp:17-27, 57)
    double fx = d.K[0], fy = d.K[4], cx = d.K[2], cy = d.K[5];
    double ifx = 1. / fx, ify = 1. / fy;
    d.nn_x.resize(W);
    d.nn_y.resize(H);
    for (int i = 0; i < W; i++) d.nn_x[i] = (float)(((double)(float)(i + 1) - cx) * ifx);
    for (int j = 0; j < H; j++) d.nn_y[j] = (float)(((double)(float)(j + 1) - cy) * ify);
    d.ztoramap.resize((size_t)H * W);
    d.mid_x.resize(W);
    d.mid_z.resize(W);
    const int mid = H / 2 - 1;                                     // :82
    for (int i = 0; i < W; i++)
        for (int j = 0; j < H; j++) {
            float x = d.nn_x[i], y = d.nn_y[j], z = 1.0f;
            float mag = std::sqrt(x * x + y * y + z * z);          // :63
            float vx = x / mag, vy = y / mag, vz = z / mag;
            float theta = atan2f(vx, vz);                          // :73
            float phi = asinf(vy / 1);                             // :74
            d.ztoramap[(size_t)j * W + i] = cosf(theta) * cosf(phi);   // :75
            if (j == mid) { d.mid_x[i] = vx; d.mid_z[i] = vz; }
        }
    d.valid_angles.resize(W);
    for (int i = 0; i < W; i++)                                    // :85-91
        d.valid_angles[i] = (float)(-((atan2_cr(d.mid_z[i], d.mid_x[i]) * 180 / M_PI) - 90));
    // camera wedge (:93-107)
    float z_max = 1;
    float t = (float)(1.25 * z_max / d.mid_z[0]);
    d.p_left_cam = {d.cam_origin.x + t * d.mid_x[0], d.cam_origin.z + t * d.mid_z[0]};
    d.p_right_cam = {d.cam_origin.x + t * d.mid_x[W - 1], d.cam_origin.z + t * d.mid_z[W - 1]};
    d.t_max = t;
    // hit LUT (:931-936; algo.cpp:377-385, 676-682)
    const int N = d.hit_N;
    d.hit_lut.resize(N);
    float h = (1.f - (-1.f)) / static_cast<float>(N - 1);
    float val = -1.f;
    for (int k = 0; k < N; k++, val += h) {
        // Eigen::pow(abs(x - mu), power) / (2 * pow(sig, power)) in float arrays; exp in float
        float a = std::pow(std::fabs(val - 0.f), d.hit_pow);
        float den = (float)(2 * std::pow(d.hit_std, d.hit_pow));
        d.hit_lut[k] = std::exp(-a / den);
    }
    float mx = d.hit_lut[0], mn = d.hit_lut[0];
    for (float v : d.hit_lut) { mx = std::max(mx, v); mn = std::min(mn, v); }
    for (float& v : d.hit_lut) v = (v - mn) / (mx - mn);
}

static void build_laser(LcoDevice& d) {
    d.laser_to_cam = mat4_inverse_general(d.cam_to_laser);        // :172
    d.laser_to_cam_inv = mat4_inverse_general(d.laser_to_cam);    // used at :189, :802
    d.laser_origin = {d.laser_to_cam.m[0][3], d.laser_to_cam.m[2][3]};
    float th[2] = {-d.fov / 2.f, d.fov / 2.f};                     // :178-181 (fov > 0)
    Vec2 out[2];
    for (int s = 0; s < 2; s++) {
        // plane through (0,0,0), (0,1,0), (sin t, 0, cos t): n = P0P1 x P0P2, d = P0.n  (:139-154)
        float sx = (float)std::sin(M_PI / 180 * th[s]), cz = (float)std::cos(M_PI / 180 * th[s]);
        float a[3] = {0, 1, 0}, b[3] = {sx, 0, cz};
        float n[4] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0], 0.f};
        // row-vector * matrix (:189)
        float c[4];
        for (int j = 0; j < 4; j++) {
            float acc = n[0] * d.laser_to_cam_inv.m[0][j];
            for (int k = 1; k < 4; k++) acc = acc + n[k] * d.laser_to_cam_inv.m[k][j];
            c[j] = acc;
        }
        float dir[3] = {0, -1, 0};
        float v0 = dir[1] * c[2] - dir[2] * c[1];
        float v2 = dir[0] * c[1] - dir[1] * c[0];
        out[s] = {d.laser_origin.x + d.t_max * v0, d.laser_origin.z + d.t_max * v2};
    }
    d.p_left_laser = out[0];
    d.p_right_laser = out[1];
}

// checkPoints, dprocessor.cpp:238-280
static std::vector<int> check_points(const LcoDevice& d, const std::vector<Vec2>& pts) {
    std::vector<int> good;
    for (int i = 0; i < (int)pts.size(); i++) {
        float x = pts[i].x, z = pts[i].z;
        float x1 = d.cam_origin.x, z1 = d.cam_origin.z;
        float x2 = d.p_left_cam.x, z2 = d.p_left_cam.z;
        float d_cam_left = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
        x2 = d.p_right_cam.x; z2 = d.p_right_cam.z;
        float d_cam_right = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
        x1 = d.laser_origin.x; z1 = d.laser_origin.z;
        x2 = d.p_left_laser.x; z2 = d.p_left_laser.z;
        float d_las_left = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
        x2 = d.p_right_laser.x; z2 = d.p_right_laser.z;
        float d_las_right = (x - x1) * (z2 - z1) - (z - z1) * (x2 - x1);
        if (d_cam_left > 0 && d_las_left > 0 && d_cam_right < 0 && d_las_right < 0) good.push_back(i);
    }
    return good;
}

// findCameraIntersectionsOpt, dprocessor.cpp:514-634.  out[4*i..] = (x, 0, z, 1 | -1)
static void camera_intersections(const LcoDevice& d, const std::vector<int>& good_in,
                                 const std::vector<Vec2>& pts, std::vector<float>& out) {
    std::vector<int> good = good_in;
    if (good.empty()) {
        good.resize(pts.size());
        for (size_t i = 0; i < good.size(); i++) good[i] = (int)i;
    }
    std::vector<float> ang(pts.size());
    for (size_t i = 0; i < pts.size(); i++)
        ang[i] = (float)(-((atan2_cr(pts[i].z, pts[i].x) * 180 / M_PI) - 90) + 0 + 0);
    out.assign((size_t)4 * d.W, 0.f);
    for (int i = 0; i < d.W; i++) {
        // Eigen Vector2f::normalize(): v /= sqrt(squaredNorm)
        float dx = d.mid_x[i], dz = d.mid_z[i];
        float nrm = std::sqrt(dx * dx + dz * dz);
        dx = dx / nrm; dz = dz / nrm;
        float cam_angle = d.valid_angles[i];
        bool found = false;
        float ix = 0, iz = 0;
        const float limits[2] = {1.f, 100.f};
        for (int li = 0; li < 2 && !found; li++) {
            float lim = limits[li];
            found = false;
            // NB the reference loops j < good.size()-1 with an unsigned size(): one point => no segment
            for (size_t j = 0; j + 1 < good.size(); j++) {
                int a = good[j], b = good[j + 1];
                float pa = ang[a], pb = ang[b];
                if (pa != pb)
                    if (std::fabs(pa - cam_angle) > lim && std::fabs(pb - cam_angle) > lim) continue;
                float p1x = pts[a].x, p1z = pts[a].z, p2x = pts[b].x, p2z = pts[b].z;
                if (p1x == p2x && p1z == p2z) continue;
                // Hyperplane<float,2>::Through(p1,p2): normal = unitOrthogonal(p2-p1), offset = -p1.normal
                float ex = p2x - p1x, ez = p2z - p1z;
                float nx = -ez, nz = ex;
                float n2 = nx * nx + nz * nz;
                if (n2 > 0.f) { float s = std::sqrt(n2); nx = nx / s; nz = nz / s; }
                float off = -(p1x * nx + p1z * nz);
                // ParametrizedLine::intersectionPoint with origin (0,0)
                float t = -(off + (nx * 0.f + nz * 0.f)) / (nx * dx + nz * dz);
                float px = 0.f + dx * t, pz = 0.f + dz * t;
                float kk = ex * ex + ez * ez;
                float kp = ex * (px - p1x) + ez * (pz - p1z);
                // :587-602 collapses to 0 <= kp <= kk (either abs overload gives the same outcome)
                if (kp < 0) found = false;
                else if (kp > kk) found = false;
                else if (kp == kp && kk == kk) { found = true; ix = px; iz = pz; }
                else found = false;
                if (found) break;
            }
        }
        float cp = (float)i / float(d.W);
        if (d.limit > 0) if (cp > d.limit) found = false;
        if (d.limit < 0) if (cp < std::fabs(d.limit)) found = false;
        if (found) { out[4 * i] = ix; out[4 * i + 1] = 0; out[4 * i + 2] = iz; out[4 * i + 3] = 1; }
        else { out[4 * i] = 0; out[4 * i + 1] = 0; out[4 * i + 2] = 0; out[4 * i + 3] = -1; }
    }
}

// Per-column plane coefficients in the camera frame: main plane + the two divergence planes.
// calculateSurface, dprocessor.cpp:751-802 and 841-863.  planes[12*i + {0..3 main, 4..7 left, 8..11 right}]
static void column_planes(const LcoDevice& d, const std::vector<float>& dp, std::vector<float>& planes) {
    const float nanv = std::numeric_limits<float>::quiet_NaN();
    planes.assign((size_t)12 * d.W, 0.f);
    for (int i = 0; i < d.W; i++) {
        // design_pts_laser = cam_to_laser * design_pts (:752)
        float pl[4];
        for (int r = 0; r < 4; r++) {
            float s = d.cam_to_laser.m[r][0] * dp[4 * i];
            for (int k = 1; k < 4; k++) s = s + d.cam_to_laser.m[r][k] * dp[4 * i + k];
            pl[r] = s;
        }
        float ang;
        if (pl[3] == -1) ang = nanv;
        else ang = (float)(-((atan2_cr(pl[2], pl[0]) * 180 / M_PI) - 90) + 0 + 0);   // :762
        float lf[4], cf[4];
        Mat4 M = transform_matrix(0, ang, 0, 0, 0, 0);
        rigid_inverse_row0(M, lf);
        plane_to_cam(d.laser_to_cam_inv, lf, cf);
        for (int k = 0; k < 4; k++) planes[12 * i + k] = cf[k];
        if (std::isnan(ang)) continue;                            // :844 (columns stay zero)
        Mat4 Ml = mat4_mul(mat4_mul(transform_matrix(0, ang, 0, 0, 0, 0),
                                    transform_matrix(0, 0, 0, (d.thickness / 2), 0, 0)),
                           transform_matrix(0, -d.divergence, 0, 0, 0, 0));
        Mat4 Mr = mat4_mul(mat4_mul(transform_matrix(0, ang, 0, 0, 0, 0),
                                    transform_matrix(0, 0, 0, -(d.thickness / 2), 0, 0)),
                           transform_matrix(0, d.divergence, 0, 0, 0, 0));
        rigid_inverse_row0(Ml, lf);
        plane_to_cam(d.laser_to_cam_inv, lf, cf);
        for (int k = 0; k < 4; k++) planes[12 * i + 4 + k] = cf[k];
        rigid_inverse_row0(Mr, lf);
        plane_to_cam(d.laser_to_cam_inv, lf, cf);
        for (int k = 0; k < 4; k++) planes[12 * i + 8 + k] = cf[k];
    }
}

// Per-pixel curtain surface, thickness and hit intensity: calculateSurface :806-887 +
// computeDepthHits :942-993.  image [H,W,4] (x,y,z,intensity 0..255), thick [H,W].
// noise: optional [H*W] additive intensity noise (stands in for the unseeded RNG).
static void surface_and_hits(const LcoDevice& d, const std::vector<float>& planes, const float* depth,
                             const float* noise, float* image, float* thick) {
    const float nanv = std::numeric_limits<float>::quiet_NaN();
    const int W = d.W, H = d.H, N = d.hit_N;
    for (int v = 0; v < H; v++)
        for (int u = 0; u < W; u++) {
            const float* P = &planes[12 * u];
            float ray[3] = {d.nn_x[u], d.nn_y[v], 1.0f};
            float* c = &image[((size_t)v * W + u) * 4];
            intersect(P[0], P[1], P[2], P[3], ray, c);
            if (c[2] < 0 || std::fabs(c[1]) > 3.0) { c[0] = c[1] = c[2] = c[3] = nanv; }
            float bot[4], top[4];
            intersect(P[4], P[5], P[6], P[7], ray, bot);
            intersect(P[8], P[9], P[10], P[11], ray, top);
            // pow(float, int) promotes to double (:884-886)
            double dx = (double)(bot[0] - top[0]), dy = (double)(bot[1] - top[1]), dz = (double)(bot[2] - top[2]);
            float unc = (float)std::sqrt(dx * dx + dy * dy + dz * dz);
            thick[(size_t)v * W + u] = unc;
            // computeDepthHits
            if (std::isnan(c[0])) continue;
            if (std::isnan(unc)) continue;
            float zval = depth[(size_t)v * W + u];
            if (std::isnan(zval)) continue;
            float srange = std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
            float drange = zval / d.ztoramap[(size_t)v * W + u];
            float inten = 0.f;
            if (d.hit_mode == 0) {
                float err = std::fabs(drange - srange);
                float color = 0;
                if (err >= 0) color = (float)(1. - 1. * (err / (unc / 2.)));
                if (err < unc / 2.) inten = color;
            } else if (d.hit_mode == 1) {
                float err = drange - srange;
                float scale = (float)(err / (unc / 2.));
                if (std::fabs(err) > unc / 2.) scale = (float)(((0.f < err) - (err < 0.f)) * 1.);
                int idx = (int)(((scale + 1.) / 2.) * (N - 1));
                inten = d.hit_lut[idx];
            }
            if (noise) {
                inten += noise[(size_t)v * W + u];
                inten = std::max(0.f, std::min(inten, 1.f));
            }
            c[3] = (float)(int)(255 * inten);
        }
}

// ---------------------------------------------------------------------------------------------
// Planner: planning.cpp:14-126 (graph + DP), 162-217 (Trajectory), 255-264 (PolarIdentity).
// Dynamic arrays instead of MAX_RAYS x MAX_NODES_PER_RAY.
// ---------------------------------------------------------------------------------------------
#define LCO_PI 3.14159265   /* planning.h:11 (truncated literal) */
static float deg2rad(float deg) { return (float)(deg * LCO_PI / 180); }
static float rad2deg(float rad) { return (float)(rad * 180 / LCO_PI); }

struct LcoPlanner {
    int W = 0, R = 0;
    std::vector<float> x, z, theta_las;                 // [W*R], node (ray i, range k) at i*R+k
    std::vector<std::vector<int>> edges;                // per node: next_i list (ascending)
    float max_d_las = 0;
};

static LcoPlanner* planner_build(const LcoDevice& d, const float* ranges, int R) {
    LcoPlanner* p = new LcoPlanner();
    p->W = d.W; p->R = R;
    p->max_d_las = d.laser_limit * d.laser_timestep;    // planning.cpp:24
    size_t n = (size_t)d.W * R;
    p->x.resize(n); p->z.resize(n); p->theta_las.resize(n); p->edges.resize(n);
    for (int i = 0; i < d.W; i++)
        for (int k = 0; k < R; k++) {
            float r = ranges[k], th = d.valid_angles[i];
            float x = r * std::sin(deg2rad(th));
            float z = r * std::cos(deg2rad(th));
            float c4[4] = {x, 0.0f, z, 1.0f}, l4[4];
            for (int a = 0; a < 4; a++) {               // laser_to_cam_ * xyz1_cam (:46)
                float s = d.laser_to_cam.m[a][0] * c4[0];
                for (int b = 1; b < 4; b++) s = s + d.laser_to_cam.m[a][b] * c4[b];
                l4[a] = s;
            }
            p->x[(size_t)i * R + k] = x;
            p->z[(size_t)i * R + k] = z;
            p->theta_las[(size_t)i * R + k] = rad2deg(std::atan2(l4[0], l4[2]));
        }
    for (int i = 0; i + 1 < d.W; i++)
        for (int a = 0; a < R; a++)
            for (int b = 0; b < R; b++) {
                float dth = p->theta_las[(size_t)(i + 1) * R + b] - p->theta_las[(size_t)i * R + a];
                if ((-p->max_d_las <= dth) && (dth <= p->max_d_las)) p->edges[(size_t)i * R + a].push_back(b);
            }
    return p;
}

// optimizedDesignPts (:75-126).  umap row-major [R, W].  Returns number of design points written.
static int planner_solve(const LcoPlanner& p, const float* umap, float* xz, int32_t* ks) {
    const int W = p.W, R = p.R;
    std::vector<float> unc((size_t)W * R), las((size_t)W * R);
    std::vector<int> sub((size_t)W * R, -1);
    for (int i = W - 1; i >= 0; i--)
        for (int k = 0; k < R; k++) {
            size_t n = (size_t)i * R + k;
            float u = umap[(size_t)k * W + i];
            float bu = u, bl = 0.0f;
            int bs = -1;
            for (int e : p.edges[n]) {
                size_t m = (size_t)(i + 1) * R + e;
                float tu = u;
                tu += unc[m];                                       // :194
                float dth = p.theta_las[m] - p.theta_las[n];
                float tl = dth * dth + las[m];                      // :197-198
                bool gt = (tu > bu) ? true : ((tu == bu) ? (tl < bl) : false);   // :210-217
                if (gt) { bu = tu; bl = tl; bs = e; }
            }
            unc[n] = bu; las[n] = bl; sub[n] = bs;
        }
    // best start on ray 0 against a default (0, 0, null) trajectory (:99-104)
    float bu = 0.0f, bl = 0.0f;
    int bk = -1;
    for (int k = 0; k < R; k++) {
        float tu = unc[k], tl = las[k];
        bool gt = (tu > bu) ? true : ((tu == bu) ? (tl < bl) : false);
        if (gt) { bu = tu; bl = tl; bk = k; }
    }
    if (bk < 0) return 0;    // the reference dereferences a null node here (:116); guarded
    int cnt = 0, i = 0, k = bk;
    while (true) {
        xz[2 * cnt] = p.x[(size_t)i * R + k];
        xz[2 * cnt + 1] = p.z[(size_t)i * R + k];
        if (ks) ks[cnt] = k;
        cnt++;
        int s = sub[(size_t)i * R + k];
        if (s < 0) break;
        k = s; i++;
    }
    return cnt;
}

// ---------------------------------------------------------------------------------------------
// C API (loaded with ctypes by oracle/lcsim.py)
// ---------------------------------------------------------------------------------------------
extern "C" {

// Algo::generateCameraAngles, algo.cpp:696-741 (zero distortion).  K row-major 3x3.
void lco_camera_angles(const float* K, int W, int H, float* out) {
    double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    double ifx = 1. / fx, ify = 1. / fy;
    int j = H / 2 - 1;
    float y = (float)(((double)(float)(j + 1) - cy) * ify);
    for (int i = 0; i < W; i++) {
        float x = (float)(((double)(float)(i + 1) - cx) * ifx);
        float z = 1.0f;
        float mag = std::sqrt(x * x + y * y + z * z);
        float vx = x / mag, vz = z / mag;
        out[i] = (float)(-((atan2f(vz, vx) * 180 / M_PI) - 90));
    }
}

LcoDevice* lco_device_create(int W, int H, const float* K, const float* lTc, float fov, float laser_timestep,
                             float laser_limit, float thickness, float divergence, int hit_mode,
                             float hit_noise, float limit) {
    LcoDevice* d = new LcoDevice();
    d->W = W; d->H = H;
    std::memcpy(d->K, K, sizeof(float) * 9);
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) d->cam_to_laser.m[i][j] = lTc[i * 4 + j];
    d->fov = fov; d->laser_timestep = laser_timestep; d->laser_limit = laser_limit;
    d->thickness = thickness; d->divergence = divergence;
    d->hit_mode = hit_mode; d->hit_noise = hit_noise; d->limit = limit;
    build_normal_map(*d);
    build_laser(*d);
    return d;
}
void lco_device_destroy(LcoDevice* d) { delete d; }
void lco_device_angles(const LcoDevice* d, float* out) { std::memcpy(out, d->valid_angles.data(), sizeof(float) * d->W); }
void lco_device_lut(const LcoDevice* d, float* out) { std::memcpy(out, d->hit_lut.data(), sizeof(float) * d->hit_N); }
// the per-pixel tables the curtain return reads (for the pin against the reference's own functions, oracle/ref_hits.py)
void lco_device_rays(const LcoDevice* d, float* nn_x, float* nn_y) {
    std::memcpy(nn_x, d->nn_x.data(), sizeof(float) * d->W);
    std::memcpy(nn_y, d->nn_y.data(), sizeof(float) * d->H);
}
void lco_device_ztoramap(const LcoDevice* d, float* out) { std::memcpy(out, d->ztoramap.data(), sizeof(float) * (size_t)d->W * d->H); }

// processPointsT (dprocessor.cpp:995-1038) via LCDevice.get_return (python/sim.py:168-211).
// pts [N,2] (x,z); depth [H,W]; noise nullable [H,W]; image [H,W,4]; thick [H,W];
// optional debug outputs: design [W,4], planes [W,12].
int lco_get_return(const LcoDevice* d, const float* depth, const float* pts, int N, const float* noise,
                   float* image, float* thick, float* design_out, float* planes_out) {
    std::vector<Vec2> p(N);
    for (int i = 0; i < N; i++) p[i] = {pts[2 * i], pts[2 * i + 1]};
    std::vector<int> good = check_points(*d, p);
    std::vector<float> design, planes;
    camera_intersections(*d, good, p, design);
    column_planes(*d, design, planes);
    surface_and_hits(*d, planes, depth, noise, image, thick);
    if (design_out) std::memcpy(design_out, design.data(), sizeof(float) * design.size());
    if (planes_out) std::memcpy(planes_out, planes.data(), sizeof(float) * planes.size());
    return (int)good.size();
}

LcoPlanner* lco_planner_create(const LcoDevice* d, const float* ranges, int R) { return planner_build(*d, ranges, R); }
void lco_planner_destroy(LcoPlanner* p) { delete p; }
int lco_planner_solve(const LcoPlanner* p, const float* umap, float* xz, int32_t* ks) {
    return planner_solve(*p, umap, xz, ks);
}
long lco_planner_edge_count(const LcoPlanner* p) {
    long n = 0;
    for (auto& e : p->edges) n += (long)e.size();
    return n;
}
void lco_planner_theta_las(const LcoPlanner* p, float* out) {
    std::memcpy(out, p->theta_las.data(), sizeof(float) * p->theta_las.size());
}

// Depth::transformPoints, depth.cpp:143-222.  pts [N,4] (x,y,z,intensity), thick [N],
// intr row-major 3x3, M row-major 4x4 -> dmap, imap, tmap [height,width].
void lco_transform_points(const float* pts, const float* thickdata, int N, const float* intr, const float* M,
                          int width, int height, int filtering, float* dmap, float* imap, float* tmap) {
    size_t n = (size_t)width * height;
    std::fill(dmap, dmap + n, 0.f); std::fill(imap, imap + n, 0.f); std::fill(tmap, tmap + n, 0.f);
    for (int i = 0; i < N; i++) {
        float q[4] = {pts[4 * i], pts[4 * i + 1], pts[4 * i + 2], 1.0f};
        float c[4];
        for (int r = 0; r < 4; r++) {
            float s = M[r * 4] * q[0];
            for (int k = 1; k < 4; k++) s = s + M[r * 4 + k] * q[k];
            c[r] = s;
        }
        float pr[3];
        for (int r = 0; r < 3; r++) pr[r] = (intr[r * 3] * c[0] + intr[r * 3 + 1] * c[1]) + intr[r * 3 + 2] * c[2];
        float pu = pr[0] / pr[2], pv = pr[1] / pr[2];
        int u = (int)(pu - 0.5), v = (int)(pv - 0.5);
        if (u < 0 || u >= width || v < 0 || v >= height) continue;
        float z = c[2];
        if (std::isnan(z)) continue;
        float cur = dmap[(size_t)v * width + u];
        if ((z < cur) || (cur == 0)) {
            dmap[(size_t)v * width + u] = z;
            imap[(size_t)v * width + u] = pts[4 * i + 3];
            tmap[(size_t)v * width + u] = thickdata[i];
        }
    }
    int off = filtering;
    if (off) {
        std::vector<float> mask(n, 1.f);
        for (int v = off; v < height - off - 1; v++)
            for (int u = off; u < width - off - 1; u++) {
                float z = dmap[(size_t)v * width + u];
                bool bad = false;
                for (int vv = v - off; vv < v + off + 1; vv++) {
                    for (int uu = u - off; uu < u + off + 1; uu++) {
                        if (vv == v && uu == u) continue;
                        float zn = dmap[(size_t)vv * width + uu];
                        if (zn == 0) continue;
                        if ((zn - z) < -1) { bad = true; break; }
                    }
                }
                if (bad) mask[(size_t)v * width + u] = 0;
            }
        for (size_t i = 0; i < n; i++) { dmap[i] *= mask[i]; imap[i] *= mask[i]; tmap[i] *= mask[i]; }
    }
}

}  // extern "C"
