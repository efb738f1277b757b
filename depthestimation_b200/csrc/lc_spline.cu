// (f)3 plan_m1's curve through the sampled points: pylc.fitBSpline = Algo::fitBSpline
// (external/lcsim/lcsim/src/algo.cpp:449-589, linspace :377-385), called by lc/light_curtain.py:580-582.
// Host-only like the reference's (a few hundred control points, a tridiagonal solve and a polyline): no kernel here.
// The arithmetic keeps the reference's mix of float and double, expression by expression (the host code of this
// library is compiled with -ffp-contract=off): the caller truncates the curve to pixels, so a last-bit difference can
// move a sample across a pixel border.
#include <cmath>
#include <vector>

#include "common.cuh"

namespace {

// algo.cpp:377-385: N values starting at a, each the previous one plus h (accumulated, not k * h)
std::vector<float> ref_linspace(float a, float b, size_t N) {
  const float h = (b - a) / static_cast<float>(N - 1);
  std::vector<float> xs(N);
  float val = a;
  for (size_t k = 0; k < N; k++, val += h) xs[k] = val;
  return xs;
}

}  // namespace

extern "C" int lc_fit_bspline(const float* pts, int n_pts, float* out, long long cap, long long* n_out) {
  DPV_REQUIRE(pts && n_out, "lc_fit_bspline: null pointer");
  // the reference reads row 1 of a one-row input (undefined behaviour): two points is the least it is defined for
  DPV_REQUIRE(n_pts >= 2, "lc_fit_bspline: needs at least 2 points, got %d", n_pts);
  // :460-471 rows reversed, two control points appended (means of the first two and of the last two rows)
  const int rows = n_pts + 2;
  std::vector<float> in((size_t)rows * 3);
  for (int i = 0; i < n_pts; i++)
    for (int k = 0; k < 3; k++) in[(size_t)i * 3 + k] = pts[(size_t)(n_pts - 1 - i) * 3 + k];
  auto I = [&](int r, int c) -> float& { return in[(size_t)r * 3 + c]; };
  const int last = n_pts - 1;
  for (int k = 0; k < 3; k++) {
    I(rows - 2, k) = (float)((I(0, k) + I(1, k)) / 2.);
    I(rows - 1, k) = (float)((I(last, k) + I(last - 1, k)) / 2.);
  }
  const int n = rows - 2;                                           // :474
  std::vector<float> dx(n, 0.f), dy(n, 0.f), Ax(n, 0.f), Ay(n, 0.f), Bi(n, 0.f);
  const float DIV = 3.f;                                            // mode 1, :485-491
  dx[0] = (I(1, 0) - I(0, 0)) / DIV;
  dy[0] = (I(1, 1) - I(0, 1)) / DIV;
  dx[n - 1] = (I(n - 1, 0) - I(n - 2, 0)) / DIV;
  dy[n - 1] = (I(n - 1, 1) - I(n - 2, 1)) / DIV;
  Bi[1] = (float)(-1. / I(1, 2));                                   // :497-499
  Ax[1] = -(I(2, 0) - I(0, 0) - dx[0]) * Bi[1];
  Ay[1] = -(I(2, 1) - I(0, 1) - dy[0]) * Bi[1];
  for (int i = 2; i < n - 1; i++) {                                 // :500-504
    Bi[i] = -1 / (I(i, 2) + Bi[i - 1]);
    Ax[i] = -(I(i + 1, 0) - I(i - 1, 0) - Ax[i - 1]) * Bi[i];
    Ay[i] = -(I(i + 1, 1) - I(i - 1, 1) - Ay[i - 1]) * Bi[i];
  }
  for (int i = n - 2; i > 0; i--) {                                 // :505-508
    dx[i] = Ax[i] + dx[i + 1] * Bi[i];
    dy[i] = Ay[i] + dy[i + 1] * Bi[i];
  }
  std::vector<float> path;                                          // x0, y0, x1, y1, ...
  path.push_back(I(0, 0));
  path.push_back(I(0, 1));
  for (int i = 0; i < n - 1; i++) {                                 // :512-543
    const float dist = (float)std::sqrt(std::pow((double)(I(i, 0) - I(i + 1, 0)), 2) + std::pow((double)(I(i, 1) - I(i + 1, 1)), 2));
    const float count = (float)((int)(dist / 0.01));
    const float p0x = I(i, 0), p0y = I(i, 1);
    const float p1x = I(i, 0) + dx[i], p1y = I(i, 1) + dy[i];
    const float p2x = I(i + 1, 0) - dx[i + 1], p2y = I(i + 1, 1) - dy[i + 1];
    const float p3x = I(i + 1, 0), p3y = I(i + 1, 1);
    const float cAx = p3x - 3 * p2x + 3 * p1x - p0x, cBx = 3 * (p2x - 2 * p1x + p0x), cCx = 3 * (p1x - p0x), cDx = p0x;
    const float cAy = p3y - 3 * p2y + 3 * p1y - p0y, cBy = 3 * (p2y - 2 * p1y + p0y), cCy = 3 * (p1y - p0y), cDy = p0y;
    if (!(count >= 1.f)) continue;                                  // (size_t)0 values: nothing to add
    const std::vector<float> ts = ref_linspace((float)0., (float)(1 + 0.0 - (1. / count)), (size_t)count);
    for (float t : ts) {
      const float X = t * t * t * cAx + t * t * cBx + t * cCx + cDx;
      const float Y = t * t * t * cAy + t * t * cBy + t * cCy + cDy;
      path.push_back(X);
      path.push_back(Y);
    }
  }
  if (path.size() == 2) {                                           // :560-567 a single point: two more beside it
    const float x = path[0], y = path[1];
    path.push_back((float)(x + 0.2)); path.push_back(y);
    path.push_back((float)(x - 0.2)); path.push_back(y);
  }
  const long long m = (long long)(path.size() / 2);
  *n_out = m;
  if (!out || cap < m) {
    DPV_REQUIRE(out == nullptr, "lc_fit_bspline: output holds %lld points, %lld needed", cap, m);
    return DPV_OK;                                                  // size query
  }
  for (long long k = 0; k < m; k++) {                               // :569-571 the path comes back reversed
    out[2 * k] = path[(size_t)(2 * (m - 1 - k))];
    out[2 * k + 1] = path[(size_t)(2 * (m - 1 - k) + 1)];
  }
  return DPV_OK;
}
