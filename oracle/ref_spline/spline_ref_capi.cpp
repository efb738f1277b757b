// TEST INFRASTRUCTURE -- builds the reference's own Algo::linspace and Algo::fitBSpline
// (external/lcsim/lcsim/src/algo.cpp:377-385, 449-589) into oracle/_ref/libspline_ref.so.
// algo.cpp as a whole needs Eigen proper, OpenCV and poly34, which this image does not have; the two functions this
// path uses touch only element access, row reversal and resizing.  oracle/ref_spline.py therefore cuts exactly those
// two function definitions out of /root/reference at build time into oracle/_ref/spline_ref_extract.inc (git-ignored,
// never committed) and this file compiles them UNMODIFIED between the stand-in types below.  Every arithmetic
// expression of the two functions is plain C++ on float / double scalars obtained through operator(): the stand-ins
// add no arithmetic of their own, so the float / double mix is the reference's.
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <memory>
#include <vector>

namespace Eigen {

template <int N>
class VectorNf {
 public:
  VectorNf() {
    for (int i = 0; i < N; i++) v_[i] = 0.0f;
  }
  VectorNf(float a, float b) : v_{a, b} { static_assert(N == 2, "two scalars"); }
  VectorNf(float a, float b, float c) : v_{a, b, c} { static_assert(N == 3, "three scalars"); }
  float& operator()(int i) { return v_[i]; }
  float operator()(int i) const { return v_[i]; }

 private:
  float v_[N];
};
typedef VectorNf<2> Vector2f;
typedef VectorNf<3> Vector3f;

class MatrixXf {
 public:
  MatrixXf() : r_(0), c_(0) {}
  MatrixXf(long r, long c) : r_(r), c_(c), v_((size_t)(r * c), 0.0f) {}
  long rows() const { return r_; }
  long cols() const { return c_; }
  float& operator()(long i, long j) { return v_[(size_t)(i * c_ + j)]; }
  float operator()(long i, long j) const { return v_[(size_t)(i * c_ + j)]; }
  void resize(long r, long c) {
    r_ = r;
    c_ = c;
    v_.assign((size_t)(r * c), 0.0f);
  }
  void conservativeResize(long r, long c) {                 // keeps the leading block
    MatrixXf n(r, c);
    for (long i = 0; i < std::min(r, r_); i++)
      for (long j = 0; j < std::min(c, c_); j++) n(i, j) = (*this)(i, j);
    *this = n;
  }
  struct ColwiseProxy {
    const MatrixXf& m;
    MatrixXf reverse() const {                              // each column reversed = the row order reversed
      MatrixXf o(m.rows(), m.cols());
      for (long i = 0; i < m.rows(); i++)
        for (long j = 0; j < m.cols(); j++) o(i, j) = m(m.rows() - 1 - i, j);
      return o;
    }
  };
  ColwiseProxy colwise() const { return ColwiseProxy{*this}; }
  struct RowProxy {
    MatrixXf& m;
    long i;
    RowProxy& operator=(const Vector3f& v) {
      for (int j = 0; j < 3; j++) m(i, j) = v(j);
      return *this;
    }
  };
  RowProxy row(long i) { return RowProxy{*this, i}; }

 private:
  long r_, c_;
  std::vector<float> v_;
};

}  // namespace Eigen

namespace lc {

struct SplineParams {                                       // include/lcsim/common.h:160-165
  Eigen::Vector2f p0, p1, p2, p3;
  float Ax, Bx, Cx, Dx;
  float Ay, By, Cy, Dy;
  float A, B, C, D, E;
};
struct SplineParamsVec {                                    // include/lcsim/common.h:167-169
  std::vector<SplineParams> splineParams;
};

class Algo {                                                // the two declarations of include/lcsim/algo.h:31,45
 public:
  static std::vector<float> linspace(float a, float b, size_t N);
  static Eigen::MatrixXf fitBSpline(const Eigen::MatrixXf& input_pts_x, std::shared_ptr<SplineParamsVec>& splineParamsVec,
                                    bool computeCurve = true);
};

}  // namespace lc

using namespace lc;                                         // algo.cpp:4

#include "spline_ref_extract.inc"                           // the reference's two definitions, cut at build time

extern "C" long long ref_fit_bspline(const float* pts, int n_pts, float* out, long long cap, float* coeffs) {
  Eigen::MatrixXf in(n_pts, 3);
  for (int i = 0; i < n_pts; i++)
    for (int j = 0; j < 3; j++) in(i, j) = pts[i * 3 + j];
  std::shared_ptr<SplineParamsVec> params = std::make_shared<SplineParamsVec>();
  Eigen::MatrixXf res = Algo::fitBSpline(in, params, true);
  const long long m = res.rows();
  if (out && cap >= m)
    for (long long i = 0; i < m; i++) {
      out[2 * i] = res(i, 0);
      out[2 * i + 1] = res(i, 1);
    }
  if (coeffs)                                               // [n_pts - 1, 8]: Ax Bx Cx Dx Ay By Cy Dy per segment
    for (size_t s = 0; s < params->splineParams.size(); s++) {
      const SplineParams& p = params->splineParams[s];
      const float c[8] = {p.Ax, p.Bx, p.Cx, p.Dx, p.Ay, p.By, p.Cy, p.Dy};
      for (int k = 0; k < 8; k++) coeffs[s * 8 + k] = c[k];
    }
  return m;
}
