// TEST INFRASTRUCTURE -- stand-in for lcsim's common.h, placed first on the include path so that the reference's
// planning.cpp (external/lcsim/lcsim/src/planning.cpp, compiled verbatim from /root/reference by oracle/ref_planner.py)
// builds without Eigen / OpenCV / PCL.  It declares exactly what planning.h and planning.cpp touch:
//   Eigen::MatrixXf (rows, cols, size, operator()), Eigen::Matrix4f (assignment from MatrixXf, product with a
//   Vector4f), Eigen::Vector4f (4-scalar constructor, operator()), lc::Laser (laser_limit, laser_timestep,
//   laser_to_cam), lc::Datum (valid_angles, laser_data).
// Arithmetic that could differ from Eigen's: only the 4x4 * 4x1 product (planning.cpp:47).  It is evaluated row by
// row, left to right, without FMA.  For a laser_to_cam that is a pure translation (the KITTI-sim rig of this path:
// lTc = translate x by 0.2 m) every product is exact and all but two terms are zero, so the result does not depend
// on the accumulation order; for a rotated laser the order is an assumption.
#ifndef DPV_REF_PLANNER_SHIM_COMMON_H
#define DPV_REF_PLANNER_SHIM_COMMON_H

#include <cmath>
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace Eigen {

class MatrixXf {                       // dynamic, column-major like Eigen's default
 public:
  MatrixXf() : r_(0), c_(0) {}
  MatrixXf(long r, long c) : r_(r), c_(c), v_((size_t)(r * c), 0.0f) {}
  long rows() const { return r_; }
  long cols() const { return c_; }
  long size() const { return r_ * c_; }
  float& operator()(long i, long j) { return v_[(size_t)(j * r_ + i)]; }
  float operator()(long i, long j) const { return v_[(size_t)(j * r_ + i)]; }

 private:
  long r_, c_;
  std::vector<float> v_;
};

class Vector4f {
 public:
  Vector4f() : v_{0, 0, 0, 0} {}
  Vector4f(float a, float b, float c, float d) : v_{a, b, c, d} {}
  float& operator()(int i) { return v_[i]; }
  float operator()(int i) const { return v_[i]; }

 private:
  float v_[4];
};

class Matrix4f {
 public:
  Matrix4f() {
    for (int i = 0; i < 16; i++) m_[i] = 0.0f;
  }
  Matrix4f(const MatrixXf& o) { *this = o; }
  Matrix4f& operator=(const MatrixXf& o) {
    if (o.rows() != 4 || o.cols() != 4) throw std::invalid_argument("Matrix4f = MatrixXf: not 4x4");
    for (int j = 0; j < 4; j++)
      for (int i = 0; i < 4; i++) m_[j * 4 + i] = o(i, j);
    return *this;
  }
  float& operator()(int i, int j) { return m_[j * 4 + i]; }
  float operator()(int i, int j) const { return m_[j * 4 + i]; }
  Vector4f operator*(const Vector4f& x) const {
    Vector4f y;
    for (int i = 0; i < 4; i++) {
      volatile float acc = (*this)(i, 0) * x(0);        // volatile: every product and sum is rounded to fp32 as written
      acc = acc + (*this)(i, 1) * x(1);
      acc = acc + (*this)(i, 2) * x(2);
      acc = acc + (*this)(i, 3) * x(3);
      y(i) = acc;
    }
    return y;
  }

 private:
  float m_[16];
};

}  // namespace Eigen

namespace lc {

class Laser {
 public:
  Eigen::MatrixXf cam_to_laser, laser_to_cam;
  float laser_limit = 14000;
  float laser_timestep = 1.5e-5f;
};

class Datum {
 public:
  std::map<std::string, Laser> laser_data;
  std::vector<float> valid_angles;
};

}  // namespace lc

#endif
