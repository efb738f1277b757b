// TEST INFRASTRUCTURE -- builds the reference's own per-pixel curtain return: Util::intersect
// (external/lcsim/lcsim/src/util.cpp:47-53) and DatumProcessor::computeDepthHits with its helpers signum / clip
// (src/dprocessor.cpp:926-928, 938-940, 942-993) into oracle/_ref/libhits_ref.so.
// dprocessor.cpp as a whole needs OpenCV, Eigen proper and PCL; these four functions use cv::Mat / cv::Vec / VectorXf
// only as CONTAINERS (size(), at<T>(v,u), operator[] / ()): every arithmetic expression is plain C++ on float / double
// scalars.  oracle/ref_hits.py cuts the four definitions UNMODIFIED out of /root/reference at build time into
// oracle/_ref/hits_ref_extract.inc (git-ignored) and this file compiles them between the stand-in containers below,
// which add no arithmetic of their own.  What this does NOT cover: the plane coefficients (Eigen 4x4 inverses and
// products inside calculateSurface) and the hit look-up table (Eigen's vectorised exp / pow in Algo::gaussian) -- they
// are inputs here.
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <map>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace Eigen {
class VectorXf {
 public:
  VectorXf() {}
  explicit VectorXf(size_t n) : v_(n, 0.0f) {}
  float& operator()(long i) { return v_[(size_t)i]; }
  float operator()(long i) const { return v_[(size_t)i]; }
  long size() const { return (long)v_.size(); }
 private:
  std::vector<float> v_;
};
}  // namespace Eigen

namespace cv {
template <int N>
struct Vecf {
  float val[N];
  float& operator[](int i) { return val[i]; }
  const float& operator[](int i) const { return val[i]; }
  float& operator()(int i) { return val[i]; }
  const float& operator()(int i) const { return val[i]; }
};
typedef Vecf<3> Vec3f;
typedef Vecf<4> Vec4f;
struct Size {
  int width, height;
  bool operator!=(const Size& o) const { return width != o.width || height != o.height; }
};
class Mat {                                                // a [rows, cols] image of T-sized elements
 public:
  Mat() : rows(0), cols(0), elem_(0) {}
  Mat(int r, int c, size_t elem) : rows(r), cols(c), elem_(elem), data_((size_t)r * c * elem) {}
  Size size() const { return Size{cols, rows}; }
  template <class T> T& at(int v, int u) { return *reinterpret_cast<T*>(&data_[((size_t)v * cols + u) * elem_]); }
  template <class T> const T& at(int v, int u) const { return *reinterpret_cast<const T*>(&data_[((size_t)v * cols + u) * elem_]); }
  int rows, cols;
 private:
  size_t elem_;
  std::vector<unsigned char> data_;
};
}  // namespace cv

namespace lc {
struct Datum {                                             // the fields computeDepthHits reads (include/lcsim/common.h:105-140)
  int hit_mode = 0;
  int hit_N = 1000;
  float hit_noise = 0.;
  Eigen::VectorXf hit_sample;
  std::normal_distribution<float> hit_dist;
  cv::Mat ztoramap;
};
class Util {
 public:
  static void intersect(float A, float B, float C, float D, const cv::Vec3f& rayEq, cv::Vec4f& coord3D);
};
class DatumProcessor {
 public:
  std::mt19937 gen_;
  void computeDepthHits(std::pair<cv::Mat, cv::Mat>& surface_data, const cv::Mat& depth_img, const Datum& cam_data);
};
}  // namespace lc

using namespace lc;                                        // dprocessor.cpp / util.cpp both open with it

#include "hits_ref_extract.inc"                            // the reference's four definitions, cut at build time

// planes [W,4] (A,B,C,D per column), rays nn_x [W], nn_y [H] -> xyz0 [H,W,4] (x, y, z, 0) = Util::intersect per pixel
extern "C" void ref_intersect(const float* planes, const float* nn_x, const float* nn_y, int W, int H, float* out) {
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++) {
      cv::Vec3f ray;
      ray[0] = nn_x[u]; ray[1] = nn_y[v]; ray[2] = 1.0f;
      cv::Vec4f c;
      Util::intersect(planes[4 * u], planes[4 * u + 1], planes[4 * u + 2], planes[4 * u + 3], ray, c);
      for (int k = 0; k < 4; k++) out[((size_t)v * W + u) * 4 + k] = c[k];
    }
}

// surface [H,W,4] (x, y, z, anything; NaN x = no return), unc [H,W], depth [H,W], ztoramap [H,W], lut [N]
// -> surface[...,3] = (int)(255 * intensity) where the reference writes it (hit_noise = 0)
extern "C" int ref_compute_depth_hits(float* surface, const float* unc, const float* depth, const float* ztoramap,
                                      const float* lut, int N, int mode, int W, int H) {
  std::pair<cv::Mat, cv::Mat> sd(cv::Mat(H, W, sizeof(cv::Vec4f)), cv::Mat(H, W, sizeof(float)));
  cv::Mat dimg(H, W, sizeof(float));
  Datum cam;
  cam.hit_mode = mode; cam.hit_N = N; cam.hit_noise = 0.f;
  cam.hit_sample = Eigen::VectorXf((size_t)N);
  for (int i = 0; i < N; i++) cam.hit_sample(i) = lut[i];
  cam.ztoramap = cv::Mat(H, W, sizeof(float));
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++) {
      const size_t p = (size_t)v * W + u;
      for (int k = 0; k < 4; k++) sd.first.at<cv::Vec4f>(v, u)[k] = surface[p * 4 + k];
      sd.second.at<float>(v, u) = unc[p];
      dimg.at<float>(v, u) = depth[p];
      cam.ztoramap.at<float>(v, u) = ztoramap[p];
    }
  DatumProcessor proc;
  try {
    proc.computeDepthHits(sd, dimg, cam);
  } catch (const std::exception&) {
    return 1;
  }
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++)
      for (int k = 0; k < 4; k++) surface[((size_t)v * W + u) * 4 + k] = sd.first.at<cv::Vec4f>(v, u)[k];
  return 0;
}
