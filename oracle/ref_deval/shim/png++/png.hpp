// TEST INFRASTRUCTURE -- stand-in for <png++/png.hpp>, first on the include path when oracle/ref_deval.py compiles the
// reference's external/deval_lib/src/evaluate_depth.h VERBATIM.  io_depth.h and utils.h (which evaluate_depth.h
// includes from its own directory) use png++ only to read / write PNG files; depthError and evaluateErrors, the two
// functions on this path, never touch a file.  This header declares just enough for those headers to compile; every
// file operation throws.
#ifndef DPV_REF_DEVAL_PNG_STANDIN
#define DPV_REF_DEVAL_PNG_STANDIN
#include <cstddef>
#include <cstdint>
#include <istream>
#include <stdexcept>
#include <string>
#include <vector>

namespace png {

enum color_type { color_type_none = -1, color_type_gray = 0, color_type_rgb = 2 };

struct rgb_pixel {
  uint8_t red, green, blue;
  rgb_pixel() : red(0), green(0), blue(0) {}
  rgb_pixel(uint8_t r, uint8_t g, uint8_t b) : red(r), green(g), blue(b) {}
};
typedef uint16_t gray_pixel_16;

template <class Pixel>
class image {
 public:
  image(size_t w, size_t h) : w_(w), h_(h), px_(w * h) {}
  explicit image(const std::string&) { throw std::runtime_error("png stand-in: no file I/O"); }
  size_t get_width() const { return w_; }
  size_t get_height() const { return h_; }
  Pixel get_pixel(size_t u, size_t v) const { return px_[v * w_ + u]; }
  void set_pixel(size_t u, size_t v, Pixel p) { px_[v * w_ + u] = p; }
  void write(const std::string&) { throw std::runtime_error("png stand-in: no file I/O"); }

 private:
  size_t w_ = 0, h_ = 0;
  std::vector<Pixel> px_;
};

template <class Stream>
class reader {
 public:
  explicit reader(Stream&) {}
  void read_info() { throw std::runtime_error("png stand-in: no file I/O"); }
  color_type get_color_type() const { return color_type_none; }
  size_t get_bit_depth() const { return 0; }
  size_t get_width() const { return 0; }
  size_t get_height() const { return 0; }
};

}  // namespace png
#endif
