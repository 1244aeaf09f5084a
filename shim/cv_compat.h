// Minimal stand-in for the handful of OpenCV core types the ORB front-end API uses
// (cv::Point2f, cv::KeyPoint, cv::Mat for CV_8UC1/3/4 and CV_32F, InputArray/OutputArray), so the drop-in
// classes in this directory compile in an image without OpenCV C++ headers.  In the reference
// tree this header is NOT used: define CUORB_SHIM_USE_OPENCV and <opencv2/core.hpp> supplies the
// real types (cuorb_keypoint is layout-identical to cv::KeyPoint, so the shim memcpy's).
#pragma once

// ---- error convention of the shim ----------------------------------------------------------------------------
// The reference's path uses no C++ exceptions: failure is -1 / empty outputs / an early return in Frame
// (src/ORBextractor.cc:1060-1062, src/Frame.cc:216-218).  The drop-in classes follow it: a libcuorb error
// is reported once on stderr (with cuorb_last_error()'s text) and the method returns its neutral value - operator()
// -1 with the outputs untouched, a search 0 matches, ComputeStereoMatches all -1.  Constructors are the exception
// (they cannot return a status): they throw std::runtime_error when no CUDA device / arena is available.
// Define CUORB_SHIM_THROW to get exceptions from every method instead.
#include <cstdio>
#include <stdexcept>
#include <string>
#ifdef CUORB_SHIM_THROW
#define CUORB_SHIM_FAIL(ExcType, msg, ...) throw ExcType(msg)
#else
#define CUORB_SHIM_FAIL(ExcType, msg, ...)                                         \
  do {                                                                             \
    std::fprintf(stderr, "[cuorb shim] %s\n", std::string(msg).c_str());           \
    return __VA_ARGS__;                                                            \
  } while (0)
#endif

#ifdef CUORB_SHIM_USE_OPENCV
#include <opencv2/core.hpp>
#else
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_8UC4 24
#define CV_32F 5
#define CV_32FC1 5

namespace cv {

struct Point2f { float x = 0, y = 0; Point2f() = default; Point2f(float x_, float y_) : x(x_), y(y_) {} };
struct Point { int x = 0, y = 0; };
typedef Point Point2i;

// same field order and size as cv::KeyPoint (28 bytes)
struct KeyPoint {
  Point2f pt;
  float   size = 0, angle = -1, response = 0;
  int     octave = 0, class_id = -1;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

// reference-counted 2-D byte/float matrix with a row step, enough for images, descriptor
// matrices (n x 32, CV_8U) and row views (Mat::row) as ORBmatcher::DescriptorDistance takes them
class Mat {
public:
  int rows = 0, cols = 0;
  size_t step = 0;
  unsigned char* data = nullptr;

  Mat() = default;
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(int r, int c, int type, void* external, size_t step_ = 0)
      : rows(r), cols(c), step(step_ ? step_ : (size_t)c * esz(type)), data((unsigned char*)external), type_(type) {}

  void create(int r, int c, int type) {
    if (r == rows && c == cols && type == type_ && data && own_) return;
    rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
    own_ = std::shared_ptr<unsigned char>(new unsigned char[(size_t)r * step + 1], std::default_delete<unsigned char[]>());
    data = own_.get();
  }
  void release() { own_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int  type() const { return type_; }
  int  channels() const { return (type_ >> 3) + 1; }
  Mat  row(int y) const { Mat m; m.rows = 1; m.cols = cols; m.step = step; m.data = data + (size_t)y * step; m.type_ = type_; m.own_ = own_; return m; }
  Mat  rowRange(int a, int b) const { Mat m = row(a); m.rows = b - a; return m; }
  Mat  clone() const { Mat m(rows, cols, type_); for (int y = 0; y < rows; y++) std::memcpy(m.data + y * m.step, data + y * step, (size_t)cols * esz(type_)); return m; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  template <typename T> T& at(int y, int x) { return ptr<T>(y)[x]; }
  template <typename T> const T& at(int y, int x) const { return ptr<T>(y)[x]; }
  Mat getMat() const { return *this; }   // so InputArray/OutputArray code reads like OpenCV's

private:
  static size_t esz(int type) { return (size_t)((type & 7) == CV_32F ? 4 : 1) * ((type >> 3) + 1); }
  int type_ = CV_8UC1;
  std::shared_ptr<unsigned char> own_;
};

typedef const Mat& InputArray;
typedef Mat&       OutputArray;

}  // namespace cv
#endif
