// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// Stand-in for <opencv2/core.hpp>, just wide enough that the reference's OWN sources
// (/root/reference/src/ORBextractor.cc, src/ORBmatcher.cc and the function bodies lifted out of
// src/Frame.cc by oracle/ref_extract.py) compile BYTE-FOR-BYTE UNMODIFIED into oracle/_ref/
// (recipe: oracle/Makefile, target `ref`).  The image has no OpenCV C++ headers; the five OpenCV
// functions the path calls are forwarded (cv_impl.cpp) to the restated primitives of
// oracle/orb_oracle.cpp, which tests/test_oracle_golden.py pins bit-for-bit against the real
// cv2 4.13.0.  Everything above those five calls - std::list / std::sort quadtree, cell loop,
// output scatter, stereo association, projection searches - is then the reference's own code.
//
// Semantics follow OpenCV where the reference relies on them: Mat is a reference-counted header
// (copy = share), ROI views alias their parent, create() is a no-op when size and type already
// match (so cv::resize / cv::copyMakeBorder write THROUGH the pyramid ROI into `temp`,
// src/ORBextractor.cc:1146-1170), Point_<int>(float, float) truncates, cvRound is round-half-even.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_PI 3.1415926535897932384626433832795
#define CV_CN_SHIFT 3
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 511) + 1)
#define CV_MAKETYPE(depth, cn) (((depth) & 7) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)

#define REF_COMPAT_UNSUPPORTED(what)                                                   \
  do {                                                                                 \
    std::fprintf(stderr, "oracle/ref_compat: unsupported OpenCV use: %s (%s:%d)\n", what, \
                 __FILE__, __LINE__);                                                  \
    std::abort();                                                                      \
  } while (0)

// cvRound = cvtss2si / cvtsd2si under the default rounding mode: round half to even
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

template <typename T> struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
  template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
  Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }  // saturate_cast<float> is the identity
  Point_& operator*=(double s) { x = (T)(x * s); y = (T)(y * s); return *this; }
  Point_& operator*=(int s) { x = (T)(x * s); y = (T)(y * s); return *this; }
  Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
  Point_& operator-=(const Point_& o) { x -= o.x; y -= o.y; return *this; }
};
template <typename T> static inline Point_<T> operator+(Point_<T> a, const Point_<T>& b) { return a += b; }
template <typename T> static inline Point_<T> operator-(Point_<T> a, const Point_<T>& b) { return a -= b; }
template <typename T> static inline bool operator==(const Point_<T>& a, const Point_<T>& b) { return a.x == b.x && a.y == b.y; }
typedef Point_<int>    Point2i;
typedef Point_<int>    Point;
typedef Point_<float>  Point2f;
typedef Point_<double> Point2d;

template <typename T> struct Point3_ {
  T x, y, z;
  Point3_() : x(0), y(0), z(0) {}
  Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

template <typename T> struct Size_ {
  T width, height;
  Size_() : width(0), height(0) {}
  Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T> struct Rect_ {
  T x, y, width, height;
  Rect_() : x(0), y(0), width(0), height(0) {}
  Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Range {
  int start, end;
  Range() : start(0), end(0) {}
  Range(int s, int e) : start(s), end(e) {}
};

struct KeyPoint {  // field order and size of the real cv::KeyPoint (28 bytes)
  Point2f pt;
  float   size, angle, response;
  int     octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0,
           int class_id_ = -1)
      : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
  KeyPoint(Point2f p, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0,
           int class_id_ = -1)
      : pt(p), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

class _InputArray;
class _OutputArray;
typedef const _InputArray&  InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;

class Mat {
public:
  int    flags = CV_8UC1;
  int    rows = 0, cols = 0;
  uchar* data = nullptr;
  size_t step = 0;

  Mat() = default;
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(Size sz, int type) { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, void* external, size_t step_ = 0)
      : flags(type), rows(r), cols(c), data((uchar*)external),
        step(step_ ? step_ : (size_t)c * esz(type)) {}
  // ROI constructors / operators: views that alias the parent
  Mat(const Mat& m, const Rect& r) : Mat(m) { roi(r.x, r.y, r.width, r.height); }
  Mat operator()(const Rect& r) const { return Mat(*this, r); }
  Mat operator()(Range rr, Range cr) const { Mat m(*this); m.roi(cr.start, rr.start, cr.end - cr.start, rr.end - rr.start); return m; }
  Mat rowRange(int a, int b) const { Mat m(*this); m.roi(0, a, cols, b - a); return m; }
  Mat colRange(int a, int b) const { Mat m(*this); m.roi(a, 0, b - a, rows); return m; }
  Mat rowRange(const Range& r) const { return rowRange(r.start, r.end); }
  Mat colRange(const Range& r) const { return colRange(r.start, r.end); }
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }

  void create(int r, int c, int type) {
    if (data && r == rows && c == cols && type == flags) return;  // OpenCV: nothing to do
    flags = type; rows = r; cols = c; step = (size_t)c * esz(type);
    own_ = std::shared_ptr<uchar>(new uchar[(size_t)r * step + 16](), std::default_delete<uchar[]>());
    data = own_.get();
    submat_ = false;
  }
  void create(Size sz, int type) { create(sz.height, sz.width, type); }
  void release() { own_.reset(); data = nullptr; rows = cols = 0; step = 0; submat_ = false; }
  Mat  clone() const {
    Mat m;
    if (empty()) return m;
    m.create(rows, cols, flags);
    for (int y = 0; y < rows; y++) std::memcpy(m.ptr(y), ptr(y), (size_t)cols * elemSize());
    return m;
  }
  void copyTo(OutputArray dst) const;  // below, needs _OutputArray
  static Mat zeros(int r, int c, int type) { Mat m(r, c, type); for (int y = 0; y < r; y++) std::memset(m.ptr(y), 0, (size_t)c * m.elemSize()); return m; }
  static Mat zeros(Size sz, int type) { return zeros(sz.height, sz.width, type); }

  bool   empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int    type() const { return flags; }
  int    depth() const { return CV_MAT_DEPTH(flags); }
  int    channels() const { return CV_MAT_CN(flags); }
  size_t elemSize() const { return esz(flags); }
  size_t elemSize1() const { return esz(CV_MAT_DEPTH(flags)); }
  size_t step1() const { return step / elemSize1(); }
  size_t total() const { return (size_t)rows * cols; }
  Size   size() const { return Size(cols, rows); }
  bool   isContinuous() const { return step == (size_t)cols * elemSize() || rows == 1; }
  bool   isSubmatrix() const { return submat_; }

  uchar*       ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
  template <typename T> T*       ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  template <typename T> T&       at(int y, int x) { return ((T*)(data + (size_t)y * step))[x]; }
  template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step))[x]; }

private:
  static size_t esz(int type) {
    static const size_t d[8] = {1, 1, 2, 2, 4, 4, 8, 2};
    return d[CV_MAT_DEPTH(type)] * CV_MAT_CN(type);
  }
  void roi(int x, int y, int w, int h) {
    if (x < 0 || y < 0 || w < 0 || h < 0 || x + w > cols || y + h > rows)
      REF_COMPAT_UNSUPPORTED("ROI outside the matrix (cv::Exception in the real library)");
    data += (size_t)y * step + (size_t)x * elemSize();
    if (w != cols || h != rows) submat_ = true;
    cols = w; rows = h;
  }
  std::shared_ptr<uchar> own_;
  bool                   submat_ = false;
};

// InputArray / OutputArray proxies holding a Mat header, as the real ones do for Mat arguments
class _InputArray {
public:
  _InputArray() {}
  _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
  Mat  getMat() const { return m_ ? *m_ : Mat(); }
  bool empty() const { return !m_ || m_->empty(); }
  Size size() const { return m_ ? m_->size() : Size(); }
  int  type() const { return m_ ? m_->type() : 0; }
protected:
  Mat* m_ = nullptr;
};
class _OutputArray : public _InputArray {
public:
  _OutputArray() {}
  _OutputArray(Mat& m) : _InputArray(m) {}
  _OutputArray(const Mat& m) : _InputArray(m) {}  // temporaries such as m.row(i), as in OpenCV
  void create(int r, int c, int type) const { need(); m_->create(r, c, type); }
  void create(Size sz, int type) const { need(); m_->create(sz, type); }
  void release() const { need(); m_->release(); }
  Mat& getMatRef() const { need(); return *m_; }
private:
  void need() const { if (!m_) REF_COMPAT_UNSUPPORTED("noArray() used as an output"); }
};
static inline InputArray  noArray() { static const _OutputArray none; return none; }

inline void Mat::copyTo(OutputArray dst) const {
  dst.create(rows, cols, flags);
  Mat& d = dst.getMatRef();
  for (int y = 0; y < rows; y++)
    if (d.ptr(y) != ptr(y)) std::memmove(d.ptr(y), ptr(y), (size_t)cols * elemSize());
}

enum BorderTypes {
  BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3,
  BORDER_REFLECT_101 = 4, BORDER_TRANSPARENT = 5, BORDER_REFLECT101 = BORDER_REFLECT_101,
  BORDER_DEFAULT = BORDER_REFLECT_101, BORDER_ISOLATED = 16
};
enum NormTypes { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4, NORM_HAMMING = 6 };

// forwarded to the cv2-certified primitives (oracle/ref_compat/cv_impl.cpp)
void   copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right,
                      int borderType);
double norm(InputArray a, InputArray b, int normType = NORM_L2);
float  fastAtan2(float y, float x);

}  // namespace cv
