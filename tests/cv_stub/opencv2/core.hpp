// TEST INFRASTRUCTURE ONLY -- a stand-in for the few OpenCV core types the
// drop-in shim (include/xpano_b200_shim.hpp) touches. OpenCV's C++ headers
// are not in this image; this header declares just enough of cv::Mat,
// cv::InputArray / OutputArray, cv::Ptr and the geometry structs -- with the
// member names and meanings of OpenCV 4.x -- for the shim to compile and run
// exactly as it would against the real headers. Nothing here is OpenCV code.
#pragma once

#include <cstddef>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_16U 2
#define CV_32F 5
#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_16UC3 CV_MAKETYPE(CV_16U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_Assert(expr)                                              \
  do {                                                               \
    if (!(expr)) throw cv::Exception("CV_Assert failed: " #expr);    \
  } while (0)

typedef unsigned char uchar;  // global, as in opencv2/core/hal/interface.h

namespace cv {

struct Exception : std::runtime_error {
  using std::runtime_error::runtime_error;
};

enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum BorderTypes { BORDER_CONSTANT = 0, BORDER_REFLECT = 2 };

struct Size {
  int width = 0, height = 0;
  Size() = default;
  Size(int w, int h) : width(w), height(h) {}
};
struct Point {
  int x = 0, y = 0;
  Point() = default;
  Point(int x_, int y_) : x(x_), y(y_) {}
};
struct Point2f {
  float x = 0, y = 0;
  Point2f() = default;
  Point2f(float x_, float y_) : x(x_), y(y_) {}
};
struct Rect {
  int x = 0, y = 0, width = 0, height = 0;
  Rect() = default;
  Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
  Size size() const { return Size(width, height); }
};

class Mat {
 public:
  int rows = 0, cols = 0;
  unsigned char* data = nullptr;
  size_t step = 0;  // bytes per row
  Mat() = default;
  Mat(int r, int c, int type) { create(r, c, type); }
  // a view of caller-owned memory, like cv::Mat(rows, cols, type, data, step)
  Mat(int r, int c, int type, void* d, size_t s = 0)
      : rows(r), cols(c), data(static_cast<unsigned char*>(d)), type_(type) {
    step = s ? s : (size_t)c * elemSize();
  }
  void create(int r, int c, int type) {
    if (r == rows && c == cols && type == type_ && data) return;
    rows = r, cols = c, type_ = type;
    step = (size_t)c * elemSize();
    buf_ = std::shared_ptr<unsigned char>(new unsigned char[step * r + 1],
                                          std::default_delete<unsigned char[]>());
    data = buf_.get();
  }
  void create(Size s, int type) { create(s.height, s.width, type); }
  int type() const { return type_; }
  int depth() const { return type_ & 7; }
  int channels() const { return (type_ >> CV_CN_SHIFT) + 1; }
  size_t elemSize1() const { return depth() == CV_8U ? 1 : depth() == CV_16U ? 2 : 4; }
  size_t elemSize() const { return elemSize1() * channels(); }
  size_t step1() const { return step / elemSize1(); }
  size_t total() const { return (size_t)rows * cols; }
  Size size() const { return Size(cols, rows); }
  bool empty() const { return !data || !rows || !cols; }
  bool isContinuous() const { return step == (size_t)cols * elemSize(); }
  template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + step * r); }
  template <class T> const T* ptr(int r = 0) const {
    return reinterpret_cast<const T*>(data + step * r);
  }

 private:
  int type_ = 0;
  std::shared_ptr<unsigned char> buf_;
};

enum AccessFlag { ACCESS_READ = 1 << 24, ACCESS_WRITE = 1 << 25 };
// cv::UMat: for the shim's purposes a Mat whose host view is asked for
// explicitly (UMat::getMat(AccessFlag)).
class UMat : public Mat {
 public:
  UMat() = default;
  UMat(const Mat& m) : Mat(m) {}
  Mat getMat(AccessFlag) const { return *this; }
};

// InputArray / OutputArray / InputOutputArray: proxies of a Mat, passed by
// const reference in OpenCV signatures.
class _InputArray {
 public:
  _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
  Mat getMat() const { return *m_; }
  int type() const { return m_->type(); }
  size_t total() const { return m_->total(); }
  Size size() const { return m_->size(); }

 protected:
  Mat* m_;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray(Mat& m) : _InputArray(m) {}
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  void create(Size s, int type) const { m_->create(s, type); }
};
using _InputOutputArray = _OutputArray;
using InputArray = const _InputArray&;
using OutputArray = const _OutputArray&;
using InputOutputArray = const _InputOutputArray&;

template <class T> using Ptr = std::shared_ptr<T>;
template <class T, class... A> Ptr<T> makePtr(A&&... a) {
  return std::make_shared<T>(std::forward<A>(a)...);
}

}  // namespace cv
