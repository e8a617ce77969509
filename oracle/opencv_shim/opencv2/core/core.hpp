// TEST INFRASTRUCTURE.  Minimal stand-in for the slice of the OpenCV C++ API that the
// reference's src/ORBextractor.cc + include/ORBextractor.h use, so that those two files compile
// UNMODIFIED here (OpenCV C++ headers/libs are absent from the image; the reference does not
// vendor OpenCV).  The image-processing entry points forward to oracle/cvprim.h, whose
// arithmetic is pinned bit-for-bit to cv2 4.13.0.  This is our own code, not OpenCV's.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iterator>
#include <list>
#include <memory>
#include <vector>

typedef unsigned char uchar;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

int cvRound(double v);
int cvFloor(double v);
int cvCeil(double v);

namespace cv {

enum { INTER_LINEAR = 1 };
enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point2i Point;
template <typename T>
inline Point_<T>& operator*=(Point_<T>& a, float b) {
  a.x = (T)(a.x * b);
  a.y = (T)(a.y * b);
  return a;
}

struct Size {
  int width, height;
  Size() : width(0), height(0) {}
  Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
  int x, y, width, height;
  Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

struct KeyPoint {  // 28 bytes, same field order as cv::KeyPoint
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0,
           int class_id_ = -1)
      : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_),
        class_id(class_id_) {}
};

struct MatZeros { int rows, cols, type; };

class Mat {
 public:
  struct Step {
    size_t v = 0;
    operator size_t() const { return v; }
  };
  int rows = 0, cols = 0;
  uchar* data = nullptr;
  Step step;

  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(Size s, int type) { create(s.height, s.width, type); }
  // wraps caller memory without owning it
  Mat(int r, int c, int /*type*/, void* ext, size_t stride) : rows(r), cols(c), data((uchar*)ext) {
    step.v = stride;
  }
  static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }

  void create(int r, int c, int /*type*/) {
    if (data && r == rows && c == cols) return;
    buf_ = std::make_shared<std::vector<uchar>>((size_t)r * c);
    data = buf_->data();
    rows = r; cols = c; step.v = (size_t)c;
  }
  void release() { buf_.reset(); data = nullptr; rows = cols = 0; step.v = 0; }
  Mat& operator=(const MatZeros& z) {  // like MatExpr assignment: in place when the shape matches
    create(z.rows, z.cols, z.type);
    for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step.v, 0, (size_t)cols);
    return *this;
  }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return CV_8UC1; }
  size_t step1() const { return step.v; }
  Mat clone() const {
    Mat m(rows, cols, CV_8UC1);
    for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step.v, data + (size_t)y * step.v, (size_t)cols);
    return m;
  }
  Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step.v; m.rows = b - a; return m; }
  Mat colRange(int a, int b) const { Mat m(*this); m.data = data + a; m.cols = b - a; return m; }
  Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
  template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step.v + (size_t)x * sizeof(T)); }
  template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step.v + (size_t)x * sizeof(T)); }
  uchar* ptr(int y = 0) { return data + (size_t)y * step.v; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step.v; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.v); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.v); }

 private:
  std::shared_ptr<std::vector<uchar>> buf_;
};

class _InputArray {
 public:
  _InputArray(const Mat& m) : m_(&m) {}
  Mat getMat() const { return *m_; }
  bool empty() const { return m_->empty(); }
 private:
  const Mat* m_;
};
class _OutputArray {
 public:
  _OutputArray(Mat& m) : m_(&m) {}
  Mat getMat() const { return *m_; }
  void release() const { m_->release(); }
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  Mat& ref() const { return *m_; }
 private:
  Mat* m_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0,
            int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right,
                    int borderType);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_REFLECT_101);
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold,
          bool nonmaxSuppression = true);
float fastAtan2(float y, float x);

}  // namespace cv
