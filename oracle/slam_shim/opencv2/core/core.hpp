// TEST INFRASTRUCTURE.  Minimal stand-in for the slice of the OpenCV C++ API that the reference's
// src/ORBmatcher.cc uses (descriptor matrices of bytes, small float matrices for poses), so that file compiles
// UNMODIFIED here: OpenCV's C++ headers/libs are absent from the image and the reference does not vendor them.
// Our own code, not OpenCV's.  Float matrix products accumulate in double and round once to float, like
// OpenCV's generic GEMM for CV_32F; only the verbatim build sees these values -- the windows the reference
// derives from them are recorded (slam_shim.h) and handed to the oracle / GPU as inputs.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <sstream>   // OpenCV's core.hpp brings these in; DBoW2's TemplatedVocabulary.h relies on it
#include <stdexcept>
#include <string>
#include <vector>

typedef unsigned char uchar;
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

namespace cv {

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;

struct KeyPoint {  // 28 bytes, same field order as cv::KeyPoint
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

class Mat {
 public:
  int rows = 0, cols = 0;
  uchar* data = nullptr;
  size_t step = 0;

  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  void create(int r, int c, int type) {
    type_ = type;
    buf_ = std::make_shared<std::vector<uchar>>((size_t)r * c * esz(), 0);
    data = buf_->data();
    rows = r; cols = c; step = (size_t)c * esz();
  }
  void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  int type() const { return type_; }
  bool empty() const { return !data || !rows || !cols; }
  size_t esz() const { return type_ == CV_32F ? 4 : 1; }
  Mat clone() const {
    Mat m(rows, cols, type_);
    for (int y = 0; y < rows; ++y) std::memcpy(m.ptr(y), ptr(y), (size_t)cols * esz());
    return m;
  }
  // OpenCV asserts on ranges that leave the matrix; here they throw, so a caller can report the case
  Mat rowRange(int a, int b) const {
    if (a < 0 || b > rows || a > b) throw std::out_of_range("Mat::rowRange");
    Mat m(*this); m.data = data + (size_t)a * step; m.rows = b - a; return m;
  }
  Mat colRange(int a, int b) const {
    if (a < 0 || b > cols || a > b) throw std::out_of_range("Mat::colRange");
    Mat m(*this); m.data = data + (size_t)a * esz(); m.cols = b - a; return m;
  }
  static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }  // create() zero-fills
  static Mat ones(int r, int c, int type) {
    Mat m(r, c, type);
    for (int y = 0; y < r; ++y) for (int x = 0; x < c; ++x) m.at<float>(y, x) = 1.0f;
    return m;
  }
  void convertTo(Mat& dst, int type) const {  // CV_8U -> CV_32F (may be in place)
    assert(type == CV_32F);
    Mat m(rows, cols, CV_32F);
    for (int y = 0; y < rows; ++y)
      for (int x = 0; x < cols; ++x) m.at<float>(y, x) = type_ == CV_32F ? at<float>(y, x) : (float)at<uchar>(y, x);
    dst = m;
  }
  void copyTo(Mat& dst) const { dst = clone(); }
  Mat reshape(int /*cn*/) const { return *this; }  // channel views are only passed to undistortPoints (a no-op here)
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }
  template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
  template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
  // single index: element i of a row or column vector
  template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  uchar* ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  Mat t() const {
    Mat m(cols, rows, CV_32F);
    for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) m.at<float>(x, y) = at<float>(y, x);
    return m;
  }
  double dot(const Mat& b) const {  // CV_32F dot product accumulates in double
    double s = 0;
    for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) s += (double)at<float>(y, x) * b.at<float>(y, x);
    return s;
  }

 private:
  int type_ = CV_8U;
  std::shared_ptr<std::vector<uchar>> buf_;
};

inline Mat operator*(const Mat& a, const Mat& b) {
  assert(a.cols == b.rows);
  Mat m(a.rows, b.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < b.cols; ++x) {
      double s = 0;
      for (int k = 0; k < a.cols; ++k) s += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
      m.at<float>(y, x) = (float)s;
    }
  return m;
}
template <class F>
inline Mat mat_map2(const Mat& a, const Mat& b, F f) {
  assert(a.rows == b.rows && a.cols == b.cols);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = f(a.at<float>(y, x), b.at<float>(y, x));
  return m;
}
template <class F>
inline Mat mat_map1(const Mat& a, F f) {
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = f(a.at<float>(y, x));
  return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return mat_map2(a, b, [](float p, float q) { return p + q; }); }
inline Mat operator-(const Mat& a, const Mat& b) { return mat_map2(a, b, [](float p, float q) { return p - q; }); }
inline Mat operator-(const Mat& a) { return mat_map1(a, [](float p) { return -p; }); }
inline Mat operator*(double s, const Mat& a) { return mat_map1(a, [s](float p) { return (float)(p * s); }); }
inline Mat operator*(const Mat& a, double s) { return s * a; }
inline Mat operator/(const Mat& a, double s) { return mat_map1(a, [s](float p) { return (float)(p / s); }); }
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }
enum { NORM_L1 = 2 };
inline double norm(const Mat& a, const Mat& b, int type) {  // CV_32F: float differences, double accumulation
  assert(type == NORM_L1 && a.rows == b.rows && a.cols == b.cols);
  double s = 0;
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) s += std::fabs(a.at<float>(y, x) - b.at<float>(y, x));
  return s;
}
template <typename T>
class Mat_ : public Mat {  // only the `(Mat_<float>(3,1) << x, y, z)` initialiser of src/Frame.cc:675
 public:
  Mat_(int r, int c) : Mat(r, c, CV_32F) {}
  struct Init {
    Mat_* m; int i;
    Init& operator,(T v) { m->template at<T>(i / m->cols, i % m->cols) = v; ++i; return *this; }
    operator Mat() const { return *m; }
  };
  Init operator<<(T v) { Init in{this, 0}; in, v; return in; }
};
// cv::FileStorage / cv::FileNode: the YAML save / load members of DBoW2's TemplatedVocabulary are virtual, so their
// bodies must compile; nothing here opens a file (the tests load vocabularies with loadFromTextFile).
class FileNode {
 public:
  size_t size() const { return 0; }
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](int) const { return FileNode(); }
  operator int() const { return 0; }
  operator double() const { return 0; }
  operator std::string() const { return std::string(); }
};
class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage(const char*, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const std::string&) const { return FileNode(); }
  template <class T> FileStorage& operator<<(const T&) { return *this; }
};
// distortion-free cameras only (the callers test mDistCoef.at<float>(0) == 0 first and never get here)
inline void undistortPoints(const Mat&, Mat&, const Mat&, const Mat&, const Mat&, const Mat&) {
  throw std::logic_error("cv::undistortPoints is not part of the shim");
}

// cv::InputArray / cv::OutputArray as the extractor's operator() takes them (include/ORBextractor.h:59-61)
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
 private:
  Mat* m_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

}  // namespace cv
