// TEST INFRASTRUCTURE ONLY — used by oracle/Makefile's `_ref` target, never by the product.
//
// The subset of the OpenCV C++ API that /root/reference/src/ORBextractor.cc and Thirdparty/DBoW2 touch, so that the
// reference's OWN sources compile here unmodified (this image has no OpenCV C++ headers). Containers (Mat with shared
// ownership and ROI views, KeyPoint, Point_, Size, Rect, Input/OutputArray) are plain re-implementations of the
// documented semantics; the five arithmetic primitives (resize INTER_LINEAR, copyMakeBorder REFLECT_101, GaussianBlur
// 7x7 s2, FAST-9/16 + NMS, fastAtan2) forward to oracle/cv_primitives.cpp, which tests/test_oracle_golden.py pins
// bit-exact against the real cv2 4.13. What the resulting binary proves: the oracle's restatement of the reference's
// control flow (cell loop, DistributeOctTree, orientation, descriptors, level bookkeeping) equals the reference's code.
#ifndef ORB_ORACLE_CVSHIM_H_
#define ORB_ORACLE_CVSHIM_H_
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "../cv_primitives.h"

typedef unsigned char uchar;
#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5

inline int cvRound(double v) { return orb_oracle::cv_round(v); }
inline int cvRound(float v) { return orb_oracle::cv_round(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }

namespace cv {

template <typename T> struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
  template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
  Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
  int width, height;
  Size() : width(0), height(0) {}
  Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
  int x, y, width, height;
  Rect() : x(0), y(0), width(0), height(0) {}
  Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
      : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };

class Mat {
 public:
  int rows, cols;
  size_t step;   // bytes per row (cv::MatStep converts to size_t)
  uchar* data;

  Mat() : rows(0), cols(0), step(0), data(nullptr), type_(0) {}
  Mat(int r, int c, int type) : Mat() { create(r, c, type); }
  Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, void* d, size_t s = 0) : rows(r), cols(c), step(s ? s : (size_t)c * esz(type)), data((uchar*)d), type_(type) {}

  void create(int r, int c, int type) {   // cv::Mat::create: no reallocation when shape and type already match
    if (data && r == rows && c == cols && type == type_) return;
    type_ = type; rows = r; cols = c;
    step = (size_t)c * esz(type);
    store_ = std::shared_ptr<uchar>((uchar*)std::malloc((size_t)r * step + 1), std::free);
    data = store_.get();
  }
  void create(Size sz, int type) { create(sz.height, sz.width, type); }
  void release() { store_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  // Mat::zeros is a MatExpr in OpenCV: assigning it to an existing Mat of the same shape and type fills that Mat IN PLACE
  // (create() keeps the buffer) — computeDescriptors relies on it to write into the caller's descriptor rows (src/ORBextractor.cc:1941)
  struct ZerosExpr { int r, c, type; };
  static ZerosExpr zeros(int r, int c, int type) { return ZerosExpr{r, c, type}; }
  Mat(const ZerosExpr& e) : Mat() { *this = e; }
  Mat& operator=(const ZerosExpr& e) {
    create(e.r, e.c, e.type);
    for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step, 0, (size_t)cols * esz(type_));
    return *this;
  }
  Mat clone() const {
    Mat m(rows, cols, type_);
    for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * esz(type_));
    return m;
  }
  Mat operator()(const Rect& r) const { Mat m(*this); m.data = data + (size_t)r.y * step + (size_t)r.x * esz(type_); m.rows = r.height; m.cols = r.width; return m; }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }
  Mat t() const;   // 32F transpose (defined below)
  double dot(const Mat& m) const {   // cv::Mat::dot for 32F: double accumulation
    double acc = 0;
    for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) acc += (double)at<float>(y, x) * (double)m.at<float>(y, x);
    return acc;
  }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  int depth() const { return type_; }
  int channels() const { return 1; }
  Size size() const { return Size(cols, rows); }
  size_t step1() const { return step / esz(type_); }
  size_t elemSize() const { return esz(type_); }
  bool isContinuous() const { return step == (size_t)cols * esz(type_) || rows <= 1; }
  template <typename T> T* ptr(int y = 0) { return reinterpret_cast<T*>(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return reinterpret_cast<const T*>(data + (size_t)y * step); }
  uchar* ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
  template <typename T> T& at(int y, int x) { return reinterpret_cast<T*>(data + (size_t)y * step)[x]; }
  template <typename T> const T& at(int y, int x) const { return reinterpret_cast<const T*>(data + (size_t)y * step)[x]; }
  // single index on a vector (cv::Mat::at(int i0)): element i of a row vector, row i of a column vector (views keep their row stride)
  template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }

  void copyTo(const class _OutputArray& o) const;   // defined after _OutputArray
  orb_oracle::Img img() const { return orb_oracle::Img{data, cols, rows, (int)step}; }

  // ---- float patches of Frame::ComputeStereoMatches (src/Frame.cc:1035-1068): convertTo(CV_32F), Mat::ones, scalar * Mat, Mat - Mat ----
  void convertTo(Mat& m, int rtype) const {   // 8U -> 32F (or a same-type copy); `m` may be *this (a view): the new buffer replaces it
    Mat out(rows, cols, rtype);
    for (int y = 0; y < rows; ++y)
      for (int x = 0; x < cols; ++x) {
        const float v = type_ == CV_32F ? at<float>(y, x) : (float)at<uchar>(y, x);
        if (rtype == CV_32F) out.at<float>(y, x) = v; else out.at<uchar>(y, x) = (uchar)v;
      }
    m = out;
  }
  static Mat ones(int r, int c, int type) {
    assert(type == CV_32F);
    Mat m(r, c, type);
    for (int y = 0; y < r; ++y) for (int x = 0; x < c; ++x) m.at<float>(y, x) = 1.f;
    return m;
  }

 private:
  static size_t esz(int type) { return type == CV_32F ? 4 : 1; }
  std::shared_ptr<uchar> store_;
  int type_;
};

inline Mat operator*(float k, const Mat& a) {   // MatExpr scale: float32 product per element
  assert(a.type() == CV_32F);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = k * a.at<float>(y, x);
  return m;
}
inline Mat operator-(const Mat& a, const Mat& b) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
  return m;
}
// ---- the pose algebra at the head of the projection matchers (src/ORBmatcher.cc:1720-1733, 1753): small 32F matrices. cv::gemm accumulates
// float products in double and rounds once; the pins feed identity rotations and power-of-two depths, for which every product and sum is exact
// in any order, so this restatement cannot differ from OpenCV on them.
inline Mat Mat::t() const {
  assert(type_ == CV_32F);
  Mat m(cols, rows, CV_32F);
  for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) m.at<float>(x, y) = at<float>(y, x);
  return m;
}
inline Mat operator-(const Mat& a) { return -1.f * a; }
inline Mat operator/(const Mat& a, float k) {   // MatExpr scale by 1 / k: cv computes a * (1.0 / k) in double and rounds once; exact for k = 1 (the pins)
  assert(a.type() == CV_32F);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = (float)((double)a.at<float>(y, x) * (1.0 / (double)k));
  return m;
}
inline Mat operator+(const Mat& a, const Mat& b) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
  return m;
}
inline Mat operator*(const Mat& a, const Mat& b) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
  Mat m(a.rows, b.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < b.cols; ++x) {
      double acc = 0;
      for (int k = 0; k < a.cols; ++k) acc += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
      m.at<float>(y, x) = (float)acc;
    }
  return m;
}
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
inline double norm(const Mat& a) {   // cv::norm(a) = L2 for 32F, accumulated in double
  double acc = 0;
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) acc += (double)a.at<float>(y, x) * (double)a.at<float>(y, x);
  return std::sqrt(acc);
}
inline double norm(const Mat& a, const Mat& b, int normType) {   // cv::norm(a, b, NORM_L1) for 32F: sum |a - b| accumulated in double
  assert(normType == NORM_L1 && a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
  (void)normType;
  double s = 0;
  for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) s += std::fabs((double)a.at<float>(y, x) - (double)b.at<float>(y, x));
  return s;
}

class _InputArray {
 public:
  _InputArray() : m_(nullptr) {}
  _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
  Mat getMat() const { return m_ ? *m_ : Mat(); }
  bool empty() const { return !m_ || m_->empty(); }
 protected:
  Mat* m_;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray() {}
  _OutputArray(Mat& m) : _InputArray(m) {}
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  void create(Size sz, int type) const { m_->create(sz, type); }
  void release() const { if (m_) m_->release(); }
};
inline void Mat::copyTo(const _OutputArray& o) const {   // cv::Mat::copyTo(OutputArray): (re)allocates the destination, then copies row by row
  o.create(rows, cols, type_);
  Mat d = o.getMat();
  for (int y = 0; y < rows; ++y) std::memcpy(d.data + (size_t)y * d.step, data + (size_t)y * step, (size_t)cols * esz(type_));
}
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline float fastAtan2(float y, float x) { return orb_oracle::fast_atan2(y, x); }

// cv::FAST(image, keypoints, threshold, nonmaxSuppression): TYPE_9_16
inline void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmax = true) {
  const Mat m = image.getMat();
  std::vector<orb_oracle::KeyPoint> kps;
  orb_oracle::fast_detect(m.img(), threshold, nonmax, kps);
  keypoints.clear();
  keypoints.reserve(kps.size());
  for (const auto& k : kps) keypoints.push_back(KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id));
}

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR); dst.create(dsize) keeps an already matching ROI (the reference relies on it)
inline void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR) {
  (void)fx; (void)fy;
  assert(interpolation == INTER_LINEAR);
  const Mat s = src.getMat();
  dst.create(dsize, s.type());
  const Mat d = dst.getMat();
  orb_oracle::resize_linear_u8(s.img(), d.img());
}

// cv::copyMakeBorder with BORDER_REFLECT_101 (+ BORDER_ISOLATED); in place when src is the interior ROI of dst
inline void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType) {
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && top == left && top == right);
  (void)bottom; (void)left; (void)right;
  const Mat s = src.getMat();
  dst.create(s.rows + 2 * top, s.cols + 2 * top, s.type());
  const Mat d = dst.getMat();
  orb_oracle::copy_make_border_reflect101(s.img(), d.img(), top);
}

inline void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sx, double sy = 0, int borderType = BORDER_REFLECT_101) {
  assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 && borderType == BORDER_REFLECT_101);
  (void)ksize; (void)sx; (void)sy; (void)borderType;
  const Mat s = src.getMat();
  dst.create(s.rows, s.cols, s.type());
  const Mat d = dst.getMat();
  orb_oracle::gaussian_blur_7x7_s2(s.img(), d.img(), std::getenv("ORB_REF_LEGACY_BLUR") != nullptr);
}

// only referenced from ComputeKeyPointsOld (dead code in the reference: its call is commented out at src/ORBextractor.cc:1987)
struct KeyPointsFilter {
  static void retainBest(std::vector<KeyPoint>&, int) { std::abort(); }
};

// cv::FileStorage: only referenced by DBoW2's yml save/load, which the ORB-SLAM2 text vocabulary path never calls
struct FileNode {
  FileNode operator[](const char*) const { std::abort(); }
  FileNode operator[](int) const { std::abort(); }
  size_t size() const { return 0; }
  operator int() const { std::abort(); }
  operator double() const { std::abort(); }
  operator std::string() const { std::abort(); }
};
struct FileStorage {
  enum { READ = 0, WRITE = 1 };
  FileStorage() {}
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  void release() {}
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
  template <typename T> FileStorage& operator<<(const T&) { return *this; }
};

}  // namespace cv
#endif
