// Minimal stand-in for the handful of OpenCV types that cross the ORBextractor / ORBmatcher boundary
// (cv::Mat as an 8-bit matrix, cv::KeyPoint, cv::Point2f, cv::InputArray / cv::OutputArray).
//
// Only used when the drop-in classes are built WITHOUT OpenCV (this image has no OpenCV C++ headers; the C++ parity test
// tests/test_cpp_dropin.py builds against this file). With OpenCV present compile with -DORB_B200_USE_OPENCV and the real
// headers are used instead; the class sources are written against the common subset:
//   Mat: empty(), type(), rows, cols, step (bytes), data, ptr<T>(row), create(rows, cols, type), release(), row(i)
//   KeyPoint: 28-byte POD {pt.x, pt.y, size, angle, response, octave, class_id}
#ifndef ORB_B200_CV_SHIM_H_
#define ORB_B200_CV_SHIM_H_
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32S 4

namespace cv {

struct Point2f {
  float x = 0.f, y = 0.f;
};

struct KeyPoint {
  Point2f pt;
  float size = 0.f;
  float angle = -1.f;
  float response = 0.f;
  int octave = 0;
  int class_id = -1;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;          // bytes per row
  unsigned char* data = nullptr;

  Mat() = default;
  Mat(int r, int c, int type) { create(r, c, type); }
  // non-owning view over caller memory (like cv::Mat(rows, cols, type, data, step))
  Mat(int r, int c, int type, void* d, size_t s = 0) : rows(r), cols(c), step(s ? s : (size_t)c * elem(type)), data((unsigned char*)d), type_(type) {}

  void create(int r, int c, int type) {
    if (r == rows && c == cols && type == type_ && store_ && data == store_->data()) return;
    type_ = type;
    rows = r; cols = c;
    step = (size_t)c * elem(type);
    store_ = std::make_shared<std::vector<unsigned char>>((size_t)r * step);
    data = store_->data();
  }
  void release() { store_.reset(); data = nullptr; rows = cols = 0; step = 0; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  bool isContinuous() const { return step == (size_t)cols * elem(type_); }
  template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
  template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
  Mat row(int r) const {
    Mat m(1, cols, type_, data + (size_t)r * step, step);
    m.store_ = store_;
    return m;
  }
  Mat clone() const {
    Mat m(rows, cols, type_);
    for (int r = 0; r < rows; ++r) std::memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols * elem(type_));
    return m;
  }

 private:
  static size_t elem(int type) { return type == CV_32S ? 4 : 1; }
  int type_ = CV_8UC1;
  std::shared_ptr<std::vector<unsigned char>> store_;
};

// The reference passes cv::Mat for both (src/Frame.cc:468-481); the proxies collapse to references here.
typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#endif  // ORB_B200_CV_SHIM_H_
