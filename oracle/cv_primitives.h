// TEST INFRASTRUCTURE ONLY — CPU oracle. Nothing under orb_slam2_detailed_comments_b200/ may
// include, link or call this. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs use it, as the checker / reported baseline.
//
// Restatement of the five OpenCV primitives the reference's hot path calls (OpenCV is an
// un-vendored dependency of /root/reference: CMakeLists.txt:49-55, so its published 8-bit
// fixed-point algorithms are restated here and pinned against cv2 4.13 in tests/test_cv_primitives.py):
//   cv::resize INTER_LINEAR 8UC1        (called at src/ORBextractor.cc:2119)
//   cv::copyMakeBorder REFLECT_101      (src/ORBextractor.cc:2138, 2159)
//   cv::GaussianBlur 7x7 sigma 2        (src/ORBextractor.cc:2042)
//   cv::FAST(th, nonmax=true) TYPE_9_16 (src/ORBextractor.cc:1510, 1521)
//   cv::fastAtan2                       (src/ORBextractor.cc:160)
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orb_oracle {

struct Img {  // non-owning 8-bit single-channel view
  uint8_t* data = nullptr;
  int cols = 0, rows = 0, step = 0;
  uint8_t* ptr(int y) const { return data + (size_t)y * step; }
  Img roi(int x, int y, int w, int h) const { return Img{data + (size_t)y * step + x, w, h, step}; }
};

// cv::KeyPoint memory layout (7 x 4 bytes): pt.x, pt.y, size, angle, response, octave, class_id
struct KeyPoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

int cv_round(float v);   // cvRound: round-half-to-even
int cv_round(double v);

void resize_linear_u8(const Img& src, const Img& dst);
// dst is (src.cols+2*b) x (src.rows+2*b); interior copied, border reflect-101 of the interior.
void copy_make_border_reflect101(const Img& src, const Img& dst, int b);
// legacy_taps=false: OpenCV >=3.4.1/4.x bit-exact path [18,34,48,56,48,34,18]; true: [18,34,49,55,49,34,18]
void gaussian_blur_7x7_s2(const Img& src, const Img& dst, bool legacy_taps = false);
// Threshold-free FAST-9/16 score S=max(A,B)-1 clamped to [0,255] for one pixel (needs 3 px margin).
int fast_corner_score(const uint8_t* p, int step);
// cv::FAST(img, kps, threshold, nonmax=true): row-major emission, pt relative to img.
void fast_detect(const Img& img, int threshold, bool nonmax, std::vector<KeyPoint>& out);
float fast_atan2(float y, float x);

}  // namespace orb_oracle
