// TEST INFRASTRUCTURE: the -DORB_B200_USE_OPENCV branch of cpp/ORBextractor.cc — the one a real ORB-SLAM2 build takes — exercised through
// cv::InputArray / cv::OutputArray (include/ORBextractor.h:122-126) against the OpenCV API model of oracle/cvshim (this image has no OpenCV C++).
// Checks operator() == CPU oracle on one frame, the empty-image rule (src/ORBextractor.cc:1966-1967) and mvImagePyramid.
//   test_dropin_cvapi frame.raw width height nfeatures
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ORBextractor.h"
#include "orb_oracle.h"

#define REQUIRE(cond, ...)                                              \
  do {                                                                  \
    if (!(cond)) {                                                      \
      std::fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__);         \
      std::fprintf(stderr, __VA_ARGS__);                                \
      std::fprintf(stderr, "\n");                                       \
      return 1;                                                         \
    }                                                                   \
  } while (0)

int main(int argc, char** argv) {
  if (argc < 5) { std::fprintf(stderr, "usage: %s frame.raw width height nfeatures\n", argv[0]); return 2; }
  const int W = std::atoi(argv[2]), H = std::atoi(argv[3]), nf = std::atoi(argv[4]);
  std::vector<unsigned char> f((size_t)W * H);
  FILE* fp = std::fopen(argv[1], "rb");
  REQUIRE(fp && std::fread(f.data(), 1, f.size(), fp) == f.size(), "cannot read %s", argv[1]);
  std::fclose(fp);

  ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
  cv::Mat img(H, W, CV_8UC1, f.data(), (size_t)W);
  std::vector<cv::KeyPoint> kps(3);   // stale content: operator() clears and refills (src/ORBextractor.cc:2016-2018)
  cv::Mat desc;
  ex(img, cv::Mat(), kps, desc);      // Frame::ExtractORB's call shape (src/Frame.cc:468-481): Mat -> InputArray, Mat -> OutputArray

  orb_oracle_extractor* oe = orb_oracle_create(nf, 1.2f, 8, 20, 7, 0);
  const int cap = nf + 1024;
  std::vector<orb_oracle_kp> ok(cap);
  std::vector<uint8_t> od((size_t)cap * 32);
  std::vector<int32_t> lc(8);
  const int n = orb_oracle_extract(oe, f.data(), W, W, H, ok.data(), od.data(), cap, lc.data());
  REQUIRE(n > 0 && (int)kps.size() == n, "keypoint count %zu vs oracle %d", kps.size(), n);
  REQUIRE(std::memcmp(kps.data(), ok.data(), (size_t)n * 28) == 0, "keypoints differ from the oracle");
  REQUIRE(desc.rows == n && desc.cols == 32 && desc.type() == CV_8U, "descriptor Mat shape %dx%d", desc.rows, desc.cols);
  for (int r = 0; r < n; ++r) REQUIRE(std::memcmp(desc.ptr<uchar>(r), &od[(size_t)r * 32], 32) == 0, "descriptor row %d differs", r);
  for (int l = 0; l < 8; ++l) {
    int c = 0, rws = 0;
    orb_oracle_level_dims(oe, l, &c, &rws);
    const cv::Mat& m = ex.mvImagePyramid[l];
    REQUIRE(m.cols == c && m.rows == rws, "mvImagePyramid[%d] size", l);
    const uint8_t* op = orb_oracle_level_ptr(oe, l);
    for (int y = 0; y < rws; ++y) REQUIRE(std::memcmp(m.ptr<uchar>(y), op + (size_t)y * c, c) == 0, "mvImagePyramid[%d] row %d", l, y);
  }
  // a second call reuses the registered pyramid staging (copy overlapped with the kernels): same answer
  ex(img, cv::Mat(), kps, desc);
  REQUIRE((int)kps.size() == n && std::memcmp(kps.data(), ok.data(), (size_t)n * 28) == 0, "second call differs");
  {
    const cv::Mat& m = ex.mvImagePyramid[7];
    int c = 0, rws = 0;
    orb_oracle_level_dims(oe, 7, &c, &rws);
    const uint8_t* op = orb_oracle_level_ptr(oe, 7);
    for (int y = 0; y < rws; ++y) REQUIRE(std::memcmp(m.ptr<uchar>(y), op + (size_t)y * c, c) == 0, "second call: mvImagePyramid[7] row %d", y);
  }
  // empty image: returns without touching the outputs (src/ORBextractor.cc:1966-1967)
  std::vector<cv::KeyPoint> keep = kps;
  ex(cv::Mat(), cv::Mat(), kps, desc);
  REQUIRE(kps.size() == keep.size() && desc.rows == n, "empty image must leave the outputs alone");
  orb_oracle_destroy(oe);
  std::printf("cvapi ok: %d keypoints through cv::InputArray / cv::OutputArray, bit-exact vs the oracle; mvImagePyramid equal\n", n);
  return 0;
}
