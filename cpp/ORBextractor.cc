// ORB_SLAM2::ORBextractor over the C ABI (include/orb_b200.h). Host glue only: argument checks with the reference's
// semantics, cv::Mat <-> plain pointers, std::vector<cv::KeyPoint> filled by memcpy (orb_keypoint == cv::KeyPoint layout).
#include "ORBextractor.h"

#include <cstring>
#include <stdexcept>
#include <string>

#include "../include/orb_b200.h"

namespace ORB_SLAM2 {

static_assert(sizeof(cv::KeyPoint) == sizeof(orb_keypoint), "cv::KeyPoint must be the 28-byte POD the C ABI writes");

static void check(int rc, const char* what) {
  if (rc < 0) throw std::runtime_error(std::string(what) + ": " + orb_last_error());
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST, int device)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST) {
  orb_params p;
  p.nfeatures = _nfeatures; p.scale_factor = _scaleFactor; p.nlevels = _nlevels;
  p.ini_th_fast = _iniThFAST; p.min_th_fast = _minThFAST; p.flags = 0;
  check(orb_extractor_create(&p, device, &mpHandle), "orb_extractor_create");
  // the tables are computed once by the library with the reference's mixed float/double arithmetic (src/ORBextractor.cc:776-845)
  mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
  mnFeaturesPerLevel.resize(nlevels);
  check(orb_extractor_get_table(mpHandle, 0, mvScaleFactor.data()), "get_table");
  check(orb_extractor_get_table(mpHandle, 1, mvInvScaleFactor.data()), "get_table");
  check(orb_extractor_get_table(mpHandle, 2, mvLevelSigma2.data()), "get_table");
  check(orb_extractor_get_table(mpHandle, 3, mvInvLevelSigma2.data()), "get_table");
  check(orb_extractor_get_features_per_level(mpHandle, mnFeaturesPerLevel.data()), "get_features_per_level");
  mvImagePyramid.resize(nlevels);
}

ORBextractor::~ORBextractor() { orb_extractor_destroy(mpHandle); }

void ORBextractor::ExtractBatch(const std::vector<cv::Mat>& images, std::vector<std::vector<cv::KeyPoint>>& keypoints,
                                std::vector<cv::Mat>& descriptors) {
  const int B = (int)images.size();
  keypoints.assign(B, {});
  descriptors.assign(B, cv::Mat());
  if (B == 0) return;
  const int W = images[0].cols, H = images[0].rows;
  const int step = (int)images[0].step;
  bool same_step = true;
  for (const cv::Mat& m : images) {
    if (m.empty() || m.type() != CV_8UC1 || m.cols != W || m.rows != H)
      throw std::runtime_error("ORBextractor::ExtractBatch: images must be non-empty CV_8UC1 of one size");
    same_step = same_step && (int)m.step == step;
  }
  std::vector<const uint8_t*> ptrs(B);
  int use_step = step;
  if (same_step) {
    for (int i = 0; i < B; ++i) ptrs[i] = images[i].data;
  } else {   // mixed pitches: repack once
    use_step = W;
    mStage.resize((size_t)B * W * H);
    for (int i = 0; i < B; ++i) {
      for (int r = 0; r < H; ++r) std::memcpy(&mStage[((size_t)i * H + r) * W], images[i].ptr<unsigned char>(r), W);
      ptrs[i] = &mStage[(size_t)i * W * H];
    }
  }
  const int cap = orb_extractor_max_keypoints(mpHandle);
  std::vector<orb_keypoint> kps((size_t)B * cap);
  std::vector<uint8_t> desc((size_t)B * cap * 32);
  std::vector<int32_t> counts(B);
  check(orb_extract_batch(mpHandle, ptrs.data(), ORB_MEM_HOST, use_step, W, H, B, kps.data(), desc.data(), ORB_MEM_HOST, cap,
                          counts.data(), nullptr),
        "orb_extract_batch");
  for (int i = 0; i < B; ++i) {
    const int n = counts[i];
    keypoints[i].resize(n);
    if (n == 0) continue;   // descriptors stay released (src/ORBextractor.cc:1998-2013)
    std::memcpy((void*)keypoints[i].data(), &kps[(size_t)i * cap], (size_t)n * sizeof(orb_keypoint));
    descriptors[i].create(n, 32, CV_8U);
    for (int r = 0; r < n; ++r) std::memcpy(descriptors[i].ptr<unsigned char>(r), &desc[((size_t)i * cap + r) * 32], 32);
  }
  if (mbDownloadPyramid) {
    for (int l = 0; l < nlevels; ++l) {
      int c = 0, r = 0;
      check(orb_extractor_pyramid_dims(mpHandle, l, &c, &r), "pyramid_dims");
      mvImagePyramid[l].create(r, c, CV_8UC1);
      // the library keeps the frames of the last internal pass resident; the last image of the call is always among them
      check(orb_extractor_pyramid_read(mpHandle, -1, l, 0, mvImagePyramid[l].ptr<unsigned char>(0), (int)mvImagePyramid[l].step),
            "pyramid_read");
    }
  }
}

void ORBextractor::ExtractStereo(const cv::Mat& left, const cv::Mat& right, float mbf, float mb, std::vector<cv::KeyPoint>& keysLeft,
                                 cv::Mat& descLeft, std::vector<cv::KeyPoint>& keysRight, cv::Mat& descRight,
                                 std::vector<float>& uRight, std::vector<float>& depth) {
  if (left.empty() || right.empty() || left.type() != CV_8UC1 || right.type() != CV_8UC1 || left.cols != right.cols || left.rows != right.rows)
    throw std::runtime_error("ORBextractor::ExtractStereo: two non-empty CV_8UC1 images of one size required");
  const int W = left.cols, H = left.rows;
  mStage.resize((size_t)2 * W * H);
  for (int r = 0; r < H; ++r) {
    std::memcpy(&mStage[(size_t)r * W], left.ptr<unsigned char>(r), W);
    std::memcpy(&mStage[((size_t)H + r) * W], right.ptr<unsigned char>(r), W);
  }
  const uint8_t* ptrs[2] = {mStage.data(), mStage.data() + (size_t)W * H};
  const int cap = orb_extractor_max_keypoints(mpHandle);
  std::vector<orb_keypoint> kps((size_t)2 * cap);
  std::vector<uint8_t> desc((size_t)2 * cap * 32);
  std::vector<float> ur(cap), dp(cap);
  int32_t counts[2] = {0, 0};
  check(orb_extract_stereo_batch(mpHandle, ptrs, ORB_MEM_HOST, W, W, H, 1, kps.data(), desc.data(), ORB_MEM_HOST, cap, counts, nullptr, mbf, mb,
                                 ur.data(), dp.data(), nullptr),
        "orb_extract_stereo_batch");
  std::vector<cv::KeyPoint>* keys[2] = {&keysLeft, &keysRight};
  cv::Mat* descs[2] = {&descLeft, &descRight};
  for (int i = 0; i < 2; ++i) {
    const int n = counts[i];
    keys[i]->resize(n);
    if (n == 0) { descs[i]->release(); continue; }
    std::memcpy((void*)keys[i]->data(), &kps[(size_t)i * cap], (size_t)n * sizeof(orb_keypoint));
    descs[i]->create(n, 32, CV_8U);
    for (int r = 0; r < n; ++r) std::memcpy(descs[i]->ptr<unsigned char>(r), &desc[((size_t)i * cap + r) * 32], 32);
  }
  uRight.assign(ur.begin(), ur.begin() + counts[0]);
  depth.assign(dp.begin(), dp.begin() + counts[0]);
  if (mbDownloadPyramid) {
    for (int l = 0; l < nlevels; ++l) {
      int cc = 0, rr = 0;
      check(orb_extractor_pyramid_dims(mpHandle, l, &cc, &rr), "pyramid_dims");
      mvImagePyramid[l].create(rr, cc, CV_8UC1);
      check(orb_extractor_pyramid_read(mpHandle, -1, l, 0, mvImagePyramid[l].ptr<unsigned char>(0), (int)mvImagePyramid[l].step), "pyramid_read");
    }
  }
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors) {
#ifdef ORB_B200_USE_OPENCV
  if (_image.empty()) return;
  cv::Mat image = _image.getMat();
#else
  if (_image.empty()) return;
  const cv::Mat& image = _image;
#endif
  if (image.type() != CV_8UC1) throw std::runtime_error("ORBextractor: image must be CV_8UC1 (src/ORBextractor.cc:1972)");
  std::vector<cv::Mat> in(1, image);
  std::vector<std::vector<cv::KeyPoint>> k;
  std::vector<cv::Mat> d;
  ExtractBatch(in, k, d);
  _keypoints.swap(k[0]);
#ifdef ORB_B200_USE_OPENCV
  if (d[0].empty()) _descriptors.release();
  else d[0].copyTo(_descriptors);
#else
  if (d[0].empty()) _descriptors.release();
  else _descriptors = d[0];
#endif
}

}  // namespace ORB_SLAM2
