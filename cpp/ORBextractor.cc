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

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST, int device, int flags)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST) {
  orb_params p;
  p.nfeatures = _nfeatures; p.scale_factor = _scaleFactor; p.nlevels = _nlevels;
  p.ini_th_fast = _iniThFAST; p.min_th_fast = _minThFAST; p.flags = flags;
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

ORBextractor::~ORBextractor() { orb_extractor_destroy(mpHandle); }   // before the Pinned members are freed: nothing is in flight any more

void ORBextractor::Pinned::ensure(size_t need) {
  if (need <= bytes) return;
  orb_host_free(p);
  p = nullptr; bytes = 0;
  check(orb_host_alloc(need, &p), "orb_host_alloc");
  bytes = need;
}
ORBextractor::Pinned::~Pinned() { orb_host_free(p); }

void ORBextractor::SetPyramidDownload(bool on) {
  mbDownloadPyramid = on;
  if (!on) {
    orb_extractor_set_pyramid_download(mpHandle, nullptr, nullptr);
    mPyrW = mPyrH = 0;
  }
}

// mvImagePyramid after a call on a W x H image. First call with this size: the level sizes are known only now, so the staging is allocated,
// the levels are fetched with one batched read and the destinations are registered; every later call finds them already filled by the
// library (copy overlapped with FAST / quadtree / descriptors of the same call). The Mats are headers over the object's pinned staging: like
// the reference's member they are overwritten by the next call (clone() what has to survive it).
void ORBextractor::FinishPyramid(int W, int H) {
  if (!mbDownloadPyramid || (W == mPyrW && H == mPyrH)) return;
  uint8_t* dst[64];
  int32_t steps[64];
  size_t total = 0;
  std::vector<int> cs(nlevels), rs(nlevels);
  for (int l = 0; l < nlevels; ++l) {
    check(orb_extractor_pyramid_dims(mpHandle, l, &cs[l], &rs[l]), "pyramid_dims");
    total += ((size_t)cs[l] * rs[l] + 63) / 64 * 64;
  }
  orb_extractor_set_pyramid_download(mpHandle, nullptr, nullptr);   // the old staging may be freed by ensure()
  mPyr.ensure(total);
  size_t off = 0;
  for (int l = 0; l < nlevels; ++l) {
    dst[l] = static_cast<uint8_t*>(mPyr.p) + off;
    steps[l] = cs[l];
    mvImagePyramid[l] = cv::Mat(rs[l], cs[l], CV_8UC1, dst[l], (size_t)cs[l]);
    off += ((size_t)cs[l] * rs[l] + 63) / 64 * 64;
  }
  check(orb_extractor_pyramid_read_all(mpHandle, -1, 0, dst, steps), "pyramid_read_all");
  check(orb_extractor_set_pyramid_download(mpHandle, dst, steps), "set_pyramid_download");
  mPyrW = W; mPyrH = H;
}

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
  mOutKps.ensure((size_t)B * cap * sizeof(orb_keypoint));
  mOutDesc.ensure((size_t)B * cap * 32);
  orb_keypoint* kps = static_cast<orb_keypoint*>(mOutKps.p);
  uint8_t* desc = static_cast<uint8_t*>(mOutDesc.p);
  std::vector<int32_t> counts(B);
  if (W != mPyrW || H != mPyrH) {   // another image size: the library drops the registration too; FinishPyramid makes the new one
    orb_extractor_set_pyramid_download(mpHandle, nullptr, nullptr);
    mPyrW = mPyrH = 0;
  }
  check(orb_extract_batch(mpHandle, ptrs.data(), ORB_MEM_HOST, use_step, W, H, B, kps, desc, ORB_MEM_HOST, cap, counts.data(), nullptr),
        "orb_extract_batch");
  for (int i = 0; i < B; ++i) {
    const int n = counts[i];
    keypoints[i].resize(n);
    if (n == 0) continue;   // descriptors stay released (src/ORBextractor.cc:1998-2013)
    std::memcpy((void*)keypoints[i].data(), &kps[(size_t)i * cap], (size_t)n * sizeof(orb_keypoint));
    descriptors[i].create(n, 32, CV_8U);
    if (descriptors[i].isContinuous()) std::memcpy(descriptors[i].ptr<unsigned char>(0), &desc[(size_t)i * cap * 32], (size_t)n * 32);
    else for (int r = 0; r < n; ++r) std::memcpy(descriptors[i].ptr<unsigned char>(r), &desc[((size_t)i * cap + r) * 32], 32);
  }
  FinishPyramid(W, H);
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
  mOutKps.ensure((size_t)2 * cap * sizeof(orb_keypoint));
  mOutDesc.ensure((size_t)2 * cap * 32);
  mOutUr.ensure((size_t)cap * sizeof(float));
  mOutDp.ensure((size_t)cap * sizeof(float));
  orb_keypoint* kps = static_cast<orb_keypoint*>(mOutKps.p);
  uint8_t* desc = static_cast<uint8_t*>(mOutDesc.p);
  float* ur = static_cast<float*>(mOutUr.p);
  float* dp = static_cast<float*>(mOutDp.p);
  int32_t counts[2] = {0, 0};
  if (W != mPyrW || H != mPyrH) {
    orb_extractor_set_pyramid_download(mpHandle, nullptr, nullptr);
    mPyrW = mPyrH = 0;
  }
  check(orb_extract_stereo_batch(mpHandle, ptrs, ORB_MEM_HOST, W, W, H, 1, kps, desc, ORB_MEM_HOST, cap, counts, nullptr, mbf, mb, ur, dp, nullptr),
        "orb_extract_stereo_batch");
  std::vector<cv::KeyPoint>* keys[2] = {&keysLeft, &keysRight};
  cv::Mat* descs[2] = {&descLeft, &descRight};
  for (int i = 0; i < 2; ++i) {
    const int n = counts[i];
    keys[i]->resize(n);
    if (n == 0) { descs[i]->release(); continue; }
    std::memcpy((void*)keys[i]->data(), &kps[(size_t)i * cap], (size_t)n * sizeof(orb_keypoint));
    descs[i]->create(n, 32, CV_8U);
    if (descs[i]->isContinuous()) std::memcpy(descs[i]->ptr<unsigned char>(0), &desc[(size_t)i * cap * 32], (size_t)n * 32);
    else for (int r = 0; r < n; ++r) std::memcpy(descs[i]->ptr<unsigned char>(r), &desc[((size_t)i * cap + r) * 32], 32);
  }
  uRight.assign(ur, ur + counts[0]);
  depth.assign(dp, dp + counts[0]);
  FinishPyramid(W, H);
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
