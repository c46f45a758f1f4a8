// Drop-in replacement for the reference's include/ORBextractor.h: same class name, namespace, constructor, operator(),
// getters and public mvImagePyramid, so Frame / Tracking compile and link against it unchanged
// (callers: src/Frame.cc:142-154, 468-481, 910, 1040; src/Tracking.cc:171-182).
// All work happens in liborb_b200.so (include/orb_b200.h) on a B200; there is no CPU implementation behind this class.
//
// Additions (not in the reference): the batched call ExtractBatch() (the stereo Frame ctor extracts left and right on two
// threads, src/Frame.cc:158-168; one batched call does both in one pass) and SetPyramidDownload() to skip the host copy
// of mvImagePyramid when the caller is monocular (only Frame::ComputeStereoMatches reads it).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <vector>

#ifdef ORB_B200_USE_OPENCV
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#else
#include "cv_shim.h"
#endif

struct orb_extractor;   // C ABI handle, include/orb_b200.h

namespace ORB_SLAM2 {

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };   // include/ORBextractor.h:87-91 (unused by the reference as well)

  // include/ORBextractor.h:104. `device` selects the GPU (default: current CUDA device 0).
  // `flags`: orb_flags of include/orb_b200.h (1 = raulmur's un-padded pyramid instead of this fork's padded one, 2 = OpenCV <= 3.4.0 blur taps).
  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0, int flags = 0);
  ~ORBextractor();
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // include/ORBextractor.h:122-126. Mask is ignored (as in the reference, :112). An empty image returns without touching
  // the outputs (src/ORBextractor.cc:1966-1967). Throws std::runtime_error on a CUDA / capacity error (the reference
  // asserts on a non-CV_8UC1 image, :1972; here that is an exception too).
  void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);

  // Batched form: all images must have the same size and type CV_8UC1.
  void ExtractBatch(const std::vector<cv::Mat>& images, std::vector<std::vector<cv::KeyPoint>>& keypoints,
                    std::vector<cv::Mat>& descriptors);

  // Stereo Frame constructor in one call (src/Frame.cc:158-168 + Frame::ComputeStereoMatches, src/Frame.cc:880-1176): extracts both
  // images and fills mvuRight / mvDepth of the left keypoints on the device. mbf = baseline * fx, mb = baseline.
  // mvImagePyramid afterwards holds the RIGHT image's levels (the last image of the call).
  void ExtractStereo(const cv::Mat& left, const cv::Mat& right, float mbf, float mb, std::vector<cv::KeyPoint>& keysLeft,
                     cv::Mat& descLeft, std::vector<cv::KeyPoint>& keysRight, cv::Mat& descRight, std::vector<float>& uRight,
                     std::vector<float>& depth);

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  // Padded CV_8UC1 levels of the last image of the last call (this fork stores the 19-px padded image,
  // src/ORBextractor.cc:2159-2166). Filled after every call unless SetPyramidDownload(false).
  std::vector<cv::Mat> mvImagePyramid;
  void SetPyramidDownload(bool on);

  orb_extractor* Handle() { return mpHandle; }

 protected:
  int nfeatures;
  double scaleFactor;   // double, like the reference member (include/ORBextractor.h:234)
  int nlevels;
  int iniThFAST;
  int minThFAST;
  std::vector<int> mnFeaturesPerLevel;
  std::vector<float> mvScaleFactor;
  std::vector<float> mvInvScaleFactor;
  std::vector<float> mvLevelSigma2;
  std::vector<float> mvInvLevelSigma2;

 private:
  orb_extractor* mpHandle = nullptr;
  bool mbDownloadPyramid = true;
  std::vector<unsigned char> mStage;   // contiguous copy of non-contiguous / differently strided batches
  // Page-locked staging owned by the object (no allocation per call): keypoints / descriptors / stereo columns of a whole batch, and the
  // padded pyramid levels that mvImagePyramid's headers point into (the library copies them home while the rest of the call still runs).
  struct Pinned {
    void* p = nullptr;
    size_t bytes = 0;
    void ensure(size_t need);
    ~Pinned();
  };
  Pinned mOutKps, mOutDesc, mOutUr, mOutDp, mPyr;
  int mPyrW = 0, mPyrH = 0;            // image size the pyramid staging / registration was made for
  void FinishPyramid(int W, int H);    // after a call: mvImagePyramid = headers over mPyr, registered for the following calls
};

}  // namespace ORB_SLAM2
#endif
