// Drop-in for the hot-path part of the reference's include/ORBmatcher.h: DescriptorDistance (include/ORBmatcher.h:62,
// src/ORBmatcher.cc:2073-2093), the public thresholds (:223-225) and the best / second-best + ratio idiom every Search*
// function repeats (canonical form src/ORBmatcher.cc:304-342) as a brute-force k=2 call on the GPU.
// The nine guided Search* / Fuse functions are NOT part of this path and are not declared here; when
// integrating, keep the reference's own ORBmatcher.cc for them and add the members below.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cstdint>
#include <vector>

#ifdef ORB_B200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif

namespace ORB_SLAM2 {

class ORBmatcher {
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true);   // include/ORBmatcher.h:54

  // Hamming distance of two 256-bit descriptors (1 x 32 CV_8U rows). Scalar host form of the reference's signature: one
  // pair is below any offload granularity; every batched form below runs on the GPU only.
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);

  // DescriptorDistance over n row pairs on the GPU: out[i] = distance(a.row(i), b.row(i)).
  static void DescriptorDistanceBatch(const cv::Mat& a, const cv::Mat& b, std::vector<int>& out);

  struct Knn2 {
    int bestDist, bestDist2;    // == sentinel where unset (256; INT_MAX in SearchForInitialization, src/ORBmatcher.cc:611)
    int64_t bestIdx, bestIdx2;  // database row, -1 where unset
  };
  // For every query row: the scan `if (d < best) {second = best; best = d; idx = i} else if (d < second) second = d`
  // over all database rows in index order (strict <: the earliest index wins ties). GPU only.
  static void BruteForceKnn2(const cv::Mat& queries, const cv::Mat& database, std::vector<Knn2>& out, int sentinel = 256);

  // The accept test of src/ORBmatcher.cc:339-342: best <= th && best < mfNNratio * second (float32 product).
  bool Accept(const Knn2& m, int th) const { return m.bestDist <= th && (float)m.bestDist < mfNNratio * (float)m.bestDist2; }
  // Convenience: indices of accepted matches (-1 = rejected), threshold TH_LOW by default.
  int MatchBruteForce(const cv::Mat& queries, const cv::Mat& database, std::vector<int64_t>& matches, int th = TH_LOW) const;

  static const int TH_LOW;        // 50
  static const int TH_HIGH;       // 100
  static const int HISTO_LENGTH;  // 30

 protected:
  float mfNNratio;
  bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2
#endif
