// Drop-in for the hot-path part of the reference's include/ORBmatcher.h: DescriptorDistance (include/ORBmatcher.h:62,
// src/ORBmatcher.cc:2073-2093), the public thresholds (:223-225) and the best / second-best + ratio idiom every Search*
// function repeats (canonical form src/ORBmatcher.cc:304-342) as a brute-force k=2 call on the GPU.
// Of the guided Search* / Fuse functions the shared core — GetFeaturesInArea, the candidate scan with
// sequential claims and the rotation histogram — is offered as one flat call, SearchByProjection(TargetFrame, points, ...);
// pose / camera projection stay in the reference's own ORBmatcher.cc (INTEGRATION.md shows the rewritten loop).
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cstdint>
#include <map>
#include <utility>
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
  // The same scan over a database that is sharded over the ranks of an NCCL communicator (one process per GPU; contiguous shards, ascending rank =
  // ascending rows; shardBase = global index of the shard's first row): every rank passes its shard and gets the result of the scan over
  // the WHOLE database (orb_hamming_knn_sharded: shard scan + one ncclAllGather + device merge). Collective call. ncclComm: an ncclComm_t of
  // the process's NCCL, or one made by orb_nccl_comm_create (include/orb_b200.h).
  static void BruteForceKnn2Sharded(const cv::Mat& queries, const cv::Mat& databaseShard, int64_t shardBase, void* ncclComm, std::vector<Knn2>& out,
                                    int sentinel = 256);

  // The accept test of src/ORBmatcher.cc:339-342: best <= th && best < mfNNratio * second (float32 product).
  bool Accept(const Knn2& m, int th) const { return m.bestDist <= th && (float)m.bestDist < mfNNratio * (float)m.bestDist2; }
  // Convenience: indices of accepted matches (-1 = rejected), threshold TH_LOW by default.
  int MatchBruteForce(const cv::Mat& queries, const cv::Mat& database, std::vector<int64_t>& matches, int th = TH_LOW) const;

  // ---- guided search (src/ORBmatcher.cc:69-184 and 1702-1866 after the projection step) ----
  struct ProjectedPoint {    // one map point projected into the target frame
    float u, v;              // mTrackProjX/Y (:101) or the pinhole projection (:1755-1756)
    float radius;            // window half-size handed to GetFeaturesInArea (:102, :1769), also the stereo tolerance
    float uRight;            // mTrackProjXR (:137) or u - mbf * invzc (:1806); only read where the keypoint has a right match
    float angle;             // angle of the point's own keypoint (rotation histogram, :1830)
    int minLevel, maxLevel;  // GetFeaturesInArea level window; maxLevel < 0 = open
    bool skip;               // !mbTrackInView / isBad() / outlier / behind the camera / outside the image
    bool claims;             // pMP->Observations() > 0 (:127-130, :1798-1801)
  };
  struct TargetFrame {       // the members of Frame the search reads
    const std::vector<cv::KeyPoint>* mvKeysUn;
    const cv::Mat* mDescriptors;                 // N x 32 CV_8U
    const std::vector<float>* mvuRight;          // nullptr for monocular
    const std::vector<unsigned char>* claimed;   // per keypoint: mvpMapPoints[i] && mvpMapPoints[i]->Observations() > 0; nullptr = none
    float mnMinX, mnMinY, mnMaxX, mnMaxY;
  };
  // Points are processed in index order, exactly like the reference's loops. assigned[i] = index of the point matched to
  // keypoint i (what the reference stores as F.mvpMapPoints[i] = pMP) or -1. useRatio selects the local-map form
  // (best / second best, level-conditional mfNNratio test, :174); otherwise best only plus the rotation histogram when
  // mbCheckOrientation (:1828-1863). Returns nmatches. GPU only.
  int SearchByProjection(const TargetFrame& F, const std::vector<ProjectedPoint>& points, const cv::Mat& pointDescriptors, int th,
                         bool useRatio, std::vector<int>& assigned, std::vector<int>* pointMatch = nullptr) const;

  // SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:570-712) with the members of F1 / F2 it
  // reads passed explicitly: F1.mvKeysUn + F1.mDescriptors, F2 as TargetFrame (mvuRight / claimed unused). vbPrevMatched is updated
  // for the matched keypoints (:707-709). Threshold TH_LOW, ratio mfNNratio. Returns nmatches. GPU only.
  int SearchForInitialization(const std::vector<cv::KeyPoint>& keys1Un, const cv::Mat& descriptors1, const TargetFrame& F2,
                              std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize) const;

  // SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:248-410) on flat inputs: keyframe descriptors / mvKeysUn /
  // "has a good map point" flags / mFeatVec, frame descriptors / mvKeys / mFeatVec (DBoW2::FeatureVector derives from this map type).
  // assigned[i] = keyframe feature matched to frame feature i (vpMapPointMatches[i] = vpMapPointsKF[assigned[i]]) or -1. Threshold
  // TH_LOW, ratio mfNNratio, rotation histogram iff mbCheckOrientation. Returns nmatches. GPU only.
  typedef std::map<unsigned int, std::vector<unsigned int>> FeatureVector;
  int SearchByBoW(const cv::Mat& kfDescriptors, const std::vector<cv::KeyPoint>& kfKeysUn, const std::vector<unsigned char>& kfHasPoint,
                  const FeatureVector& kfFeatVec, const cv::Mat& fDescriptors, const std::vector<cv::KeyPoint>& fKeys,
                  const FeatureVector& fFeatVec, std::vector<int>& assigned) const;

  // The candidate search of Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:1100-1280), Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
  // (:1282-1430) and of SearchBySim3's two projection passes (:1433-1700): every projected point on its own (these loops hold no
  // claims between points). points[i].minLevel / maxLevel = nPredictedLevel - 1 / nPredictedLevel (:1188); skip = the reference's
  // `continue`s before GetFeaturesInArea; invLevelSigma2 (pKF->mvInvLevelSigma2) switches the chi-square reprojection test of the
  // first Fuse form on (:1193-1218), nullptr = off. bestIdx[i] = the keypoint the reference would fuse / match (bestDist <= th) or
  // -1; bestDist[i] = 256 without a candidate. The Replace / AddObservation bookkeeping stays with the caller. GPU only.
  void FuseSearch(const TargetFrame& KF, const std::vector<ProjectedPoint>& points, const cv::Mat& pointDescriptors,
                  const std::vector<float>* invLevelSigma2, int th, std::vector<int>& bestIdx, std::vector<int>* bestDist = nullptr) const;
  // SearchBySim3's matching (:1528-1690) after the caller projected KF1's points into KF2 (points1in2, SKIP where the reference
  // `continue`s, vbAlreadyMatched1 included) and vice versa: vnMatch1 / vnMatch2 by FuseSearch with TH_HIGH, then the mutual check
  // "vnMatch2[vnMatch1[i1]] == i1" (:1672-1686). matches12[i1] = index in KF2 or -1; returns nFound.
  int SearchBySim3(const TargetFrame& KF1, const TargetFrame& KF2, const std::vector<ProjectedPoint>& points1in2, const cv::Mat& descriptors1,
                   const std::vector<ProjectedPoint>& points2in1, const cv::Mat& descriptors2, std::vector<int>& matches12) const;

  // SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:879-1090) on flat inputs: per keyframe
  // mDescriptors / mvKeysUn / "GetMapPoint(idx) != NULL" flags / mvuRight / mFeatVec; F12 (3x3 CV_32F); the epipole (ex, ey) of camera 1
  // in image 2 computed by the caller as in :885-897; pKF2's mvLevelSigma2 and mvScaleFactors. Returns nmatches. GPU only.
  int SearchForTriangulation(const cv::Mat& descriptors1, const std::vector<cv::KeyPoint>& keys1Un, const std::vector<unsigned char>& hasPoint1,
                             const std::vector<float>& uRight1, const FeatureVector& featVec1, const cv::Mat& descriptors2,
                             const std::vector<cv::KeyPoint>& keys2Un, const std::vector<unsigned char>& hasPoint2, const std::vector<float>& uRight2,
                             const FeatureVector& featVec2, const float F12[9], float ex, float ey, const std::vector<float>& levelSigma2,
                             const std::vector<float>& scaleFactors, std::vector<std::pair<size_t, size_t>>& vMatchedPairs, bool bOnlyStereo) const;

  // MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517) for many map points in one call: rows
  // [offsets[p], offsets[p+1]) of `observations` (total x 32 CV_8U) are point p's descriptors in the reference's map order;
  // bestIdx[p] = the row (relative to offsets[p]) the reference would clone into mDescriptor, -1 for a point without rows.
  static void ComputeDistinctiveDescriptors(const cv::Mat& observations, const std::vector<int64_t>& offsets, std::vector<int>& bestIdx);

  static const int TH_LOW;        // 50
  static const int TH_HIGH;       // 100
  static const int HISTO_LENGTH;  // 30

 protected:
  float mfNNratio;
  bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2
#endif
