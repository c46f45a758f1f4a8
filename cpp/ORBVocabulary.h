// Drop-in for include/ORBVocabulary.h (typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary):
// the two members ORB-SLAM2 uses on the per-frame path — loadFromTextFile (System.cc, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1368-1440)
// and transform(features, BowVector&, FeatureVector&, levelsup) (Frame::ComputeBoW src/Frame.cc:750-765, KeyFrame::ComputeBoW) — over the
// C ABI (orb_vocabulary_create / orb_bow_transform). DBoW2::BowVector and DBoW2::FeatureVector derive from the std::map types used here, so
// `mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)` compiles unchanged. score() and the KeyFrameDatabase queries stay with DBoW2.
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

#include <map>
#include <string>
#include <vector>

#ifdef ORB_B200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif

struct orb_vocabulary;   // C ABI handle, include/orb_b200.h

namespace ORB_SLAM2 {

class ORBVocabulary {
 public:
  typedef std::map<unsigned int, double> BowVector;                           // base of DBoW2::BowVector
  typedef std::map<unsigned int, std::vector<unsigned int>> FeatureVector;    // base of DBoW2::FeatureVector

  explicit ORBVocabulary(int device = 0) : mDevice(device) {}
  ~ORBVocabulary();
  ORBVocabulary(const ORBVocabulary&) = delete;
  ORBVocabulary& operator=(const ORBVocabulary&) = delete;

  // ORBvoc.txt: "k L scoring weighting" then "parent is_leaf d0..d31 weight" per node. Only the ORB configuration
  // (L1_NORM scoring, TF_IDF weighting) is accepted. Returns false on a malformed file, throws on a CUDA error.
  bool loadFromTextFile(const std::string& filename);
  // Same from the node table (entry 0 = root, ignored).
  void create(int k, int L, const std::vector<int>& parent, const std::vector<unsigned char>& isLeaf, const std::vector<unsigned char>& desc,
              const std::vector<double>& weight);

  bool empty() const { return mpHandle == nullptr; }
  unsigned int size() const;

  // Reference signature (Converter::toDescriptorVector gives one 1 x 32 row per feature).
  void transform(const std::vector<cv::Mat>& features, BowVector& v, FeatureVector& fv, int levelsup) const;
  // Same on the N x 32 descriptor matrix directly (no per-row Mat headers).
  void transform(const cv::Mat& descriptors, BowVector& v, FeatureVector& fv, int levelsup) const;

 private:
  orb_vocabulary* mpHandle = nullptr;
  int mDevice;
};

}  // namespace ORB_SLAM2
#endif
