// ORB_SLAM2::ORBmatcher hot-path members over the C ABI (include/orb_b200.h).
#include "ORBmatcher.h"

#include <cstring>
#include <stdexcept>
#include <string>

#include "../include/orb_b200.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;     // src/ORBmatcher.cc:49-54
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

static void check(int rc, const char* what) {
  if (rc < 0) throw std::runtime_error(std::string(what) + ": " + orb_last_error());
}

// rows of a CV_8U N x 32 matrix as one contiguous block (copy only if the matrix is strided)
static const uint8_t* rows32(const cv::Mat& m, std::vector<uint8_t>& tmp) {
  if (m.cols != 32 || m.type() != CV_8U) throw std::runtime_error("descriptor matrix must be N x 32 CV_8U");
  if (m.isContinuous()) return m.ptr<unsigned char>(0);
  tmp.resize((size_t)m.rows * 32);
  for (int r = 0; r < m.rows; ++r) std::memcpy(&tmp[(size_t)r * 32], m.ptr<unsigned char>(r), 32);
  return tmp.data();
}

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
  // 256 bits = four 64-bit words; the reference's SWAR popcount over eight 32-bit words gives the same integer
  uint64_t wa[4], wb[4];
  std::memcpy(wa, a.ptr<unsigned char>(0), 32);
  std::memcpy(wb, b.ptr<unsigned char>(0), 32);
  int dist = 0;
  for (int i = 0; i < 4; ++i) dist += __builtin_popcountll(wa[i] ^ wb[i]);
  return dist;
}

void ORBmatcher::DescriptorDistanceBatch(const cv::Mat& a, const cv::Mat& b, std::vector<int>& out) {
  if (a.rows != b.rows) throw std::runtime_error("DescriptorDistanceBatch: row counts differ");
  out.assign(a.rows, 0);
  if (a.rows == 0) return;
  std::vector<uint8_t> ta, tb;
  const uint8_t* pa = rows32(a, ta);
  const uint8_t* pb = rows32(b, tb);
  check(orb_descriptor_distance_host(pa, pb, a.rows, out.data()), "orb_descriptor_distance_host");
}

void ORBmatcher::BruteForceKnn2(const cv::Mat& queries, const cv::Mat& database, std::vector<Knn2>& out, int sentinel) {
  out.clear();
  if (queries.rows == 0) return;
  std::vector<uint8_t> tq, td;
  const uint8_t* pq = rows32(queries, tq);
  const uint8_t* pd = database.rows ? rows32(database, td) : nullptr;
  std::vector<orb_knn2> res(queries.rows);
  check(orb_hamming_knn(pq, ORB_MEM_HOST, queries.rows, pd, ORB_MEM_HOST, database.rows, 0, sentinel, res.data(), ORB_MEM_HOST, nullptr),
        "orb_hamming_knn");
  out.resize(queries.rows);
  for (int i = 0; i < queries.rows; ++i) out[i] = Knn2{res[i].d1, res[i].d2, res[i].i1, res[i].i2};
}

void ORBmatcher::BruteForceKnn2Sharded(const cv::Mat& queries, const cv::Mat& databaseShard, int64_t shardBase, void* ncclComm, std::vector<Knn2>& out,
                                       int sentinel) {
  out.clear();
  if (queries.rows == 0) return;   // every rank sees the same queries, so every rank returns here together
  std::vector<uint8_t> tq, td;
  const uint8_t* pq = rows32(queries, tq);
  const uint8_t* pd = databaseShard.rows ? rows32(databaseShard, td) : nullptr;
  std::vector<orb_knn2> res(queries.rows);
  check(orb_hamming_knn_sharded(pq, ORB_MEM_HOST, queries.rows, pd, ORB_MEM_HOST, databaseShard.rows, shardBase, sentinel, res.data(), ORB_MEM_HOST,
                                ncclComm, nullptr),
        "orb_hamming_knn_sharded");
  out.resize(queries.rows);
  for (int i = 0; i < queries.rows; ++i) out[i] = Knn2{res[i].d1, res[i].d2, res[i].i1, res[i].i2};
}

int ORBmatcher::MatchBruteForce(const cv::Mat& queries, const cv::Mat& database, std::vector<int64_t>& matches, int th) const {
  std::vector<Knn2> r;
  BruteForceKnn2(queries, database, r);
  matches.assign(r.size(), -1);
  int n = 0;
  for (size_t i = 0; i < r.size(); ++i)
    if (r[i].bestIdx >= 0 && Accept(r[i], th)) { matches[i] = r[i].bestIdx; ++n; }
  return n;
}

int ORBmatcher::SearchByProjection(const TargetFrame& F, const std::vector<ProjectedPoint>& points, const cv::Mat& pointDescriptors, int th,
                                   bool useRatio, std::vector<int>& assigned, std::vector<int>* pointMatch) const {
  static_assert(sizeof(cv::KeyPoint) == sizeof(orb_keypoint), "cv::KeyPoint must be the 28-byte POD the C ABI reads");
  const int n = (int)F.mvKeysUn->size(), m = (int)points.size();
  assigned.assign(n, -1);
  if (pointMatch) pointMatch->assign(m, -1);
  if (n == 0 || m == 0) return 0;
  if (F.mDescriptors->rows != n || pointDescriptors.rows != m) throw std::runtime_error("SearchByProjection: descriptor rows != keypoints / points");
  std::vector<uint8_t> tmp_d, tmp_q;
  const uint8_t* d = rows32(*F.mDescriptors, tmp_d);
  const uint8_t* qd = rows32(pointDescriptors, tmp_q);
  std::vector<orb_proj_query> q(m);
  for (int i = 0; i < m; ++i) {
    const ProjectedPoint& p = points[i];
    q[i].u = p.u; q[i].v = p.v; q[i].radius = p.radius; q[i].ur = p.uRight; q[i].angle = p.angle;
    q[i].min_level = (int16_t)p.minLevel; q[i].max_level = (int16_t)p.maxLevel;
    q[i].flags = (p.skip ? ORB_QUERY_SKIP : 0) | (p.claims ? ORB_QUERY_CLAIMS : 0);
    q[i].reserved = 0;
  }
  orb_search_params prm;
  prm.min_x = F.mnMinX; prm.min_y = F.mnMinY; prm.max_x = F.mnMaxX; prm.max_y = F.mnMaxY;
  prm.th_dist = th;
  prm.nn_ratio = useRatio ? mfNNratio : 0.f;
  prm.check_orientation = (!useRatio && mbCheckOrientation) ? 1 : 0;
  prm.reserved = 0;
  std::vector<int32_t> qm(m);
  int32_t nmatches = 0;
  check(orb_search_by_projection(ORB_MEM_HOST, reinterpret_cast<const orb_keypoint*>(F.mvKeysUn->data()), d,
                                 F.mvuRight ? F.mvuRight->data() : nullptr, F.claimed ? F.claimed->data() : nullptr, &n, n, q.data(), qd,
                                 &m, m, 1, &prm, assigned.data(), qm.data(), &nmatches, nullptr),
        "orb_search_by_projection");
  if (pointMatch) pointMatch->assign(qm.begin(), qm.end());
  return nmatches;
}

void ORBmatcher::FuseSearch(const TargetFrame& KF, const std::vector<ProjectedPoint>& points, const cv::Mat& pointDescriptors,
                            const std::vector<float>* invLevelSigma2, int th, std::vector<int>& bestIdx, std::vector<int>* bestDist) const {
  const int n = (int)KF.mvKeysUn->size(), m = (int)points.size();
  bestIdx.assign(m, -1);
  if (bestDist) bestDist->assign(m, 256);
  if (n == 0 || m == 0) return;
  if (KF.mDescriptors->rows != n || pointDescriptors.rows != m) throw std::runtime_error("FuseSearch: descriptor rows != keypoints / points");
  if (invLevelSigma2 && invLevelSigma2->size() > 16) throw std::runtime_error("FuseSearch: more than 16 pyramid levels");
  std::vector<uint8_t> tmp_d, tmp_q;
  const uint8_t* d = rows32(*KF.mDescriptors, tmp_d);
  const uint8_t* qd = rows32(pointDescriptors, tmp_q);
  std::vector<orb_proj_query> q(m);
  for (int i = 0; i < m; ++i) {
    const ProjectedPoint& p = points[i];
    q[i].u = p.u; q[i].v = p.v; q[i].radius = p.radius; q[i].ur = p.uRight; q[i].angle = p.angle;
    q[i].min_level = (int16_t)p.minLevel; q[i].max_level = (int16_t)p.maxLevel;
    q[i].flags = p.skip ? ORB_QUERY_SKIP : 0;
    q[i].reserved = 0;
  }
  orb_search_params prm;
  prm.min_x = KF.mnMinX; prm.min_y = KF.mnMinY; prm.max_x = KF.mnMaxX; prm.max_y = KF.mnMaxY;
  prm.th_dist = th; prm.nn_ratio = 0.f; prm.check_orientation = 0; prm.reserved = 0;
  float sig[16] = {0.f};
  if (invLevelSigma2) std::copy(invLevelSigma2->begin(), invLevelSigma2->end(), sig);
  std::vector<int32_t> bd(m);
  check(orb_fuse_search(ORB_MEM_HOST, reinterpret_cast<const orb_keypoint*>(KF.mvKeysUn->data()), d, KF.mvuRight ? KF.mvuRight->data() : nullptr, &n, n,
                        q.data(), qd, &m, m, 1, &prm, invLevelSigma2 ? sig : nullptr, bestIdx.data(), bd.data(), nullptr),
        "orb_fuse_search");
  if (bestDist) bestDist->assign(bd.begin(), bd.end());
}

int ORBmatcher::SearchBySim3(const TargetFrame& KF1, const TargetFrame& KF2, const std::vector<ProjectedPoint>& points1in2, const cv::Mat& descriptors1,
                             const std::vector<ProjectedPoint>& points2in1, const cv::Mat& descriptors2, std::vector<int>& matches12) const {
  std::vector<int> vnMatch1, vnMatch2;
  FuseSearch(KF2, points1in2, descriptors1, nullptr, TH_HIGH, vnMatch1);   // :1528-1580
  FuseSearch(KF1, points2in1, descriptors2, nullptr, TH_HIGH, vnMatch2);   // :1590-1660
  const int N1 = (int)points1in2.size();
  matches12.assign(N1, -1);
  int nFound = 0;
  for (int i1 = 0; i1 < N1; ++i1) {                                        // :1672-1686
    const int idx2 = vnMatch1[i1];
    if (idx2 >= 0 && idx2 < (int)vnMatch2.size() && vnMatch2[idx2] == i1) {
      matches12[i1] = idx2;
      ++nFound;
    }
  }
  return nFound;
}

void ORBmatcher::ComputeDistinctiveDescriptors(const cv::Mat& observations, const std::vector<int64_t>& offsets, std::vector<int>& bestIdx) {
  const int points = (int)offsets.size() - 1;
  bestIdx.assign(points > 0 ? points : 0, -1);
  if (points <= 0) return;
  std::vector<uint8_t> tmp;
  const uint8_t* d = observations.rows > 0 ? rows32(observations, tmp) : nullptr;
  check(orb_distinctive_descriptors(ORB_MEM_HOST, d, offsets.data(), points, bestIdx.data(), nullptr, nullptr), "orb_distinctive_descriptors");
}

static void to_csr(const ORBmatcher::FeatureVector& fv, int cap, std::vector<int32_t>& node, std::vector<int32_t>& start, std::vector<int32_t>& feat,
                   int32_t& count) {
  node.assign(cap, 0); start.assign(cap + 1, 0); feat.assign(cap, 0);
  count = 0;
  int pos = 0;
  for (const auto& e : fv) {
    if (count >= cap || pos + (int)e.second.size() > cap) throw std::runtime_error("SearchByBoW: FeatureVector larger than the feature count");
    node[count] = (int32_t)e.first;
    start[count] = pos;
    for (unsigned int i : e.second) feat[pos++] = (int32_t)i;
    ++count;
  }
  start[count] = pos;
}

int ORBmatcher::SearchByBoW(const cv::Mat& kfDescriptors, const std::vector<cv::KeyPoint>& kfKeysUn, const std::vector<unsigned char>& kfHasPoint,
                            const FeatureVector& kfFeatVec, const cv::Mat& fDescriptors, const std::vector<cv::KeyPoint>& fKeys,
                            const FeatureVector& fFeatVec, std::vector<int>& assigned) const {
  const int nk = (int)kfKeysUn.size(), nf = (int)fKeys.size();
  assigned.assign(nf, -1);
  if (nk == 0 || nf == 0) return 0;
  if (kfDescriptors.rows != nk || fDescriptors.rows != nf || (int)kfHasPoint.size() != nk) throw std::runtime_error("SearchByBoW: array sizes differ");
  std::vector<uint8_t> tk, tf;
  const uint8_t* dk = rows32(kfDescriptors, tk);
  const uint8_t* df = rows32(fDescriptors, tf);
  std::vector<int32_t> kn, ks, kfe, fn, fs, ffe;
  int32_t kc = 0, fc = 0;
  to_csr(kfFeatVec, nk, kn, ks, kfe, kc);
  to_csr(fFeatVec, nf, fn, fs, ffe, fc);
  int32_t nm = 0;
  check(orb_search_by_bow(ORB_MEM_HOST, dk, reinterpret_cast<const orb_keypoint*>(kfKeysUn.data()), kfHasPoint.data(), kn.data(), ks.data(), kfe.data(),
                          &kc, nk, df, reinterpret_cast<const orb_keypoint*>(fKeys.data()), nullptr, fn.data(), fs.data(), ffe.data(), &fc, nf, 1,
                          TH_LOW, mfNNratio, mbCheckOrientation ? 1 : 0, assigned.data(), nullptr, &nm, nullptr),
        "orb_search_by_bow");
  return nm;
}

int ORBmatcher::SearchForTriangulation(const cv::Mat& descriptors1, const std::vector<cv::KeyPoint>& keys1Un, const std::vector<unsigned char>& hasPoint1,
                                       const std::vector<float>& uRight1, const FeatureVector& featVec1, const cv::Mat& descriptors2,
                                       const std::vector<cv::KeyPoint>& keys2Un, const std::vector<unsigned char>& hasPoint2,
                                       const std::vector<float>& uRight2, const FeatureVector& featVec2, const float F12[9], float ex, float ey,
                                       const std::vector<float>& levelSigma2, const std::vector<float>& scaleFactors,
                                       std::vector<std::pair<size_t, size_t>>& vMatchedPairs, bool bOnlyStereo) const {
  const int n1 = (int)keys1Un.size(), n2 = (int)keys2Un.size();
  vMatchedPairs.clear();
  if (n1 == 0 || n2 == 0) return 0;
  if (descriptors1.rows != n1 || descriptors2.rows != n2 || (int)hasPoint1.size() != n1 || (int)hasPoint2.size() != n2 || (int)uRight1.size() != n1 ||
      (int)uRight2.size() != n2 || levelSigma2.size() > 16 || scaleFactors.size() != levelSigma2.size())
    throw std::runtime_error("SearchForTriangulation: array sizes differ");
  std::vector<uint8_t> t1, t2;
  const uint8_t* d1 = rows32(descriptors1, t1);
  const uint8_t* d2 = rows32(descriptors2, t2);
  std::vector<int32_t> an, as, af, bn, bs, bf;
  int32_t ac = 0, bc = 0;
  to_csr(featVec1, n1, an, as, af, ac);
  to_csr(featVec2, n2, bn, bs, bf, bc);
  std::vector<unsigned char> free1(n1), blocked2(n2);
  for (int i = 0; i < n1; ++i) free1[i] = !hasPoint1[i] && (!bOnlyStereo || uRight1[i] >= 0);        // :925-934
  for (int i = 0; i < n2; ++i) blocked2[i] = hasPoint2[i] || (bOnlyStereo && !(uRight2[i] >= 0));    // :957-966
  orb_epipolar epi;
  std::memset(&epi, 0, sizeof(epi));
  std::memcpy(epi.f12, F12, sizeof(epi.f12));
  epi.ex = ex; epi.ey = ey;
  std::copy(levelSigma2.begin(), levelSigma2.end(), epi.level_sigma2);
  std::copy(scaleFactors.begin(), scaleFactors.end(), epi.scale_factor);
  std::vector<int32_t> m12(n1, -1);
  int32_t nm = 0;
  check(orb_search_for_triangulation(ORB_MEM_HOST, d1, reinterpret_cast<const orb_keypoint*>(keys1Un.data()), free1.data(), uRight1.data(), an.data(),
                                     as.data(), af.data(), &ac, n1, d2, reinterpret_cast<const orb_keypoint*>(keys2Un.data()), blocked2.data(),
                                     uRight2.data(), bn.data(), bs.data(), bf.data(), &bc, n2, 1, &epi, TH_LOW, mbCheckOrientation ? 1 : 0, m12.data(),
                                     &nm, nullptr),
        "orb_search_for_triangulation");
  vMatchedPairs.reserve(nm);
  for (int i = 0; i < n1; ++i)                                                                       // :1078-1086
    if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));
  return nm;
}

int ORBmatcher::SearchForInitialization(const std::vector<cv::KeyPoint>& keys1Un, const cv::Mat& descriptors1, const TargetFrame& F2,
                                        std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize) const {
  const int m = (int)keys1Un.size(), n = (int)F2.mvKeysUn->size();
  vnMatches12.assign(m, -1);   // :574
  if (m == 0 || n == 0) return 0;
  if (descriptors1.rows != m || F2.mDescriptors->rows != n || (int)vbPrevMatched.size() != m)
    throw std::runtime_error("SearchForInitialization: array sizes differ");
  std::vector<uint8_t> t1, t2;
  const uint8_t* d1 = rows32(descriptors1, t1);
  const uint8_t* d2 = rows32(*F2.mDescriptors, t2);
  std::vector<orb_proj_query> q(m);
  for (int i = 0; i < m; ++i) {
    q[i].u = vbPrevMatched[i].x; q[i].v = vbPrevMatched[i].y; q[i].radius = (float)windowSize; q[i].ur = -1.f;
    q[i].angle = keys1Un[i].angle;
    q[i].min_level = 0; q[i].max_level = 0;                      // GetFeaturesInArea(x, y, windowSize, level1, level1) with level1 == 0 (:602)
    q[i].flags = keys1Un[i].octave > 0 ? ORB_QUERY_SKIP : 0;     // :596-597
    q[i].reserved = 0;
  }
  orb_search_params prm;
  prm.min_x = F2.mnMinX; prm.min_y = F2.mnMinY; prm.max_x = F2.mnMaxX; prm.max_y = F2.mnMaxY;
  prm.th_dist = TH_LOW; prm.nn_ratio = mfNNratio; prm.check_orientation = mbCheckOrientation ? 1 : 0; prm.reserved = 0;
  int32_t nm = 0;
  check(orb_search_for_initialization(ORB_MEM_HOST, reinterpret_cast<const orb_keypoint*>(F2.mvKeysUn->data()), d2, &n, n, q.data(), d1, &m, m, 1,
                                      &prm, vnMatches12.data(), nullptr, &nm, nullptr),
        "orb_search_for_initialization");
  for (int i = 0; i < m; ++i)
    if (vnMatches12[i] >= 0) vbPrevMatched[i] = (*F2.mvKeysUn)[vnMatches12[i]].pt;   // :707-709
  return nm;
}

}  // namespace ORB_SLAM2
