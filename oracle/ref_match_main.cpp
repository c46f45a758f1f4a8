// TEST INFRASTRUCTURE ONLY — pins the matcher-side oracles (M1, M2, N1, N4, ComputeThreeMaxima) to the reference's OWN code.
//
// src/ORBmatcher.cc, src/Frame.cc and src/MapPoint.cc cannot be compiled whole here (they include Frame / KeyFrame / MapPoint / Map,
// Eigen, g2o and Pangolin headers the image lacks), but the function bodies on the hot path touch none of that. oracle/Makefile
// therefore `sed -n`-extracts exactly those line ranges from /root/reference into oracle/_ref/gen_*.inc at build time (generated,
// git-ignored, never committed) and this file supplies the few declarations they need: stub classes with the members the bodies
// name (same names, same types) over oracle/cvshim. The text that runs is the reference's, byte for byte.
//
//   gen_consts.inc        src/ORBmatcher.cc:49-51        TH_HIGH / TH_LOW / HISTO_LENGTH definitions
//   gen_distance.inc      src/ORBmatcher.cc:2073-2093    ORBmatcher::DescriptorDistance
//   gen_three_maxima.inc  src/ORBmatcher.cc:2026-2068    ORBmatcher::ComputeThreeMaxima
//   gen_scan.inc          src/ORBmatcher.cc:304-335      best / second-best scan of SearchByBoW (sentinel 256)
//   gen_accept.inc        src/ORBmatcher.cc:339-343      `if (best <= TH_LOW) { if ((float)best < mfNNratio * (float)second) {`
//   gen_scan_init.inc     src/ORBmatcher.cc:611-637      scan of SearchForInitialization (sentinel INT_MAX, vMatchedDistance skip)
//   gen_accept_init.inc   src/ORBmatcher.cc:641-645      `if (best <= TH_LOW) { if (best < (float)second * mfNNratio) {`
//   gen_distinctive.inc   src/MapPoint.cc:468-510        all-pairs distances + least-median row of ComputeDistinctiveDescriptors
//   gen_stereo.inc        src/Frame.cc:880-1176          Frame::ComputeStereoMatches
//   gen_sbp.inc           src/ORBmatcher.cc:69-194       SearchByProjection(Frame&, vector<MapPoint*>&, th) + RadiusByViewingCos (N2, local-map form)
//   gen_sbow.inc          src/ORBmatcher.cc:248-410      SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (N2: FeatureVector walk, rotation filter)
//   gen_sinit.inc         src/ORBmatcher.cc:570-712      SearchForInitialization (N2: vMatchedDistance / eviction bookkeeping, rotation filter)
//   gen_grid.inc          src/Frame.cc:432-460           Frame::AssignFeaturesToGrid
//   gen_area.inc          src/Frame.cc:633-717           Frame::GetFeaturesInArea
//   gen_posingrid.inc     src/Frame.cc:729-745           Frame::PosInGrid
//   gen_griddefs.inc      include/Frame.h:54, 59         #define FRAME_GRID_ROWS 48 / FRAME_GRID_COLS 64
// FORB::distance (Thirdparty/DBoW2/DBoW2/FORB.cpp:81-103, the verbatim twin of DescriptorDistance) is linked in compiled unmodified.
//
// usage: orb_ref_match MODE IN.bin OUT.bin     (formats: tests/ref_binary.py)
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <list>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "FORB.h"
#include "FeatureVector.h"
#include "cvshim/cvshim.h"

using namespace std;   // the reference's sources say `vector`, `sort`, `pair` unqualified (src/ORBmatcher.cc:43)

namespace ORB_SLAM2 {

class ORBextractor {   // include/ORBextractor.h:185: the one member ComputeStereoMatches reads
 public:
  std::vector<cv::Mat> mvImagePyramid;
};

#include "_ref/gen_griddefs.inc"

class KeyFrame;
class Frame;
class MapPoint;
static MapPoint* g_cur_point = nullptr;   // the map point whose candidate scan is running (set by GetDescriptor, read by KeyFrame::GetMapPoint)

class MapPoint {       // include/MapPoint.h: the members the matcher loops name
 public:
  bool mbTrackInView = true;
  int mnTrackScaleLevel = 0;
  float mTrackViewCos = 1.f, mTrackProjX = 0.f, mTrackProjY = 0.f, mTrackProjXR = 0.f;
  bool isBad() { return bad; }
  int Observations() { return nObs; }
  cv::Mat GetDescriptor() { g_cur_point = this; return mDescriptor.clone(); }   // include/MapPoint.h: returns mDescriptor.clone()
  cv::Mat GetWorldPos() { return mWorldPos.clone(); }       // include/MapPoint.h / src/MapPoint.cc:117-121
  // what Fuse asks of a map point (src/ORBmatcher.cc:1123-1160): host-side quantities here, plain stored values
  bool IsInKeyFrame(KeyFrame*) { return inKF; }
  float GetMaxDistanceInvariance() { return maxD; }
  float GetMinDistanceInvariance() { return minD; }
  cv::Mat GetNormal() { return mNormal.clone(); }
  int PredictScale(const float&, KeyFrame*) { return predLevel; }
  int PredictScale(const float&, Frame*) { return predLevel; }
  int GetIndexInKeyFrame(KeyFrame*) { return idxInKF2; }     // SearchBySim3 asks it of the pre-matched points of keyframe 2 (:1469)
  int idxInKF2 = -1;
  void Replace(MapPoint*) {}
  void AddObservation(KeyFrame*, size_t) {}
  bool inKF = false;
  float maxD = 3.0e38f, minD = 0.f;
  int predLevel = 0, fusedIdx = -1;
  cv::Mat mNormal;
  bool bad = false;
  int nObs = 0;
  cv::Mat mDescriptor, mWorldPos;
};

class Frame {          // include/Frame.h: the members src/Frame.cc:880-1176, 432-460, 633-745 and the matcher loops name
 public:
  void ComputeStereoMatches();
  void AssignFeaturesToGrid();
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
  vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1) const;
  int N;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  cv::Mat mDescriptors, mDescriptorsRight;
  ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
  std::vector<float> mvScaleFactors, mvInvScaleFactors;
  std::vector<float> mvuRight, mvDepth;
  std::vector<MapPoint*> mvpMapPoints;
  DBoW2::FeatureVector mFeatVec;
  float mb, mbf;
  cv::Mat mTcw;                                              // include/Frame.h:160
  std::vector<bool> mvbOutlier;
  float fx, fy, cx, cy;                                      // static in the reference (include/Frame.h:113-118)
  // static members of the reference's class (include/Frame.h:190-200): plain members here, named unqualified inside the member functions
  float mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
};

class KeyFrame {       // include/KeyFrame.h: what SearchByBoW and Fuse read
 public:
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  std::set<MapPoint*> GetMapPoints() { std::set<MapPoint*> s; for (MapPoint* p : mvpMapPoints) if (p) s.insert(p); return s; }   // src/KeyFrame.cc:548-563 without isBad (stub points are good)
  cv::Mat GetRotation() { return Rcw.clone(); }
  cv::Mat GetTranslation() { return tcw.clone(); }
  cv::Mat GetCameraCenter() { return Ow.clone(); }
  vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const;
  bool IsInImage(const float& x, const float& y) const;
  MapPoint* GetMapPoint(const size_t& idx) { if (g_cur_point) g_cur_point->fusedIdx = (int)idx; return mvpMapPoints[idx]; }   // called once per accepted point (:1243)
  void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
  cv::Mat Rcw, tcw, Ow;
  int N = 0, mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;
  float fx, fy, cx, cy, mbf, mnMinX, mnMinY, mnMaxX, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<float> mvScaleFactors, mvInvLevelSigma2, mvLevelSigma2, mvuRight;
  std::vector<std::vector<std::vector<size_t>>> mGrid;
  DBoW2::FeatureVector mFeatVec;
  cv::Mat mDescriptors;
  std::vector<cv::KeyPoint> mvKeysUn;
  std::vector<MapPoint*> mvpMapPoints;
};

class ORBmatcher {     // include/ORBmatcher.h:54-62, 223-236
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
  static const int TH_LOW;
  static const int TH_HIGH;
  static const int HISTO_LENGTH;
  void ComputeThreeMaxima(vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
  int SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th = 3);
  int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
  int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, const float th, const int ORBdist);
  int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, int th);
  int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12, const cv::Mat& t12, const float th);
  int SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches);
  int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12);
  int Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th = 3.0);
  int Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint);
  bool CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12, const KeyFrame* pKF);
  int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, vector<pair<size_t, size_t>>& vMatchedPairs, const bool bOnlyStereo);
  int SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched, vector<int>& vnMatches12, int windowSize = 10);
  float RadiusByViewingCos(const float& viewCos);
  // harness entry points around the extracted loop bodies (declared here so the bodies see the members unqualified)
  void Scan(const cv::Mat& Q, const cv::Mat& D, int32_t* out);
  void ScanInit(const cv::Mat& Q, const cv::Mat& D, const std::vector<int>& matched_distance, int32_t* out);
  float mfNNratio;
  bool mbCheckOrientation;
};

#include "_ref/gen_consts.inc"
#include "_ref/gen_distance.inc"
#include "_ref/gen_three_maxima.inc"
#include "_ref/gen_stereo.inc"
#include "_ref/gen_grid.inc"
#include "_ref/gen_area.inc"
#include "_ref/gen_posingrid.inc"
#include "_ref/gen_sbp.inc"
#include "_ref/gen_sbow.inc"
#include "_ref/gen_sinit.inc"
#include "_ref/gen_sbpf.inc"
#include "_ref/gen_sbow2.inc"
#include "_ref/gen_kfarea.inc"
#include "_ref/gen_fuse.inc"
#include "_ref/gen_fuse2.inc"
#include "_ref/gen_epiline.inc"
#include "_ref/gen_stri.inc"
#include "_ref/gen_sbpk.inc"
#include "_ref/gen_sbps.inc"
#include "_ref/gen_sim3.inc"

struct FrameDescs { cv::Mat mDescriptors; };

// src/ORBmatcher.cc:304-343 with every candidate of the frame in one vocabulary node and no map point matched yet
void ORBmatcher::Scan(const cv::Mat& Q, const cv::Mat& D, int32_t* out) {
  FrameDescs F;
  F.mDescriptors = D;
  std::vector<unsigned int> vIndicesF(D.rows);
  for (int i = 0; i < D.rows; ++i) vIndicesF[i] = (unsigned int)i;
  std::vector<void*> vpMapPointMatches(D.rows, (void*)nullptr);
  for (int q = 0; q < Q.rows; ++q) {
    const cv::Mat& dKF = Q.row(q);
#include "_ref/gen_scan.inc"
    int accept = 0;
#include "_ref/gen_accept.inc"
    accept = 1; } }
    out[4 * q] = bestDist1; out[4 * q + 1] = bestDist2; out[4 * q + 2] = bestIdxF; out[4 * q + 3] = accept;
  }
}

// src/ORBmatcher.cc:611-645 with every keypoint of F2 in the search window
void ORBmatcher::ScanInit(const cv::Mat& Q, const cv::Mat& D, const std::vector<int>& vMatchedDistance, int32_t* out) {
  FrameDescs F2;
  F2.mDescriptors = D;
  std::vector<size_t> vIndices2(D.rows);
  for (int i = 0; i < D.rows; ++i) vIndices2[i] = (size_t)i;
  for (int q = 0; q < Q.rows; ++q) {
    cv::Mat d1 = Q.row(q);
#include "_ref/gen_scan_init.inc"
    int accept = 0;
#include "_ref/gen_accept_init.inc"
    accept = 1; } }
    out[4 * q] = bestDist; out[4 * q + 1] = bestDist2; out[4 * q + 2] = bestIdx2; out[4 * q + 3] = accept;
  }
}

// src/MapPoint.cc:468-510
static void Distinctive(std::vector<cv::Mat>& vDescriptors, int32_t* out) {
#include "_ref/gen_distinctive.inc"
  out[0] = BestIdx; out[1] = BestMedian;
}

}  // namespace ORB_SLAM2

static std::vector<uint8_t> slurp(const char* path) {
  FILE* f = std::fopen(path, "rb");
  if (!f) { std::perror(path); std::exit(1); }
  std::fseek(f, 0, SEEK_END);
  const long n = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> b((size_t)n);
  if (n > 0 && std::fread(b.data(), 1, (size_t)n, f) != (size_t)n) { std::perror("read"); std::exit(1); }
  std::fclose(f);
  return b;
}
struct Reader {
  const uint8_t* p;
  template <typename T> T get() { T v; std::memcpy(&v, p, sizeof(T)); p += sizeof(T); return v; }
  const uint8_t* bytes(size_t n) { const uint8_t* q = p; p += n; return q; }
};
static cv::Mat mat_u8(const uint8_t* src, int rows, int cols) {
  cv::Mat m(rows > 0 ? rows : 1, cols, CV_8U);
  if (rows > 0) std::memcpy(m.data, src, (size_t)rows * cols);
  m.rows = rows;
  return m;
}

int main(int argc, char** argv) {
  if (argc != 4) { std::fprintf(stderr, "usage: %s distance|scan|scan_init|three_maxima|distinctive|stereo|sbp|sbow|sinit IN OUT\n", argv[0]); return 1; }
  const std::string mode = argv[1];
  std::vector<uint8_t> in = slurp(argv[2]);
  Reader r{in.data()};
  std::vector<int32_t> out;
  std::vector<float> outf;
  if (mode == "distance") {   // int32 n; n x (32 B, 32 B) -> n x {FORB::distance, ORBmatcher::DescriptorDistance}
    const int n = r.get<int32_t>();
    out.resize(2 * (size_t)n);
    for (int i = 0; i < n; ++i) {
      cv::Mat a = mat_u8(r.bytes(32), 1, 32), b = mat_u8(r.bytes(32), 1, 32);
      out[2 * i] = DBoW2::FORB::distance(a, b);
      out[2 * i + 1] = ORB_SLAM2::ORBmatcher::DescriptorDistance(a, b);
    }
  } else if (mode == "scan" || mode == "scan_init") {   // int32 Q, D; float ratio; Q x 32; D x 32; (scan_init: D x int32 vMatchedDistance)
    const int Q = r.get<int32_t>(), D = r.get<int32_t>();
    const float ratio = r.get<float>();
    cv::Mat q = mat_u8(r.bytes((size_t)Q * 32), Q, 32), d = mat_u8(r.bytes((size_t)D * 32), D, 32);
    ORB_SLAM2::ORBmatcher m(ratio, true);
    out.resize(4 * (size_t)Q);
    if (mode == "scan") m.Scan(q, d, out.data());
    else {
      std::vector<int> md(D);
      for (int i = 0; i < D; ++i) md[i] = r.get<int32_t>();
      m.ScanInit(q, d, md, out.data());
    }
  } else if (mode == "three_maxima") {   // int32 n, L; n x L bin sizes -> n x {ind1, ind2, ind3}
    const int n = r.get<int32_t>(), L = r.get<int32_t>();
    ORB_SLAM2::ORBmatcher m;
    out.resize(3 * (size_t)n);
    for (int i = 0; i < n; ++i) {
      std::vector<std::vector<int>> histo(L);
      for (int b = 0; b < L; ++b) histo[b].resize((size_t)r.get<int32_t>());
      int i1 = -1, i2 = -1, i3 = -1;
      m.ComputeThreeMaxima(histo.data(), L, i1, i2, i3);
      out[3 * i] = i1; out[3 * i + 1] = i2; out[3 * i + 2] = i3;
    }
  } else if (mode == "distinctive") {   // int32 P; per point int32 N, N x 32 -> P x {BestIdx, BestMedian}
    const int P = r.get<int32_t>();
    out.resize(2 * (size_t)P);
    for (int p = 0; p < P; ++p) {
      const int N = r.get<int32_t>();
      cv::Mat all = mat_u8(r.bytes((size_t)N * 32), N, 32);
      std::vector<cv::Mat> rows;
      for (int i = 0; i < N; ++i) rows.push_back(all.row(i));
      ORB_SLAM2::Distinctive(rows, &out[2 * p]);
    }
  } else if (mode == "stereo") {
    // int32 nlevels; float bf, b; nlevels x float scale; per side (L, R): nlevels x {int32 cols, rows, bytes}; int32 n; n x 28-B KeyPoint; n x 32 B
    // -> N x float mvuRight, N x float mvDepth
    ORB_SLAM2::Frame F;
    ORB_SLAM2::ORBextractor exl, exr;
    const int nl = r.get<int32_t>();
    F.mbf = r.get<float>(); F.mb = r.get<float>();
    for (int l = 0; l < nl; ++l) { const float s = r.get<float>(); F.mvScaleFactors.push_back(s); F.mvInvScaleFactors.push_back(1.0f / s); }   // src/Frame.cc:147-148
    for (int side = 0; side < 2; ++side) {
      ORB_SLAM2::ORBextractor& ex = side ? exr : exl;
      for (int l = 0; l < nl; ++l) {
        const int c = r.get<int32_t>(), rw = r.get<int32_t>();
        ex.mvImagePyramid.push_back(mat_u8(r.bytes((size_t)c * rw), rw, c));
      }
      const int n = r.get<int32_t>();
      std::vector<cv::KeyPoint> k((size_t)n);
      std::memcpy(k.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
      cv::Mat d = mat_u8(r.bytes((size_t)n * 32), n, 32);
      if (side == 0) { F.mvKeys = k; F.mDescriptors = d; F.N = n; } else { F.mvKeysRight = k; F.mDescriptorsRight = d; }
    }
    F.mpORBextractorLeft = &exl; F.mpORBextractorRight = &exr;
    F.ComputeStereoMatches();
    outf = F.mvuRight;
    outf.insert(outf.end(), F.mvDepth.begin(), F.mvDepth.end());
  } else if (mode == "sbp") {
    // float bounds[4] (mnMinX, mnMinY, mnMaxX, mnMaxY), th, nnratio; int nlevels; nlevels x float scale; int n; n x KeyPoint (mvKeysUn); n x 32 B;
    // n x float mvuRight; n x u8 "already holds a map point with observations"; int m; m x {float projX, projY, projXR, viewCos; int level, flags
    // (1 = !mbTrackInView, 2 = Observations() > 0)} ; m x 32 B  ->  int nmatches; n x int32 index of the map point assigned by this call (-1 none)
    ORB_SLAM2::Frame F;
    F.mnMinX = r.get<float>(); F.mnMinY = r.get<float>(); F.mnMaxX = r.get<float>(); F.mnMaxY = r.get<float>();
    const float th = r.get<float>(), nnratio = r.get<float>();
    const int nl = r.get<int32_t>();
    for (int l = 0; l < nl; ++l) F.mvScaleFactors.push_back(r.get<float>());
    F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);     // src/Frame.cc:200-202
    F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
    const int n = r.get<int32_t>();
    F.N = n;
    F.mvKeysUn.resize(n);
    std::memcpy(F.mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
    F.mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
    F.mvuRight.resize(n);
    std::memcpy(F.mvuRight.data(), r.bytes((size_t)n * 4), (size_t)n * 4);
    ORB_SLAM2::MapPoint taken;   // a map point with observations, standing for whatever the keypoint was matched to before the call
    taken.nObs = 1;
    F.mvpMapPoints.assign(n, nullptr);
    const uint8_t* cl = r.bytes(n);
    for (int i = 0; i < n; ++i) if (cl[i]) F.mvpMapPoints[i] = &taken;
    F.AssignFeaturesToGrid();
    const int m = r.get<int32_t>();
    std::vector<ORB_SLAM2::MapPoint> pts(m);
    for (int i = 0; i < m; ++i) {
      pts[i].mTrackProjX = r.get<float>(); pts[i].mTrackProjY = r.get<float>(); pts[i].mTrackProjXR = r.get<float>(); pts[i].mTrackViewCos = r.get<float>();
      pts[i].mnTrackScaleLevel = r.get<int32_t>();
      const int fl = r.get<int32_t>();
      pts[i].mbTrackInView = !(fl & 1);
      pts[i].nObs = (fl & 2) ? 1 : 0;
    }
    std::vector<ORB_SLAM2::MapPoint*> vp(m);
    for (int i = 0; i < m; ++i) { pts[i].mDescriptor = mat_u8(r.bytes(32), 1, 32); vp[i] = &pts[i]; }
    ORB_SLAM2::ORBmatcher matcher(nnratio, true);
    out.push_back(matcher.SearchByProjection(F, vp, th));
    for (int i = 0; i < n; ++i) {
      ORB_SLAM2::MapPoint* p = F.mvpMapPoints[i];
      out.push_back(p && p != &taken ? (int32_t)(p - pts.data()) : -1);
    }
  } else if (mode == "sbpf") {
    // ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1702-1871), the TrackWithMotionModel matcher.
    // float bounds[4], th, nnratio (unused by this form), mb, mbf, fx, fy, cx, cy, last_tz; int bMono, check_orientation, nlevels; nlevels x float
    // scale; current frame: int n; n x KeyPoint (mvKeysUn); n x 32 B; n x float mvuRight; n x u8 claimed; last frame: int m; m x {float X, Y, Z,
    // angle; int octave, flags (1 = no map point, 2 = Observations() > 0, 4 = outlier)}; m x 32 B
    //   ->  int nmatches; n x int32 index of the last-frame point assigned by this call (-1 none).
    // Poses: CurrentFrame.mTcw = identity, LastFrame.mTcw = translation (0, 0, last_tz): tlc(2) = last_tz decides forward / backward.
    ORB_SLAM2::Frame F, LF;
    F.mnMinX = r.get<float>(); F.mnMinY = r.get<float>(); F.mnMaxX = r.get<float>(); F.mnMaxY = r.get<float>();
    const float th = r.get<float>(), nnratio = r.get<float>();
    F.mb = r.get<float>(); F.mbf = r.get<float>();
    F.fx = r.get<float>(); F.fy = r.get<float>(); F.cx = r.get<float>(); F.cy = r.get<float>();
    const float last_tz = r.get<float>();
    const int bMono = r.get<int32_t>(), ori = r.get<int32_t>();
    const int nl = r.get<int32_t>();
    for (int l = 0; l < nl; ++l) F.mvScaleFactors.push_back(r.get<float>());
    F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);     // src/Frame.cc:200-202
    F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
    const int n = r.get<int32_t>();
    F.N = n;
    F.mvKeysUn.resize(n);
    std::memcpy(F.mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
    F.mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
    F.mvuRight.resize(n);
    std::memcpy(F.mvuRight.data(), r.bytes((size_t)n * 4), (size_t)n * 4);
    ORB_SLAM2::MapPoint taken;
    taken.nObs = 1;
    F.mvpMapPoints.assign(n, nullptr);
    const uint8_t* cl = r.bytes(n);
    for (int i = 0; i < n; ++i) if (cl[i]) F.mvpMapPoints[i] = &taken;
    F.AssignFeaturesToGrid();
    auto pose = [](float tz) {
      cv::Mat T(4, 4, CV_32F);
      for (int y = 0; y < 4; ++y) for (int x = 0; x < 4; ++x) T.at<float>(y, x) = y == x ? 1.f : 0.f;
      T.at<float>(2, 3) = tz;
      return T;
    };
    F.mTcw = pose(0.f);
    LF.mTcw = pose(last_tz);
    const int m = r.get<int32_t>();
    LF.N = m;
    std::vector<ORB_SLAM2::MapPoint> pts(m);
    LF.mvpMapPoints.assign(m, nullptr);
    LF.mvbOutlier.assign(m, false);
    LF.mvKeys.resize(m); LF.mvKeysUn.resize(m);
    for (int i = 0; i < m; ++i) {
      pts[i].mWorldPos = cv::Mat(3, 1, CV_32F);
      for (int k = 0; k < 3; ++k) pts[i].mWorldPos.at<float>(k) = r.get<float>();
      LF.mvKeysUn[i].angle = r.get<float>();
      LF.mvKeys[i].octave = r.get<int32_t>();
      const int fl = r.get<int32_t>();
      if (!(fl & 1)) LF.mvpMapPoints[i] = &pts[i];
      pts[i].nObs = (fl & 2) ? 1 : 0;
      LF.mvbOutlier[i] = (fl & 4) != 0;
    }
    for (int i = 0; i < m; ++i) pts[i].mDescriptor = mat_u8(r.bytes(32), 1, 32);
    ORB_SLAM2::ORBmatcher matcher(nnratio, ori != 0);
    out.push_back(matcher.SearchByProjection(F, LF, th, bMono != 0));
    for (int i = 0; i < n; ++i) {
      ORB_SLAM2::MapPoint* p = F.mvpMapPoints[i];
      out.push_back(p && p != &taken ? (int32_t)(p - pts.data()) : -1);
    }
  } else if (mode == "sbow") {
    // float nnratio; int checkOri; keyframe: int nkf; nkf x 32 B; nkf x KeyPoint (mvKeysUn); nkf x u8 "holds a good map point"; FeatureVector
    // {int count; count x {uint32 node; int m; m x uint32 feature}}; frame: int nf; nf x 32 B; nf x KeyPoint (mvKeys); FeatureVector
    // -> int nmatches; nf x int32 keyframe feature whose map point was assigned (-1 none)
    const float nnratio = r.get<float>();
    const int ori = r.get<int32_t>();
    ORB_SLAM2::KeyFrame KF;
    ORB_SLAM2::Frame F;
    auto read_fv = [&](DBoW2::FeatureVector& fv) {
      const int cnt = r.get<int32_t>();
      for (int j = 0; j < cnt; ++j) {
        const uint32_t node = r.get<uint32_t>();
        const int mm = r.get<int32_t>();
        std::vector<unsigned int>& v = fv[node];
        for (int k = 0; k < mm; ++k) v.push_back(r.get<uint32_t>());
      }
    };
    const int nkf = r.get<int32_t>();
    KF.mDescriptors = mat_u8(r.bytes((size_t)nkf * 32), nkf, 32);
    KF.mvKeysUn.resize(nkf);
    std::memcpy(KF.mvKeysUn.data(), r.bytes((size_t)nkf * 28), (size_t)nkf * 28);
    std::vector<ORB_SLAM2::MapPoint> pts(nkf);
    const uint8_t* valid = r.bytes(nkf);
    KF.mvpMapPoints.assign(nkf, nullptr);
    for (int i = 0; i < nkf; ++i) if (valid[i]) KF.mvpMapPoints[i] = &pts[i];
    read_fv(KF.mFeatVec);
    const int nf = r.get<int32_t>();
    F.N = nf;
    F.mDescriptors = mat_u8(r.bytes((size_t)nf * 32), nf, 32);
    F.mvKeys.resize(nf);
    std::memcpy(F.mvKeys.data(), r.bytes((size_t)nf * 28), (size_t)nf * 28);
    read_fv(F.mFeatVec);
    ORB_SLAM2::ORBmatcher matcher(nnratio, ori != 0);
    std::vector<ORB_SLAM2::MapPoint*> matches;
    out.push_back(matcher.SearchByBoW(&KF, F, matches));
    for (int i = 0; i < nf; ++i) out.push_back(matches[i] ? (int32_t)(matches[i] - pts.data()) : -1);
  } else if (mode == "sbow2") {
    // ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (src/ORBmatcher.cc:724-867), the loop-closing matcher.
    // float nnratio; int checkOri; twice {int n; n x 32 B; n x KeyPoint (mvKeysUn); n x u8 "holds a good map point"; FeatureVector}
    // -> int nmatches; n1 x int32 feature of keyframe 2 whose map point vpMatches12[i] is (-1 none)
    const float nnratio = r.get<float>();
    const int ori = r.get<int32_t>();
    ORB_SLAM2::KeyFrame K[2];
    std::vector<ORB_SLAM2::MapPoint> pts[2];
    for (int k = 0; k < 2; ++k) {
      const int n = r.get<int32_t>();
      K[k].mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
      K[k].mvKeysUn.resize(n);
      std::memcpy(K[k].mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
      pts[k].resize(n);
      const uint8_t* valid = r.bytes(n);
      K[k].mvpMapPoints.assign(n, nullptr);
      for (int i = 0; i < n; ++i) if (valid[i]) K[k].mvpMapPoints[i] = &pts[k][i];
      const int cnt = r.get<int32_t>();
      for (int j = 0; j < cnt; ++j) {
        const uint32_t node = r.get<uint32_t>();
        const int mm = r.get<int32_t>();
        std::vector<unsigned int>& v = K[k].mFeatVec[node];
        for (int q = 0; q < mm; ++q) v.push_back(r.get<uint32_t>());
      }
    }
    ORB_SLAM2::ORBmatcher matcher(nnratio, ori != 0);
    std::vector<ORB_SLAM2::MapPoint*> matches;
    out.push_back(matcher.SearchByBoW(&K[0], &K[1], matches));
    for (size_t i = 0; i < matches.size(); ++i) out.push_back(matches[i] ? (int32_t)(matches[i] - pts[1].data()) : -1);
  } else if (mode == "fuse" || mode == "fuse2") {
    // fuse2: Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:1282-1421) with Scw = identity; flag 4 = the point is one of the keyframe's own
    // map points (spAlreadyFound), flag 1 is not used (the reference dereferences every entry).
    // ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:1100-1270) with KeyFrame::GetFeaturesInArea / IsInImage (src/KeyFrame.cc:879-929) on a grid
    // built by Frame::AssignFeaturesToGrid and copied as the KeyFrame constructor does (src/KeyFrame.cc:85-92). Pose = identity, camera centre = 0.
    // float bounds[4], th, mbf, fx, fy, cx, cy; int nlevels; nlevels x float scale; nlevels x float invLevelSigma2; int n; n x KeyPoint; n x 32 B;
    // n x float mvuRight; n x u8 "slot holds a map point"; int m; m x {float X, Y, Z, nz; int level, flags (1 null, 2 bad, 4 IsInKeyFrame)}; m x 32 B
    //   -> int nFused; m x int32 bestIdx of the points that were fused (-1 otherwise)
    ORB_SLAM2::Frame F;
    ORB_SLAM2::KeyFrame KF;
    F.mnMinX = r.get<float>(); F.mnMinY = r.get<float>(); F.mnMaxX = r.get<float>(); F.mnMaxY = r.get<float>();
    const float th = r.get<float>();
    KF.mbf = r.get<float>(); KF.fx = r.get<float>(); KF.fy = r.get<float>(); KF.cx = r.get<float>(); KF.cy = r.get<float>();
    const int nl = r.get<int32_t>();
    for (int l = 0; l < nl; ++l) KF.mvScaleFactors.push_back(r.get<float>());
    for (int l = 0; l < nl; ++l) KF.mvInvLevelSigma2.push_back(r.get<float>());
    F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);     // src/Frame.cc:200-202
    F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
    const int n = r.get<int32_t>();
    F.N = n;
    F.mvKeysUn.resize(n);
    std::memcpy(F.mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
    KF.mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
    KF.mvuRight.resize(n);
    std::memcpy(KF.mvuRight.data(), r.bytes((size_t)n * 4), (size_t)n * 4);
    ORB_SLAM2::MapPoint held;   // whatever a keyframe slot held before the call
    held.nObs = 1;
    KF.mvpMapPoints.assign(n, nullptr);
    const uint8_t* cl = r.bytes(n);
    for (int i = 0; i < n; ++i) if (cl[i]) KF.mvpMapPoints[i] = &held;
    F.AssignFeaturesToGrid();
    KF.N = n; KF.mvKeysUn = F.mvKeysUn;
    KF.mnMinX = F.mnMinX; KF.mnMinY = F.mnMinY; KF.mnMaxX = F.mnMaxX; KF.mnMaxY = F.mnMaxY;
    KF.mfGridElementWidthInv = F.mfGridElementWidthInv; KF.mfGridElementHeightInv = F.mfGridElementHeightInv;
    KF.mGrid.resize(KF.mnGridCols);
    for (int i = 0; i < KF.mnGridCols; ++i) {
      KF.mGrid[i].resize(KF.mnGridRows);
      for (int j = 0; j < KF.mnGridRows; ++j) KF.mGrid[i][j] = F.mGrid[i][j];
    }
    KF.Rcw = cv::Mat(3, 3, CV_32F); KF.tcw = cv::Mat(3, 1, CV_32F); KF.Ow = cv::Mat(3, 1, CV_32F);
    for (int y = 0; y < 3; ++y) { for (int x = 0; x < 3; ++x) KF.Rcw.at<float>(y, x) = y == x ? 1.f : 0.f; KF.tcw.at<float>(y) = 0.f; KF.Ow.at<float>(y) = 0.f; }
    const int m = r.get<int32_t>();
    std::vector<ORB_SLAM2::MapPoint> pts(m);
    std::vector<ORB_SLAM2::MapPoint*> vp(m, nullptr);
    for (int i = 0; i < m; ++i) {
      pts[i].mWorldPos = cv::Mat(3, 1, CV_32F);
      for (int k = 0; k < 3; ++k) pts[i].mWorldPos.at<float>(k) = r.get<float>();
      pts[i].mNormal = cv::Mat(3, 1, CV_32F);
      pts[i].mNormal.at<float>(0) = 0.f; pts[i].mNormal.at<float>(1) = 0.f; pts[i].mNormal.at<float>(2) = r.get<float>();
      pts[i].predLevel = r.get<int32_t>();
      const int fl = r.get<int32_t>();
      if (!(fl & 1)) vp[i] = &pts[i];
      pts[i].bad = (fl & 2) != 0;
      pts[i].inKF = (fl & 4) != 0;
    }
    for (int i = 0; i < m; ++i) pts[i].mDescriptor = mat_u8(r.bytes(32), 1, 32);
    ORB_SLAM2::ORBmatcher matcher(0.6f, true);
    if (mode == "fuse") {
      out.push_back(matcher.Fuse(&KF, vp, th));
    } else {
      for (int i = 0, slot = 0; i < m; ++i)   // points the keyframe already holds sit in slots that hold a map point
        if (pts[i].inKF) { while (slot < n && !cl[slot]) ++slot; if (slot < n) KF.mvpMapPoints[slot++] = &pts[i]; }
      cv::Mat Scw(4, 4, CV_32F);
      for (int y = 0; y < 4; ++y) for (int x = 0; x < 4; ++x) Scw.at<float>(y, x) = y == x ? 1.f : 0.f;
      std::vector<ORB_SLAM2::MapPoint*> repl(m, nullptr);
      out.push_back(matcher.Fuse(&KF, Scw, vp, th, repl));
    }
    for (int i = 0; i < m; ++i) out.push_back(pts[i].fusedIdx);
  } else if (mode == "stri") {
    // ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) + CheckDistEpipolarLine (src/ORBmatcher.cc:879-1088, 206-233).
    // int onlyStereo, checkOri, nlevels; nlevels x float scale; nlevels x float levelSigma2; 9 x float F12 (row-major); 3 x float Cw (camera centre of
    // keyframe 1); 4 x float fx, fy, cx, cy of keyframe 2 (pose of keyframe 2 = identity); twice {int n; n x 32 B; n x KeyPoint (mvKeysUn); n x u8 "holds a
    // map point"; n x float mvuRight; FeatureVector}  ->  int nmatches; n1 x int32 vMatches12 (from vMatchedPairs)
    const int only_stereo = r.get<int32_t>(), ori = r.get<int32_t>(), nl = r.get<int32_t>();
    ORB_SLAM2::KeyFrame K[2];
    for (int l = 0; l < nl; ++l) K[1].mvScaleFactors.push_back(r.get<float>());
    for (int l = 0; l < nl; ++l) K[1].mvLevelSigma2.push_back(r.get<float>());
    cv::Mat F12(3, 3, CV_32F);
    for (int y = 0; y < 3; ++y) for (int x = 0; x < 3; ++x) F12.at<float>(y, x) = r.get<float>();
    K[0].Ow = cv::Mat(3, 1, CV_32F);
    for (int k = 0; k < 3; ++k) K[0].Ow.at<float>(k) = r.get<float>();
    K[1].fx = r.get<float>(); K[1].fy = r.get<float>(); K[1].cx = r.get<float>(); K[1].cy = r.get<float>();
    K[1].Rcw = cv::Mat(3, 3, CV_32F); K[1].tcw = cv::Mat(3, 1, CV_32F);
    for (int y = 0; y < 3; ++y) { for (int x = 0; x < 3; ++x) K[1].Rcw.at<float>(y, x) = y == x ? 1.f : 0.f; K[1].tcw.at<float>(y) = 0.f; }
    ORB_SLAM2::MapPoint held;
    for (int k = 0; k < 2; ++k) {
      const int n = r.get<int32_t>();
      K[k].N = n;
      K[k].mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
      K[k].mvKeysUn.resize(n);
      std::memcpy(K[k].mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
      const uint8_t* has = r.bytes(n);
      K[k].mvpMapPoints.assign(n, nullptr);
      for (int i = 0; i < n; ++i) if (has[i]) K[k].mvpMapPoints[i] = &held;
      K[k].mvuRight.resize(n);
      std::memcpy(K[k].mvuRight.data(), r.bytes((size_t)n * 4), (size_t)n * 4);
      const int cnt = r.get<int32_t>();
      for (int j = 0; j < cnt; ++j) {
        const uint32_t node = r.get<uint32_t>();
        const int mm = r.get<int32_t>();
        std::vector<unsigned int>& v = K[k].mFeatVec[node];
        for (int q = 0; q < mm; ++q) v.push_back(r.get<uint32_t>());
      }
    }
    ORB_SLAM2::g_cur_point = nullptr;
    ORB_SLAM2::ORBmatcher matcher(0.6f, ori != 0);
    std::vector<std::pair<size_t, size_t>> pairs;
    out.push_back(matcher.SearchForTriangulation(&K[0], &K[1], F12, pairs, only_stereo != 0));
    std::vector<int32_t> m12(K[0].N, -1);
    for (const auto& pr : pairs) m12[pr.first] = (int32_t)pr.second;
    out.insert(out.end(), m12.begin(), m12.end());
  } else if (mode == "sbpk" || mode == "sbps") {
    // sbpk: SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1883-2015, relocalisation), CurrentFrame.mTcw = identity;
    // sbps: SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:422-551, loop closing), Scw = identity, target = a KeyFrame (its own GetFeaturesInArea).
    // float bounds[4], th, fx, fy, cx, cy; int ORBdist, checkOri, nlevels; nlevels x float scale; target: int n; n x KeyPoint; n x 32 B; n x u8 "slot
    // holds a map point"; points: int m; m x {float X, Y, Z, angle; int level, flags (1 null [sbpk only], 2 bad, 4 already found)}; m x 32 B
    //   -> int nmatches; n x int32 point assigned to the keypoint by this call (-1 none)
    const bool kf_target = mode == "sbps";
    ORB_SLAM2::Frame F;
    ORB_SLAM2::KeyFrame KF;    // sbps: the target; sbpk: the source of the map points
    F.mnMinX = r.get<float>(); F.mnMinY = r.get<float>(); F.mnMaxX = r.get<float>(); F.mnMaxY = r.get<float>();
    const float th = r.get<float>();
    F.fx = r.get<float>(); F.fy = r.get<float>(); F.cx = r.get<float>(); F.cy = r.get<float>();
    const int orbdist = r.get<int32_t>(), ori = r.get<int32_t>(), nl = r.get<int32_t>();
    for (int l = 0; l < nl; ++l) F.mvScaleFactors.push_back(r.get<float>());
    F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);     // src/Frame.cc:200-202
    F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
    const int n = r.get<int32_t>();
    F.N = n;
    F.mvKeysUn.resize(n);
    std::memcpy(F.mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
    F.mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
    ORB_SLAM2::MapPoint taken;
    taken.nObs = 1;
    F.mvpMapPoints.assign(n, nullptr);
    const uint8_t* cl = r.bytes(n);
    for (int i = 0; i < n; ++i) if (cl[i]) F.mvpMapPoints[i] = &taken;
    F.AssignFeaturesToGrid();
    cv::Mat I4(4, 4, CV_32F);
    for (int y = 0; y < 4; ++y) for (int x = 0; x < 4; ++x) I4.at<float>(y, x) = y == x ? 1.f : 0.f;
    F.mTcw = I4;
    const int m = r.get<int32_t>();
    std::vector<ORB_SLAM2::MapPoint> pts(m);
    std::vector<ORB_SLAM2::MapPoint*> vp(m, nullptr);
    std::vector<float> angle(m);
    std::set<ORB_SLAM2::MapPoint*> found;
    for (int i = 0; i < m; ++i) {
      pts[i].mWorldPos = cv::Mat(3, 1, CV_32F);
      for (int k = 0; k < 3; ++k) pts[i].mWorldPos.at<float>(k) = r.get<float>();
      pts[i].mNormal = cv::Mat(3, 1, CV_32F);
      pts[i].mNormal.at<float>(0) = 0.f; pts[i].mNormal.at<float>(1) = 0.f; pts[i].mNormal.at<float>(2) = 1.f;
      angle[i] = r.get<float>();
      pts[i].predLevel = r.get<int32_t>();
      const int fl = r.get<int32_t>();
      if (!(fl & 1)) vp[i] = &pts[i];
      pts[i].bad = (fl & 2) != 0;
      if (fl & 4) found.insert(&pts[i]);
    }
    for (int i = 0; i < m; ++i) pts[i].mDescriptor = mat_u8(r.bytes(32), 1, 32);
    ORB_SLAM2::ORBmatcher matcher(0.6f, ori != 0);
    std::vector<ORB_SLAM2::MapPoint*> result;
    if (!kf_target) {
      KF.mvpMapPoints = vp;
      KF.mvKeysUn.resize(m);
      for (int i = 0; i < m; ++i) KF.mvKeysUn[i].angle = angle[i];
      out.push_back(matcher.SearchByProjection(F, &KF, found, th, orbdist));
      result = F.mvpMapPoints;
    } else {
      KF.N = n; KF.mvKeysUn = F.mvKeysUn; KF.mDescriptors = F.mDescriptors;
      KF.fx = F.fx; KF.fy = F.fy; KF.cx = F.cx; KF.cy = F.cy;
      KF.mvScaleFactors = F.mvScaleFactors;
      KF.mnMinX = F.mnMinX; KF.mnMinY = F.mnMinY; KF.mnMaxX = F.mnMaxX; KF.mnMaxY = F.mnMaxY;
      KF.mfGridElementWidthInv = F.mfGridElementWidthInv; KF.mfGridElementHeightInv = F.mfGridElementHeightInv;
      KF.mGrid.resize(KF.mnGridCols);
      for (int i = 0; i < KF.mnGridCols; ++i) {
        KF.mGrid[i].resize(KF.mnGridRows);
        for (int j = 0; j < KF.mnGridRows; ++j) KF.mGrid[i][j] = F.mGrid[i][j];
      }
      result = F.mvpMapPoints;                      // vpMatched: slots that already hold a match
      int slot = 0;
      for (int i = 0; i < m; ++i)                   // "already found" points are the ones sitting in vpMatched (:434-435)
        if (found.count(&pts[i])) { while (slot < n && !cl[slot]) ++slot; if (slot < n) result[slot++] = &pts[i]; }
      out.push_back(matcher.SearchByProjection(&KF, I4, vp, result, (int)th));
    }
    for (int i = 0; i < n; ++i) {
      ORB_SLAM2::MapPoint* p = result[i];
      out.push_back(p && p != &taken && !(cl[i]) ? (int32_t)(p - pts.data()) : -1);
    }
  } else if (mode == "sim3") {
    // ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1433-1684) with identity poses and s12 = 1, R12 = I, t12 = 0.
    // float bounds[4], th, fx, fy, cx, cy; int nlevels; nlevels x float scale; twice {int n; n x KeyPoint; n x 32 B; n x {float X, Y, Z; int level,
    // flags (1 no map point, 2 bad), prematch (keyframe 1: feature of keyframe 2 it is already matched to, -1 none; keyframe 2: unused)}}
    //   -> int nFound; n1 x int32 feature of keyframe 2 newly matched to feature i1 (-1 none)
    float bounds[4];
    for (float& b : bounds) b = r.get<float>();
    const float th = r.get<float>();
    float cam[4];
    for (float& c : cam) c = r.get<float>();
    const int nl = r.get<int32_t>();
    std::vector<float> scale(nl);
    for (float& v : scale) v = r.get<float>();
    ORB_SLAM2::KeyFrame K[2];
    std::vector<ORB_SLAM2::MapPoint> pts[2];
    std::vector<int> prematch;
    for (int k = 0; k < 2; ++k) {
      ORB_SLAM2::Frame F;
      F.mnMinX = bounds[0]; F.mnMinY = bounds[1]; F.mnMaxX = bounds[2]; F.mnMaxY = bounds[3];
      F.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F.mnMaxX - F.mnMinX);
      F.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F.mnMaxY - F.mnMinY);
      const int n = r.get<int32_t>();
      F.N = n;
      F.mvKeysUn.resize(n);
      std::memcpy(F.mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
      K[k].mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
      F.AssignFeaturesToGrid();
      ORB_SLAM2::KeyFrame& KF = K[k];
      KF.N = n; KF.mvKeysUn = F.mvKeysUn;
      KF.fx = cam[0]; KF.fy = cam[1]; KF.cx = cam[2]; KF.cy = cam[3];
      KF.mvScaleFactors = scale;
      KF.mnMinX = F.mnMinX; KF.mnMinY = F.mnMinY; KF.mnMaxX = F.mnMaxX; KF.mnMaxY = F.mnMaxY;
      KF.mfGridElementWidthInv = F.mfGridElementWidthInv; KF.mfGridElementHeightInv = F.mfGridElementHeightInv;
      KF.mGrid.resize(KF.mnGridCols);
      for (int i = 0; i < KF.mnGridCols; ++i) {
        KF.mGrid[i].resize(KF.mnGridRows);
        for (int j = 0; j < KF.mnGridRows; ++j) KF.mGrid[i][j] = F.mGrid[i][j];
      }
      KF.Rcw = cv::Mat(3, 3, CV_32F); KF.tcw = cv::Mat(3, 1, CV_32F);
      for (int y = 0; y < 3; ++y) { for (int x = 0; x < 3; ++x) KF.Rcw.at<float>(y, x) = y == x ? 1.f : 0.f; KF.tcw.at<float>(y) = 0.f; }
      pts[k].resize(n);
      KF.mvpMapPoints.assign(n, nullptr);
      for (int i = 0; i < n; ++i) {
        pts[k][i].mWorldPos = cv::Mat(3, 1, CV_32F);
        for (int c = 0; c < 3; ++c) pts[k][i].mWorldPos.at<float>(c) = r.get<float>();
        pts[k][i].predLevel = r.get<int32_t>();
        const int fl = r.get<int32_t>();
        const int pm = r.get<int32_t>();
        if (!(fl & 1)) KF.mvpMapPoints[i] = &pts[k][i];
        pts[k][i].bad = (fl & 2) != 0;
        if (k == 0) prematch.push_back(pm);
      }
    }
    for (int i = 0; i < K[0].N; ++i) pts[0][i].mDescriptor = K[0].mDescriptors.row(i);   // a map point's descriptor = its feature's here
    for (int i = 0; i < K[1].N; ++i) pts[1][i].mDescriptor = K[1].mDescriptors.row(i);
    std::vector<ORB_SLAM2::MapPoint*> m12(K[0].N, nullptr);
    for (int i = 0; i < K[0].N; ++i)
      if (prematch[i] >= 0) { m12[i] = &pts[1][prematch[i]]; pts[1][prematch[i]].idxInKF2 = prematch[i]; }
    cv::Mat R12(3, 3, CV_32F), t12(3, 1, CV_32F);
    for (int y = 0; y < 3; ++y) { for (int x = 0; x < 3; ++x) R12.at<float>(y, x) = y == x ? 1.f : 0.f; t12.at<float>(y) = 0.f; }
    ORB_SLAM2::ORBmatcher matcher(0.6f, true);
    const float s12 = 1.f;
    out.push_back(matcher.SearchBySim3(&K[0], &K[1], m12, s12, R12, t12, th));
    for (int i = 0; i < K[0].N; ++i) out.push_back(m12[i] && prematch[i] < 0 ? (int32_t)(m12[i] - pts[1].data()) : -1);
  } else if (mode == "sinit") {
    // float bounds[4], nnratio; int checkOri, windowSize; F2: int n; n x KeyPoint (mvKeysUn); n x 32 B; F1: int m; m x KeyPoint (mvKeysUn: octave,
    // angle); m x 32 B; m x {float x, y} vbPrevMatched  ->  int nmatches; m x int32 vnMatches12
    ORB_SLAM2::Frame F1, F2;
    F2.mnMinX = r.get<float>(); F2.mnMinY = r.get<float>(); F2.mnMaxX = r.get<float>(); F2.mnMaxY = r.get<float>();
    const float nnratio = r.get<float>();
    const int ori = r.get<int32_t>(), window = r.get<int32_t>();
    F2.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(F2.mnMaxX - F2.mnMinX);   // src/Frame.cc:200-202
    F2.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(F2.mnMaxY - F2.mnMinY);
    const int n = r.get<int32_t>();
    F2.N = n;
    F2.mvKeysUn.resize(n);
    std::memcpy(F2.mvKeysUn.data(), r.bytes((size_t)n * 28), (size_t)n * 28);
    F2.mDescriptors = mat_u8(r.bytes((size_t)n * 32), n, 32);
    F2.AssignFeaturesToGrid();
    const int m = r.get<int32_t>();
    F1.N = m;
    F1.mvKeysUn.resize(m);
    std::memcpy(F1.mvKeysUn.data(), r.bytes((size_t)m * 28), (size_t)m * 28);
    F1.mDescriptors = mat_u8(r.bytes((size_t)m * 32), m, 32);
    std::vector<cv::Point2f> prev(m);
    for (int i = 0; i < m; ++i) { prev[i].x = r.get<float>(); prev[i].y = r.get<float>(); }
    ORB_SLAM2::ORBmatcher matcher(nnratio, ori != 0);
    std::vector<int> m12;
    out.push_back(matcher.SearchForInitialization(F1, F2, prev, m12, window));
    for (int i = 0; i < m; ++i) out.push_back(m12[i]);
  } else {
    std::fprintf(stderr, "unknown mode %s\n", mode.c_str());
    return 1;
  }
  FILE* fo = std::fopen(argv[3], "wb");
  if (!fo) { std::perror(argv[3]); return 1; }
  if (!out.empty()) std::fwrite(out.data(), 4, out.size(), fo);
  if (!outf.empty()) std::fwrite(outf.data(), 4, outf.size(), fo);
  std::fclose(fo);
  return 0;
}
