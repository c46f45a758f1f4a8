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
#include "cvshim/cvshim.h"

using namespace std;   // the reference's sources say `vector`, `sort`, `pair` unqualified (src/ORBmatcher.cc:43)

namespace ORB_SLAM2 {

class ORBextractor {   // include/ORBextractor.h:185: the one member ComputeStereoMatches reads
 public:
  std::vector<cv::Mat> mvImagePyramid;
};

class ORBmatcher {     // include/ORBmatcher.h:54-62, 223-236
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
  static const int TH_LOW;
  static const int TH_HIGH;
  static const int HISTO_LENGTH;
  void ComputeThreeMaxima(vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
  // harness entry points around the extracted loop bodies (declared here so the bodies see the members unqualified)
  void Scan(const cv::Mat& Q, const cv::Mat& D, int32_t* out);
  void ScanInit(const cv::Mat& Q, const cv::Mat& D, const std::vector<int>& matched_distance, int32_t* out);
  float mfNNratio;
  bool mbCheckOrientation;
};

class Frame {          // include/Frame.h: the members src/Frame.cc:880-1176 names
 public:
  void ComputeStereoMatches();
  int N;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
  cv::Mat mDescriptors, mDescriptorsRight;
  ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
  std::vector<float> mvScaleFactors, mvInvScaleFactors;
  std::vector<float> mvuRight, mvDepth;
  float mb, mbf;
};

#include "_ref/gen_consts.inc"
#include "_ref/gen_distance.inc"
#include "_ref/gen_three_maxima.inc"
#include "_ref/gen_stereo.inc"

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
  if (argc != 4) { std::fprintf(stderr, "usage: %s distance|scan|scan_init|three_maxima|distinctive|stereo IN OUT\n", argv[0]); return 1; }
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
