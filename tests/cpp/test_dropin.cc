// C++ parity test of the drop-in classes (built against cv_shim.h): ORB_SLAM2::ORBextractor / ORBmatcher on the GPU versus
// the CPU oracle (oracle/orb_oracle.h, test infrastructure). Run by tests/test_cpp_dropin.py on a B200; exit code 0 = pass.
//   usage: test_dropin <frame.raw> <width> <height> <nfeatures> [<frame2.raw>]
#include <algorithm>
#include <cmath>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../oracle/orb_oracle.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary.h"
#include "FrameOps.h"

#define REQUIRE(cond, ...)                      \
  do {                                          \
    if (!(cond)) {                              \
      std::fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); \
      std::fprintf(stderr, __VA_ARGS__);        \
      std::fprintf(stderr, "\n");               \
      return 1;                                 \
    }                                           \
  } while (0)

static bool read_raw(const char* path, std::vector<unsigned char>& buf, size_t n) {
  buf.resize(n);
  FILE* f = std::fopen(path, "rb");
  if (!f) return false;
  const size_t got = std::fread(buf.data(), 1, n, f);
  std::fclose(f);
  return got == n;
}

static int compare_with_oracle(orb_oracle_extractor* oe, const cv::Mat& img, const std::vector<cv::KeyPoint>& kps, const cv::Mat& desc,
                               ORB_SLAM2::ORBextractor& ex, bool check_pyramid) {
  const int cap = 1 << 16;
  std::vector<orb_oracle_kp> okp(cap);
  std::vector<uint8_t> odesc((size_t)cap * 32);
  const int n = orb_oracle_extract(oe, img.data, (int)img.step, img.cols, img.rows, okp.data(), odesc.data(), cap, nullptr);
  REQUIRE(n == (int)kps.size(), "keypoint count %zu vs oracle %d", kps.size(), n);
  REQUIRE(n > 100, "too few keypoints (%d) for a meaningful test", n);
  REQUIRE(desc.rows == n && desc.cols == 32, "descriptor shape %dx%d", desc.rows, desc.cols);
  REQUIRE(std::memcmp(kps.data(), okp.data(), (size_t)n * 28) == 0, "cv::KeyPoint records differ from the oracle");
  for (int r = 0; r < n; ++r) REQUIRE(std::memcmp(desc.ptr<unsigned char>(r), &odesc[(size_t)r * 32], 32) == 0, "descriptor row %d differs", r);
  if (check_pyramid) {
    for (int l = 0; l < ex.GetLevels(); ++l) {
      int c = 0, r = 0;
      orb_oracle_level_dims(oe, l, &c, &r);
      const cv::Mat& m = ex.mvImagePyramid[l];
      REQUIRE(m.cols == c && m.rows == r, "mvImagePyramid[%d] is %dx%d, oracle %dx%d", l, m.cols, m.rows, c, r);
      const uint8_t* op = orb_oracle_level_ptr(oe, l);
      for (int y = 0; y < r; ++y) REQUIRE(std::memcmp(m.ptr<unsigned char>(y), op + (size_t)y * c, c) == 0, "mvImagePyramid[%d] row %d differs", l, y);
    }
  }
  return 0;
}

// --latency frame.raw W H nfeatures [calls]: wall clock of ORBextractor::operator() the way Frame::ExtractORB calls it (src/Frame.cc:468-481: one
// pageable cv::Mat in, std::vector<cv::KeyPoint> + cv::Mat out), with mvImagePyramid filled (the drop-in's default, what the stereo Frame needs)
// and without. One JSON line; no oracle involved.
static int latency_mode(int argc, char** argv) {
  if (argc < 6) { std::fprintf(stderr, "usage: %s --latency frame.raw width height nfeatures [calls]\n", argv[0]); return 2; }
  const int W = std::atoi(argv[3]), H = std::atoi(argv[4]), nf = std::atoi(argv[5]);
  const int calls = argc > 6 ? std::atoi(argv[6]) : 300;
  std::vector<unsigned char> f0;
  REQUIRE(read_raw(argv[2], f0, (size_t)W * H), "cannot read %s", argv[2]);
  // 8 shifted copies so that consecutive calls do not see identical pixels
  std::vector<std::vector<unsigned char>> frames(8, f0);
  for (int k = 1; k < 8; ++k) std::rotate(frames[k].begin(), frames[k].begin() + 131 * k, frames[k].end());
  std::printf("{\"metric\": \"wall clock of one C++ ORB_SLAM2::ORBextractor::operator() call (cpp/ORBextractor.cc over the C ABI; 1 pageable image in, keypoints + descriptors out)\", \"unit\": \"ms\", \"width\": %d, \"height\": %d, \"nfeatures\": %d, \"calls\": %d", W, H, nf, calls);
  for (int mode = 0; mode < 2; ++mode) {
    ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
    ex.SetPyramidDownload(mode == 0);
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc;
    size_t nk = 0;
    for (int i = 0; i < 30; ++i) { cv::Mat img(H, W, CV_8UC1, frames[i & 7].data(), (size_t)W); ex(img, cv::Mat(), kps, desc); }
    std::vector<double> t(calls);
    for (int i = 0; i < calls; ++i) {
      cv::Mat img(H, W, CV_8UC1, frames[i & 7].data(), (size_t)W);
      const auto t0 = std::chrono::steady_clock::now();
      ex(img, cv::Mat(), kps, desc);
      t[i] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      nk += kps.size();
    }
    std::sort(t.begin(), t.end());
    std::printf(", \"%s\": {\"median\": %.4f, \"p99\": %.4f, \"min\": %.4f, \"keypoints_per_call\": %.1f}", mode == 0 ? "with_pyramid_download" : "without_pyramid_download",
                t[calls / 2], t[(size_t)(calls * 0.99)], t[0], (double)nk / calls);
  }
  std::printf("}\n");
  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1 && std::strcmp(argv[1], "--latency") == 0) return latency_mode(argc, argv);
  if (argc < 5) {
    std::fprintf(stderr, "usage: %s frame.raw width height nfeatures [frame2.raw]\n", argv[0]);
    return 2;
  }
  const int W = std::atoi(argv[2]), H = std::atoi(argv[3]), nf = std::atoi(argv[4]);
  std::vector<unsigned char> f0, f1;
  REQUIRE(read_raw(argv[1], f0, (size_t)W * H), "cannot read %s", argv[1]);
  const bool two = argc > 5;
  if (two) REQUIRE(read_raw(argv[5], f1, (size_t)W * H), "cannot read %s", argv[5]);

  // ---- constructor tables and getters (include/ORBextractor.h:134-182)
  ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
  orb_oracle_extractor* oe = orb_oracle_create(nf, 1.2f, 8, 20, 7, 0);
  REQUIRE(ex.GetLevels() == 8, "GetLevels");
  REQUIRE(ex.GetScaleFactor() == 1.2f, "GetScaleFactor");
  for (int what = 0; what < 4; ++what) {
    float t[8];
    orb_oracle_get_table(oe, what, t);
    const std::vector<float> got = what == 0 ? ex.GetScaleFactors() : what == 1 ? ex.GetInverseScaleFactors()
                                   : what == 2 ? ex.GetScaleSigmaSquares() : ex.GetInverseScaleSigmaSquares();
    REQUIRE(got.size() == 8 && std::memcmp(got.data(), t, sizeof(t)) == 0, "table %d differs from the oracle", what);
  }

  // ---- operator(): strided input (ROI of a wider buffer), empty mask, like Frame::ExtractORB (src/Frame.cc:468-481)
  std::vector<unsigned char> wide((size_t)(W + 13) * H, 0);
  for (int y = 0; y < H; ++y) std::memcpy(&wide[(size_t)y * (W + 13)], &f0[(size_t)y * W], W);
  cv::Mat img(H, W, CV_8UC1, wide.data(), (size_t)W + 13);
  std::vector<cv::KeyPoint> kps;
  cv::Mat desc;
  ex(img, cv::Mat(), kps, desc);
  cv::Mat contiguous(H, W, CV_8UC1, f0.data());
  if (compare_with_oracle(oe, contiguous, kps, desc, ex, true)) return 1;
  std::printf("operator(): %zu keypoints, descriptors and mvImagePyramid bit-exact vs oracle\n", kps.size());

  // empty image: outputs untouched (src/ORBextractor.cc:1966-1967)
  const size_t before = kps.size();
  ex(cv::Mat(), cv::Mat(), kps, desc);
  REQUIRE(kps.size() == before && desc.rows == (int)before, "empty image must leave the outputs untouched");

  // ---- batched stereo pair + brute-force matching left -> right
  if (two) {
    cv::Mat right(H, W, CV_8UC1, f1.data());
    std::vector<cv::Mat> pair{contiguous, right};
    std::vector<std::vector<cv::KeyPoint>> bk;
    std::vector<cv::Mat> bd;
    ex.ExtractBatch(pair, bk, bd);
    if (compare_with_oracle(oe, right, bk[1], bd[1], ex, true)) return 1;   // oracle state now = right image
    if (compare_with_oracle(oe, contiguous, bk[0], bd[0], ex, false)) return 1;
    std::printf("ExtractBatch: %zu + %zu keypoints bit-exact vs oracle\n", bk[0].size(), bk[1].size());

    std::vector<ORB_SLAM2::ORBmatcher::Knn2> knn;
    ORB_SLAM2::ORBmatcher::BruteForceKnn2(bd[0], bd[1], knn);
    std::vector<orb_oracle_knn2> want(bd[0].rows);
    orb_oracle_hamming_knn(bd[0].ptr<unsigned char>(0), bd[0].rows, bd[1].ptr<unsigned char>(0), bd[1].rows, 0, 256, want.data(), 4);
    for (int i = 0; i < bd[0].rows; ++i)
      REQUIRE(knn[i].bestDist == want[i].d1 && knn[i].bestDist2 == want[i].d2 && knn[i].bestIdx == want[i].i1 && knn[i].bestIdx2 == want[i].i2,
              "kNN row %d: (%d,%d,%lld,%lld) vs oracle (%d,%d,%lld,%lld)", i, knn[i].bestDist, knn[i].bestDist2, (long long)knn[i].bestIdx,
              (long long)knn[i].bestIdx2, want[i].d1, want[i].d2, (long long)want[i].i1, (long long)want[i].i2);
    ORB_SLAM2::ORBmatcher matcher(0.6f, true);
    std::vector<int64_t> matches;
    const int nacc = matcher.MatchBruteForce(bd[0], bd[1], matches);
    int want_acc = 0;
    for (int i = 0; i < bd[0].rows; ++i)
      want_acc += want[i].d1 <= ORB_SLAM2::ORBmatcher::TH_LOW && (float)want[i].d1 < 0.6f * (float)want[i].d2;
    REQUIRE(nacc == want_acc, "accepted matches %d vs oracle %d", nacc, want_acc);
    // DescriptorDistance: scalar form, batch form and oracle agree on the matched pairs
    std::vector<int> dist;
    const int m = bd[0].rows < bd[1].rows ? bd[0].rows : bd[1].rows;
    cv::Mat a(m, 32, CV_8U, bd[0].ptr<unsigned char>(0)), b(m, 32, CV_8U, bd[1].ptr<unsigned char>(0));
    ORB_SLAM2::ORBmatcher::DescriptorDistanceBatch(a, b, dist);
    for (int i = 0; i < m; ++i) {
      const int o = orb_oracle_descriptor_distance(a.ptr<unsigned char>(i), b.ptr<unsigned char>(i));
      REQUIRE(dist[i] == o && ORB_SLAM2::ORBmatcher::DescriptorDistance(a.row(i), b.row(i)) == o, "DescriptorDistance row %d", i);
    }
    // two extractors driven from two std::threads, like the stereo Frame constructor (src/Frame.cc:158-168)
    {
      ORB_SLAM2::ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
      std::vector<cv::KeyPoint> kl, kr;
      cv::Mat dl, dr;
      bool same = true;
      for (int round = 0; round < 5 && same; ++round) {
        std::thread tl([&]() { exL(contiguous, cv::Mat(), kl, dl); });
        std::thread tr([&]() { exR(right, cv::Mat(), kr, dr); });
        tl.join();
        tr.join();
        same = kl.size() == bk[0].size() && kr.size() == bk[1].size() && dl.rows == bd[0].rows && dr.rows == bd[1].rows &&
               std::memcmp(dl.ptr<unsigned char>(0), bd[0].ptr<unsigned char>(0), (size_t)dl.rows * 32) == 0 &&
               std::memcmp(dr.ptr<unsigned char>(0), bd[1].ptr<unsigned char>(0), (size_t)dr.rows * 32) == 0;
        for (size_t i = 0; same && i < kl.size(); ++i)
          same = kl[i].pt.x == bk[0][i].pt.x && kl[i].pt.y == bk[0][i].pt.y && kl[i].angle == bk[0][i].angle && kl[i].octave == bk[0][i].octave;
        for (size_t i = 0; same && i < kr.size(); ++i)
          same = kr[i].pt.x == bk[1][i].pt.x && kr[i].pt.y == bk[1][i].pt.y && kr[i].angle == bk[1][i].angle && kr[i].octave == bk[1][i].octave;
      }
      REQUIRE(same, "left / right extractors on two threads differ from the batched call");
      std::printf("two extractors on two std::threads: equal to the batched call\n");
    }
    // stereo Frame constructor in one call: keypoints as above, mvuRight / mvDepth equal to the oracle's ComputeStereoMatches
    {
      std::vector<cv::KeyPoint> kl, kr;
      cv::Mat dl, dr;
      std::vector<float> ur, dp;
      const float bf = 386.1448f, mb = 0.5372f;
      ex.ExtractStereo(contiguous, right, bf, mb, kl, dl, kr, dr, ur, dp);
      orb_oracle_extractor* oR = orb_oracle_create(nf, 1.2f, 8, 20, 7, 0);
      const int capo = 1 << 16;
      std::vector<orb_oracle_kp> okl(capo), okr(capo);
      std::vector<uint8_t> odl((size_t)capo * 32), odr((size_t)capo * 32);
      const int nl = orb_oracle_extract(oe, contiguous.data, (int)contiguous.step, W, H, okl.data(), odl.data(), capo, nullptr);
      const int nr = orb_oracle_extract(oR, right.data, (int)right.step, W, H, okr.data(), odr.data(), capo, nullptr);
      REQUIRE(nl == (int)kl.size() && nr == (int)kr.size() && (int)ur.size() == nl && (int)dp.size() == nl, "ExtractStereo sizes");
      REQUIRE(std::memcmp(kl.data(), okl.data(), (size_t)nl * 28) == 0 && std::memcmp(kr.data(), okr.data(), (size_t)nr * 28) == 0, "ExtractStereo keypoints");
      std::vector<float> our(nl), odp(nl);
      const int kept = orb_oracle_stereo_match(oe, oR, okl.data(), odl.data(), nl, okr.data(), odr.data(), nr, bf, mb, our.data(), odp.data());
      REQUIRE(std::memcmp(ur.data(), our.data(), (size_t)nl * 4) == 0, "mvuRight differs from the oracle");
      REQUIRE(std::memcmp(dp.data(), odp.data(), (size_t)nl * 4) == 0, "mvDepth differs from the oracle");
      REQUIRE(kept > 50, "too few stereo matches (%d)", kept);
      orb_oracle_destroy(oR);
      std::printf("ExtractStereo: %d stereo matches, mvuRight / mvDepth bit-exact vs oracle\n", kept);
    }
    // guided search: the left keypoints projected into the right image (constant shift unknown to the matcher: window 40 px),
    // frame-to-frame form with rotation histogram, then the local-map form with the ratio test
    {
      std::vector<ORB_SLAM2::ORBmatcher::ProjectedPoint> pts(bk[0].size());
      for (size_t i = 0; i < pts.size(); ++i) {
        const cv::KeyPoint& k = bk[0][i];
        pts[i] = {k.pt.x, k.pt.y, 40.f, -1.f, k.angle, k.octave - 1, k.octave + 1, (i % 17) == 3, (i % 5) != 0};
      }
      std::vector<unsigned char> claimed(bk[1].size(), 0);
      for (size_t i = 0; i < claimed.size(); i += 11) claimed[i] = 1;
      ORB_SLAM2::ORBmatcher::TargetFrame F{&bk[1], &bd[1], nullptr, &claimed, 0.f, 0.f, (float)W, (float)H};
      std::vector<orb_oracle_query> oq(pts.size());
      for (size_t i = 0; i < pts.size(); ++i) {
        oq[i] = {pts[i].u, pts[i].v, pts[i].radius, pts[i].uRight, pts[i].angle, (int16_t)pts[i].minLevel, (int16_t)pts[i].maxLevel,
                 (pts[i].skip ? 1 : 0) | (pts[i].claims ? 2 : 0), 0};
      }
      for (int form = 0; form < 2; ++form) {
        ORB_SLAM2::ORBmatcher gm(0.8f, true);
        std::vector<int> assigned, pm;
        const int nm = gm.SearchByProjection(F, pts, bd[0], ORB_SLAM2::ORBmatcher::TH_HIGH, form == 1, assigned, &pm);
        orb_oracle_search_params op{0.f, 0.f, (float)W, (float)H, ORB_SLAM2::ORBmatcher::TH_HIGH, form == 1 ? 0.8f : 0.f, form == 0 ? 1 : 0, 0};
        std::vector<int32_t> oa(bk[1].size()), oqm(pts.size());
        const int onm = orb_oracle_search_by_projection(reinterpret_cast<const orb_oracle_kp*>(bk[1].data()), bd[1].ptr<unsigned char>(0), nullptr,
                                                        claimed.data(), (int)bk[1].size(), oq.data(), bd[0].ptr<unsigned char>(0), (int)pts.size(),
                                                        &op, oa.data(), oqm.data());
        REQUIRE(nm == onm, "SearchByProjection form %d: nmatches %d vs oracle %d", form, nm, onm);
        REQUIRE(std::memcmp(assigned.data(), oa.data(), oa.size() * 4) == 0 && std::memcmp(pm.data(), oqm.data(), oqm.size() * 4) == 0,
                "SearchByProjection form %d: assignment differs from the oracle", form);
        REQUIRE(nm > 50, "SearchByProjection form %d: too few matches (%d)", form, nm);
        std::printf("SearchByProjection (%s): %d matches, assignment bit-exact vs oracle\n", form ? "local-map form" : "frame-to-frame form", nm);
      }
    }
    // Fuse candidate search (with and without the chi-square test) and SearchBySim3's mutual check on the same two feature sets
    {
      typedef ORB_SLAM2::ORBmatcher::ProjectedPoint PP;
      std::vector<PP> p12(bk[0].size()), p21(bk[1].size());
      for (size_t i = 0; i < p12.size(); ++i) { const cv::KeyPoint& k = bk[0][i]; p12[i] = {k.pt.x, k.pt.y, 40.f, k.pt.x - 12.f, k.angle, k.octave - 1, k.octave, (i % 13) == 5, false}; }
      for (size_t i = 0; i < p21.size(); ++i) { const cv::KeyPoint& k = bk[1][i]; p21[i] = {k.pt.x, k.pt.y, 40.f, k.pt.x - 12.f, k.angle, k.octave - 1, k.octave, (i % 19) == 2, false}; }
      std::vector<float> ur2(bk[1].size(), -1.f), inv(8);
      for (size_t i = 0; i < ur2.size(); i += 3) ur2[i] = bk[1][i].pt.x - 10.f;
      for (int l = 0; l < 8; ++l) inv[l] = 1.f / (400.f * (float)std::pow(1.44, l));   // loose table: the synthetic pair is not registered
      ORB_SLAM2::ORBmatcher::TargetFrame K1{&bk[0], &bd[0], nullptr, nullptr, 0.f, 0.f, (float)W, (float)H};
      ORB_SLAM2::ORBmatcher::TargetFrame K2{&bk[1], &bd[1], &ur2, nullptr, 0.f, 0.f, (float)W, (float)H};
      auto to_oq = [](const std::vector<PP>& p) {
        std::vector<orb_oracle_query> q(p.size());
        for (size_t i = 0; i < p.size(); ++i) q[i] = {p[i].u, p[i].v, p[i].radius, p[i].uRight, p[i].angle, (int16_t)p[i].minLevel, (int16_t)p[i].maxLevel, p[i].skip ? 1 : 0, 0};
        return q;
      };
      const std::vector<orb_oracle_query> q12 = to_oq(p12), q21 = to_oq(p21);
      ORB_SLAM2::ORBmatcher fm(0.6f, true);
      float sig16[16] = {0.f};
      std::copy(inv.begin(), inv.end(), sig16);
      for (int chi = 0; chi < 2; ++chi) {
        std::vector<int> bi, bdist;
        fm.FuseSearch(K2, p12, bd[0], chi ? &inv : nullptr, ORB_SLAM2::ORBmatcher::TH_LOW, bi, &bdist);
        orb_oracle_search_params op{0.f, 0.f, (float)W, (float)H, ORB_SLAM2::ORBmatcher::TH_LOW, 0.f, 0, 0};
        std::vector<int32_t> wi(p12.size()), wd(p12.size());
        orb_oracle_fuse_search(reinterpret_cast<const orb_oracle_kp*>(bk[1].data()), bd[1].ptr<unsigned char>(0), ur2.data(), (int)bk[1].size(), q12.data(),
                               bd[0].ptr<unsigned char>(0), (int)q12.size(), &op, chi ? sig16 : nullptr, wi.data(), wd.data());
        int nf = 0;
        for (int v : bi) nf += v >= 0;
        REQUIRE(std::memcmp(bi.data(), wi.data(), wi.size() * 4) == 0 && std::memcmp(bdist.data(), wd.data(), wd.size() * 4) == 0,
                "FuseSearch (chi-square %d) differs from the oracle", chi);
        REQUIRE(nf > 20, "FuseSearch: too few fused points (%d)", nf);
        std::printf("FuseSearch (%s): %d points with a fuse candidate, bestIdx / bestDist bit-exact vs oracle\n", chi ? "chi-square test" : "no chi-square test", nf);
      }
      std::vector<int> m12;
      const int nfound = fm.SearchBySim3(K1, K2, p12, bd[0], p21, bd[1], m12);
      orb_oracle_search_params oh{0.f, 0.f, (float)W, (float)H, ORB_SLAM2::ORBmatcher::TH_HIGH, 0.f, 0, 0};
      std::vector<int32_t> v1(p12.size()), v2(p21.size()), dd(std::max(p12.size(), p21.size()));
      orb_oracle_fuse_search(reinterpret_cast<const orb_oracle_kp*>(bk[1].data()), bd[1].ptr<unsigned char>(0), ur2.data(), (int)bk[1].size(), q12.data(),
                             bd[0].ptr<unsigned char>(0), (int)q12.size(), &oh, nullptr, v1.data(), dd.data());
      orb_oracle_fuse_search(reinterpret_cast<const orb_oracle_kp*>(bk[0].data()), bd[0].ptr<unsigned char>(0), nullptr, (int)bk[0].size(), q21.data(),
                             bd[1].ptr<unsigned char>(0), (int)q21.size(), &oh, nullptr, v2.data(), dd.data());
      int want_found = 0;
      bool same = true;
      for (size_t i1 = 0; i1 < v1.size(); ++i1) {
        const bool mutual = v1[i1] >= 0 && v2[v1[i1]] == (int)i1;
        want_found += mutual;
        same = same && m12[i1] == (mutual ? v1[i1] : -1);
      }
      REQUIRE(same && nfound == want_found, "SearchBySim3: %d mutual matches vs oracle %d", nfound, want_found);
      REQUIRE(nfound > 20, "SearchBySim3: too few mutual matches (%d)", nfound);
      std::printf("SearchBySim3: %d mutual matches bit-exact vs oracle\n", nfound);
    }
    // SearchForInitialization: left keypoints search a 100-px window around their own position in the right image
    {
      std::vector<cv::Point2f> prev(bk[0].size());
      for (size_t i = 0; i < prev.size(); ++i) prev[i] = bk[0][i].pt;
      ORB_SLAM2::ORBmatcher::TargetFrame F2{&bk[1], &bd[1], nullptr, nullptr, 0.f, 0.f, (float)W, (float)H};
      ORB_SLAM2::ORBmatcher im(0.9f, true);
      std::vector<int> m12;
      const int nmi = im.SearchForInitialization(bk[0], bd[0], F2, prev, m12, 100);
      std::vector<orb_oracle_query> oq(bk[0].size());
      for (size_t i = 0; i < oq.size(); ++i)
        oq[i] = {bk[0][i].pt.x, bk[0][i].pt.y, 100.f, -1.f, bk[0][i].angle, 0, 0, bk[0][i].octave > 0 ? 1 : 0, 0};
      orb_oracle_search_params op{0.f, 0.f, (float)W, (float)H, ORB_SLAM2::ORBmatcher::TH_LOW, 0.9f, 1, 0};
      std::vector<int32_t> want(oq.size());
      const int onm = orb_oracle_search_for_initialization(reinterpret_cast<const orb_oracle_kp*>(bk[1].data()), bd[1].ptr<unsigned char>(0), (int)bk[1].size(),
                                                           oq.data(), bd[0].ptr<unsigned char>(0), (int)oq.size(), &op, want.data(), nullptr);
      REQUIRE(nmi == onm && std::memcmp(m12.data(), want.data(), want.size() * 4) == 0, "SearchForInitialization: %d matches vs oracle %d", nmi, onm);
      REQUIRE(nmi > 30, "SearchForInitialization: too few matches (%d)", nmi);
      for (size_t i = 0; i < m12.size(); ++i)
        if (m12[i] >= 0) REQUIRE(prev[i].x == bk[1][m12[i]].pt.x && prev[i].y == bk[1][m12[i]].pt.y, "vbPrevMatched %zu not updated", i);
      std::printf("SearchForInitialization: %d matches, vnMatches12 bit-exact vs oracle\n", nmi);
    }
    // distinctive descriptors: groups of consecutive left descriptors stand in for the observations of map points
    {
      std::vector<int64_t> off{0};
      for (int n = 1; off.back() + n <= bd[0].rows && n <= 40; ++n) off.push_back(off.back() + n);
      cv::Mat obs((int)off.back(), 32, CV_8U, bd[0].ptr<unsigned char>(0));
      std::vector<int> best;
      ORB_SLAM2::ORBmatcher::ComputeDistinctiveDescriptors(obs, off, best);
      std::vector<int32_t> want(off.size() - 1);
      orb_oracle_distinctive_descriptors(obs.ptr<unsigned char>(0), off.data(), (int)want.size(), want.data(), nullptr, 2);
      REQUIRE(best.size() == want.size() && std::memcmp(best.data(), want.data(), want.size() * 4) == 0, "ComputeDistinctiveDescriptors differs from the oracle");
      std::printf("ComputeDistinctiveDescriptors: %zu map points exact vs oracle\n", want.size());
    }
    // ComputeBoW: a random k = 8, L = 3 vocabulary whose nodes are drawn from the frame's own descriptors
    {
      const int k = 8, L = 3;
      std::vector<int> parent(1, 0);
      std::vector<unsigned char> leaf(1, 0), vd(32, 0);
      std::vector<double> weight(1, 0.0);
      unsigned lcg = 12345u;
      auto rnd = [&]() { lcg = lcg * 1664525u + 1013904223u; return lcg >> 8; };
      std::vector<int> frontier(1, 0);
      for (int lvl = 1; lvl <= L; ++lvl) {
        std::vector<int> next;
        for (int pid : frontier)
          for (int c = 0; c < k; ++c) {
            const int nid = (int)parent.size();
            parent.push_back(pid);
            leaf.push_back(lvl == L);
            const unsigned char* src = bd[1].ptr<unsigned char>((int)(rnd() % (unsigned)bd[1].rows));
            vd.insert(vd.end(), src, src + 32);
            weight.push_back(lvl == L ? ((rnd() % 19u) == 0 ? 0.0 : 0.25 + (double)(rnd() % 1000u) / 137.0) : 0.0);
            next.push_back(nid);
          }
        frontier.swap(next);
      }
      ORB_SLAM2::ORBVocabulary voc;
      voc.create(k, L, parent, leaf, vd, weight);
      orb_oracle_vocabulary* ov = orb_oracle_vocabulary_create(k, L, (int)parent.size(), parent.data(), leaf.data(), vd.data(), weight.data());
      REQUIRE((int)voc.size() == k * k * k, "vocabulary size %u", voc.size());
      ORB_SLAM2::ORBVocabulary::BowVector bow;
      ORB_SLAM2::ORBVocabulary::FeatureVector fv;
      std::vector<cv::Mat> rows;
      for (int r = 0; r < bd[0].rows; ++r) rows.push_back(bd[0].row(r));   // Converter::toDescriptorVector
      voc.transform(rows, bow, fv, 2);
      const int n = bd[0].rows;
      std::vector<int32_t> bw(n), fn(n), fs(n + 1), ff(n);
      std::vector<double> bv(n);
      int32_t bc = 0, fc = 0;
      orb_oracle_bow_transform(ov, bd[0].ptr<unsigned char>(0), n, 2, nullptr, nullptr, bw.data(), bv.data(), &bc, fn.data(), fs.data(), ff.data(), &fc);
      REQUIRE((int)bow.size() == bc && (int)fv.size() == fc && bc > 20, "BowVector / FeatureVector sizes %zu / %zu vs oracle %d / %d", bow.size(), fv.size(), bc, fc);
      int i = 0;
      for (const auto& e : bow) {
        REQUIRE((int)e.first == bw[i] && std::memcmp(&e.second, &bv[i], 8) == 0, "BowVector entry %d differs from the oracle", i);
        ++i;
      }
      i = 0;
      for (const auto& e : fv) {
        REQUIRE((int)e.first == fn[i] && (int)e.second.size() == fs[i + 1] - fs[i] &&
                    std::memcmp(e.second.data(), &ff[fs[i]], e.second.size() * 4) == 0, "FeatureVector node %d differs from the oracle", i);
        ++i;
      }
      // SearchByBoW between the two images through their FeatureVectors
      {
        ORB_SLAM2::ORBVocabulary::BowVector bow2;
        ORB_SLAM2::ORBVocabulary::FeatureVector fv2;
        voc.transform(bd[1], bow2, fv2, 2);
        std::vector<unsigned char> has(bk[0].size(), 1);
        for (size_t q = 0; q < has.size(); q += 9) has[q] = 0;
        ORB_SLAM2::ORBmatcher bm(0.7f, true);
        std::vector<int> asg;
        const int nmb = bm.SearchByBoW(bd[0], bk[0], has, fv, bd[1], bk[1], fv2, asg);
        auto csr = [](const ORB_SLAM2::ORBVocabulary::FeatureVector& m, int cap, std::vector<int32_t>& nd, std::vector<int32_t>& st, std::vector<int32_t>& fe) {
          nd.assign(cap, 0); st.assign(cap + 1, 0); fe.assign(cap, 0);
          int c = 0, pos = 0;
          for (const auto& e : m) { nd[c] = (int)e.first; st[c] = pos; for (unsigned v : e.second) fe[pos++] = (int)v; ++c; }
          st[c] = pos;
          return c;
        };
        std::vector<int32_t> kn, ks, ke, f2n, f2s, f2e, want(bk[1].size());
        const int kc = csr(fv, (int)bk[0].size(), kn, ks, ke), fc2 = csr(fv2, (int)bk[1].size(), f2n, f2s, f2e);
        const int onm = orb_oracle_search_by_bow(bd[0].ptr<unsigned char>(0), reinterpret_cast<const orb_oracle_kp*>(bk[0].data()), has.data(), kn.data(),
                                                 ks.data(), ke.data(), kc, bd[1].ptr<unsigned char>(0), reinterpret_cast<const orb_oracle_kp*>(bk[1].data()),
                                                 (int)bk[1].size(), f2n.data(), f2s.data(), f2e.data(), fc2, ORB_SLAM2::ORBmatcher::TH_LOW, 0.7f, 1, want.data(), nullptr, nullptr, 0);
        REQUIRE(nmb == onm && std::memcmp(asg.data(), want.data(), want.size() * 4) == 0, "SearchByBoW: %d matches vs oracle %d", nmb, onm);
        REQUIRE(nmb > 30, "SearchByBoW: too few matches (%d)", nmb);
        std::printf("SearchByBoW: %d matches, assignment bit-exact vs oracle\n", nmb);
        // SearchForTriangulation between the same two feature sets: sideways-translation F12 (x1^T F12 x2 = y2 - y1), an epipole inside
        // the image, a loose chi-square table so that the rows of the two synthetic images need not agree
        {
          const float F12[9] = {0.f, 0.f, 0.f, 0.f, 0.f, -1.f, 0.f, 1.f, 0.f};
          std::vector<float> sig(8), scl(8), ur1(bk[0].size(), -1.f), ur2(bk[1].size(), -1.f);
          for (int l = 0; l < 8; ++l) { scl[l] = l ? scl[l - 1] * 1.2f : 1.f; sig[l] = scl[l] * scl[l] * 2.0e4f; }
          for (size_t q = 0; q < ur1.size(); q += 4) ur1[q] = bk[0][q].pt.x - 10.f;
          for (size_t q = 1; q < ur2.size(); q += 5) ur2[q] = bk[1][q].pt.x - 10.f;
          std::vector<unsigned char> has2(bk[1].size(), 0);
          for (size_t q = 3; q < has2.size(); q += 7) has2[q] = 1;
          std::vector<unsigned char> hp1(has.size());
          for (size_t q = 0; q < has.size(); ++q) hp1[q] = !has[q];
          std::vector<std::pair<size_t, size_t>> pairs;
          const int nt = bm.SearchForTriangulation(bd[0], bk[0], hp1, ur1, fv, bd[1], bk[1], has2, ur2, fv2, F12, 300.5f, 200.25f, sig, scl, pairs, false);
          std::vector<int32_t> want12(bk[0].size());
          const int ont = orb_oracle_search_for_triangulation(
              bd[0].ptr<unsigned char>(0), reinterpret_cast<const orb_oracle_kp*>(bk[0].data()), hp1.data(), ur1.data(), (int)bk[0].size(), kn.data(),
              ks.data(), ke.data(), kc, bd[1].ptr<unsigned char>(0), reinterpret_cast<const orb_oracle_kp*>(bk[1].data()), has2.data(), ur2.data(),
              (int)bk[1].size(), f2n.data(), f2s.data(), f2e.data(), fc2, F12, 300.5f, 200.25f, sig.data(), scl.data(), 0, ORB_SLAM2::ORBmatcher::TH_LOW, 1,
              want12.data());
          size_t np = 0;
          bool same = nt == ont;
          for (size_t q = 0; q < want12.size() && same; ++q)
            if (want12[q] >= 0) { same = np < pairs.size() && pairs[np].first == q && pairs[np].second == (size_t)want12[q]; ++np; }
          REQUIRE(same && np == pairs.size(), "SearchForTriangulation: %d matches vs oracle %d", nt, ont);
          REQUIRE(nt > 10, "SearchForTriangulation: too few matches (%d)", nt);
          std::printf("SearchForTriangulation: %d matched pairs bit-exact vs oracle\n", nt);
        }
      }
      orb_oracle_vocabulary_destroy(ov);
      std::printf("ORBVocabulary::transform: %d words, %d nodes, doubles bit-exact vs oracle\n", bc, fc);
    }
    // UndistortKeyPoints / ComputeImageBounds with the TUM1 camera (Examples/Monocular/TUM1.yaml)
    {
      const std::vector<float> dc{0.262383f, -0.953104f, -0.005358f, 0.002628f, 1.163314f};
      const float fx = 517.306408f, fy = 516.469215f, cx = 318.643040f, cy = 255.313989f;
      std::vector<cv::KeyPoint> un;
      ORB_SLAM2::UndistortKeyPoints(bk[0], fx, fy, cx, cy, dc, un);
      std::vector<float> xy(bk[0].size() * 2);
      for (size_t i = 0; i < bk[0].size(); ++i) { xy[2 * i] = bk[0][i].pt.x; xy[2 * i + 1] = bk[0][i].pt.y; }
      orb_oracle_undistort_points(fx, fy, cx, cy, dc.data(), (int)dc.size(), xy.data(), (int)bk[0].size());
      for (size_t i = 0; i < un.size(); ++i)
        REQUIRE(std::memcmp(&un[i].pt.x, &xy[2 * i], 8) == 0 && un[i].octave == bk[0][i].octave && un[i].angle == bk[0][i].angle, "UndistortKeyPoints %zu", i);
      float x0, x1, y0, y1;
      ORB_SLAM2::ComputeImageBounds(W, H, fx, fy, cx, cy, dc, x0, x1, y0, y1);
      float c4[8] = {0.f, 0.f, (float)W, 0.f, 0.f, (float)H, (float)W, (float)H};
      orb_oracle_undistort_points(fx, fy, cx, cy, dc.data(), (int)dc.size(), c4, 4);
      REQUIRE(x0 == std::min(c4[0], c4[4]) && x1 == std::max(c4[2], c4[6]) && y0 == std::min(c4[1], c4[3]) && y1 == std::max(c4[5], c4[7]), "ComputeImageBounds");
      std::printf("UndistortKeyPoints: %zu keypoints bit-exact vs oracle, bounds [%g, %g] x [%g, %g]\n", un.size(), x0, x1, y0, y1);
    }
    std::printf("ORBmatcher: kNN over %d x %d bit-exact, %d ratio-test accepts, DescriptorDistance x %d exact\n", bd[0].rows, bd[1].rows, nacc, m);
  }
  orb_oracle_destroy(oe);
  std::printf("PASS\n");
  return 0;
}
