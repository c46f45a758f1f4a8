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

int ORBmatcher::MatchBruteForce(const cv::Mat& queries, const cv::Mat& database, std::vector<int64_t>& matches, int th) const {
  std::vector<Knn2> r;
  BruteForceKnn2(queries, database, r);
  matches.assign(r.size(), -1);
  int n = 0;
  for (size_t i = 0; i < r.size(); ++i)
    if (r[i].bestIdx >= 0 && Accept(r[i], th)) { matches[i] = r[i].bestIdx; ++n; }
  return n;
}

}  // namespace ORB_SLAM2
