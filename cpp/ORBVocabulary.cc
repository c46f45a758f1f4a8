// ORB_SLAM2::ORBVocabulary over the C ABI (include/orb_b200.h): host glue only.
#include "ORBVocabulary.h"

#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "../include/orb_b200.h"

namespace ORB_SLAM2 {

static void check(int rc, const char* what) {
  if (rc < 0) throw std::runtime_error(std::string(what) + ": " + orb_last_error());
}

ORBVocabulary::~ORBVocabulary() { orb_vocabulary_destroy(mpHandle); }

unsigned int ORBVocabulary::size() const { return (unsigned int)orb_vocabulary_words(mpHandle); }

void ORBVocabulary::create(int k, int L, const std::vector<int>& parent, const std::vector<unsigned char>& isLeaf,
                           const std::vector<unsigned char>& desc, const std::vector<double>& weight) {
  orb_vocabulary_destroy(mpHandle);
  mpHandle = nullptr;
  const int n = (int)parent.size();
  if ((int)isLeaf.size() != n || (int)weight.size() != n || desc.size() != (size_t)n * 32) throw std::runtime_error("ORBVocabulary::create: array sizes differ");
  check(orb_vocabulary_create(k, L, n, parent.data(), isLeaf.data(), desc.data(), weight.data(), mDevice, &mpHandle), "orb_vocabulary_create");
}

bool ORBVocabulary::loadFromTextFile(const std::string& filename) {
  std::ifstream f(filename.c_str());
  if (!f.is_open()) return false;
  std::string line;
  if (!std::getline(f, line)) return false;
  int k = -1, L = -1, n1 = -1, n2 = -1;
  std::stringstream hs(line);
  hs >> k >> L >> n1 >> n2;
  if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return false;   // TemplatedVocabulary.h:1389
  if (n1 != 0 || n2 != 0) return false;   // only L1_NORM + TF_IDF (the ORB vocabulary) on this path
  std::vector<int> parent(1, 0);
  std::vector<unsigned char> leaf(1, 0), desc(32, 0);
  std::vector<double> weight(1, 0.0);
  while (std::getline(f, line)) {
    if (line.find_first_not_of(" \t\r\n") == std::string::npos) continue;
    std::stringstream ss(line);
    int pid = 0, isleaf = 0;
    ss >> pid >> isleaf;
    unsigned char d[32];
    for (int i = 0; i < 32; ++i) { int b = 0; ss >> b; d[i] = (unsigned char)b; }
    double w = 0.0;
    ss >> w;
    if (ss.fail()) return false;
    parent.push_back(pid);
    leaf.push_back(isleaf > 0);
    desc.insert(desc.end(), d, d + 32);
    weight.push_back(w);
  }
  if (parent.size() < 2) return false;
  create(k, L, parent, leaf, desc, weight);
  return true;
}

void ORBVocabulary::transform(const cv::Mat& descriptors, BowVector& v, FeatureVector& fv, int levelsup) const {
  v.clear();
  fv.clear();
  if (empty() || descriptors.rows == 0) return;   // "if(empty()) return" (TemplatedVocabulary.h:1143)
  if (descriptors.cols != 32 || descriptors.type() != CV_8U) throw std::runtime_error("ORBVocabulary::transform: descriptors must be N x 32 CV_8U");
  const int n = descriptors.rows;
  std::vector<unsigned char> rows;
  const unsigned char* d = descriptors.ptr<unsigned char>(0);
  if (!descriptors.isContinuous()) {
    rows.resize((size_t)n * 32);
    for (int r = 0; r < n; ++r) std::memcpy(&rows[(size_t)r * 32], descriptors.ptr<unsigned char>(r), 32);
    d = rows.data();
  }
  std::vector<int32_t> bw(n), fn(n), fs(n + 1), ff(n);
  std::vector<double> bv(n);
  int32_t bc = 0, fc = 0;
  check(orb_bow_transform(mpHandle, ORB_MEM_HOST, d, &n, n, 1, levelsup, nullptr, nullptr, bw.data(), bv.data(), &bc, fn.data(), fs.data(), ff.data(),
                          &fc, nullptr),
        "orb_bow_transform");
  for (int i = 0; i < bc; ++i) v.insert(v.end(), BowVector::value_type((unsigned)bw[i], bv[i]));   // ascending: O(1) inserts
  for (int j = 0; j < fc; ++j) {
    std::vector<unsigned int>& dst = fv.insert(fv.end(), FeatureVector::value_type((unsigned)fn[j], std::vector<unsigned int>()))->second;
    dst.assign(ff.begin() + fs[j], ff.begin() + fs[j + 1]);
  }
}

void ORBVocabulary::transform(const std::vector<cv::Mat>& features, BowVector& v, FeatureVector& fv, int levelsup) const {
  cv::Mat all((int)features.size(), 32, CV_8U);
  for (size_t i = 0; i < features.size(); ++i) std::memcpy(all.ptr<unsigned char>((int)i), features[i].ptr<unsigned char>(0), 32);
  transform(all, v, fv, levelsup);
}

}  // namespace ORB_SLAM2
