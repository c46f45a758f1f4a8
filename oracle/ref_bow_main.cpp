// TEST INFRASTRUCTURE ONLY — driver around the reference's own DBoW2 (Thirdparty/DBoW2/DBoW2/*.cpp + TemplatedVocabulary.h,
// compiled unmodified against oracle/cvshim; see oracle/Makefile target `_ref`): ORBVocabulary::loadFromTextFile +
// transform(features, BowVector, FeatureVector, levelsup) exactly as Frame::ComputeBoW calls them (src/Frame.cc:750-765).
//
//   orb_ref_bow VOC.txt DESC.bin OUT.bin
// DESC: int32 {frames, levelsup}, per frame int32 n + n x 32 B descriptors
// OUT : per frame int32 nb, nb x {uint32 word, float64 value} (BowVector in map order),
//                 int32 nf, nf x {uint32 node, int32 m, m x uint32 feature index} (FeatureVector in map order)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "FORB.h"
#include "TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;   // include/ORBVocabulary.h:41-42

static void put(FILE* f, const void* p, size_t n) { if (std::fwrite(p, 1, n, f) != n) { std::perror("write"); std::exit(2); } }

int main(int argc, char** argv) {
  if (argc != 4) { std::fprintf(stderr, "usage: %s VOC.txt DESC.bin OUT.bin\n", argv[0]); return 1; }
  ORBVocabulary voc;
  if (!voc.loadFromTextFile(argv[1])) { std::fprintf(stderr, "loadFromTextFile failed\n"); return 3; }
  FILE* fi = std::fopen(argv[2], "rb");
  FILE* fo = std::fopen(argv[3], "wb");
  if (!fi || !fo) { std::perror("open"); return 1; }
  int32_t h[2];
  if (std::fread(h, 4, 2, fi) != 2) return 1;
  for (int f = 0; f < h[0]; ++f) {
    int32_t n;
    if (std::fread(&n, 4, 1, fi) != 1) return 1;
    cv::Mat desc(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0 && std::fread(desc.data, 32, n, fi) != (size_t)n) return 1;
    std::vector<cv::Mat> rows;   // Converter::toDescriptorVector (src/Converter.cc): one 1x32 Mat per descriptor
    rows.reserve(n);
    for (int j = 0; j < n; ++j) rows.push_back(desc.row(j));
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    voc.transform(rows, bv, fv, h[1]);
    const int32_t nb = (int32_t)bv.size();
    put(fo, &nb, 4);
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it) {
      const uint32_t w = it->first;
      const double v = it->second;
      put(fo, &w, 4); put(fo, &v, 8);
    }
    const int32_t nf = (int32_t)fv.size();
    put(fo, &nf, 4);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
      const uint32_t node = it->first;
      const int32_t m = (int32_t)it->second.size();
      put(fo, &node, 4); put(fo, &m, 4);
      for (int j = 0; j < m; ++j) { const uint32_t idx = it->second[j]; put(fo, &idx, 4); }
    }
  }
  std::fclose(fo);
  std::fclose(fi);
  return 0;
}
