// TEST INFRASTRUCTURE ONLY — driver around the reference's own ORBextractor (compiled unmodified from
// /root/reference/src/ORBextractor.cc against oracle/cvshim, see oracle/Makefile target `_ref`).
//
//   orb_ref_extract IN OUT
//   orb_ref_extract --time IN THREADS REPEAT   (stock allocator, one ORBextractor per std::thread over the frames of IN, REPEAT passes;
//                                               prints "images seconds" — bench.py's --impl reference leg)
// IN : int32 {magic 0x4f524231, n, width, height, nfeatures, nlevels, iniThFAST, minThFAST}, float32 scaleFactor, n*width*height u8
// OUT: int32 nlevels, 4 x nlevels float32 (GetScaleFactors, GetInverseScaleFactors, GetScaleSigmaSquares, GetInverseScaleSigmaSquares),
//      per frame: int32 N, N x 28 B cv::KeyPoint, N x 32 B descriptors, per level int32 {cols, rows} + rows*cols u8 of mvImagePyramid[level]
//
// Determinism of the reference: DistributeOctTree sorts (size, ExtractorNode*) pairs (src/ORBextractor.cc:1295-1308), so ties are
// ordered by heap address. Here operator new is a monotonic bump arena (reset between frames): addresses grow in allocation
// order, which makes the reference's tie order "later-created node first" — the order the oracle canonicalises.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <chrono>
#include <new>
#include <thread>
#include <vector>

#include "ORBextractor.h"

namespace {
char* g_arena = nullptr;
size_t g_arena_size = 0, g_arena_pos = 0;
bool g_arena_on = false;
}  // namespace

void* operator new(size_t n) {
  if (g_arena_on) {
    const size_t a = (g_arena_pos + 15) & ~(size_t)15;
    if (a + n > g_arena_size) { std::fprintf(stderr, "orb_ref_extract: arena exhausted\n"); std::abort(); }
    g_arena_pos = a + n;
    return g_arena + a;
  }
  void* p = std::malloc(n ? n : 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept {
  if (p >= (void*)g_arena && p < (void*)(g_arena + g_arena_size)) return;   // arena memory is reclaimed by the per-frame reset
  std::free(p);
}
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

static void put(FILE* f, const void* p, size_t n) { if (std::fwrite(p, 1, n, f) != n) { std::perror("write"); std::exit(2); } }

static int time_mode(const char* in, int threads, int repeat) {
  FILE* fi = std::fopen(in, "rb");
  int32_t h[8];
  float scale;
  if (!fi || std::fread(h, 4, 8, fi) != 8 || h[0] != 0x4f524231 || std::fread(&scale, 4, 1, fi) != 1) { std::fprintf(stderr, "bad input\n"); return 1; }
  const int n = h[1], W = h[2], H = h[3];
  std::vector<unsigned char> frames((size_t)n * W * H);
  if (std::fread(frames.data(), 1, frames.size(), fi) != frames.size()) return 1;
  std::fclose(fi);
  std::atomic<int> next(0);
  std::atomic<long long> total(0);
  auto work = [&](int passes) {
    ORB_SLAM2::ORBextractor ex(h[4], scale, h[5], h[6], h[7]);   // one instance per thread, as in Frame's stereo ctor (src/Frame.cc:158-168)
    for (int i; (i = next.fetch_add(1)) < n * passes;) {
      cv::Mat img(H, W, CV_8UC1, frames.data() + (size_t)(i % n) * W * H), desc;
      std::vector<cv::KeyPoint> kps;
      ex(img, cv::Mat(), kps, desc);
      total += (long long)kps.size();
    }
  };
  auto run = [&](int passes) {
    next = 0;
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t) th.emplace_back(work, passes);
    for (auto& t : th) t.join();
  };
  run(1);   // warm
  const auto t0 = std::chrono::steady_clock::now();
  run(repeat);
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::printf("%d %.6f %lld\n", n * repeat, dt, (long long)total);
  return 0;
}

int main(int argc, char** argv) {
  if (argc == 5 && !std::strcmp(argv[1], "--time")) return time_mode(argv[2], std::atoi(argv[3]), std::atoi(argv[4]));
  if (argc != 3) { std::fprintf(stderr, "usage: %s IN OUT\n", argv[0]); return 1; }
  FILE* fi = std::fopen(argv[1], "rb");
  FILE* fo = std::fopen(argv[2], "wb");
  if (!fi || !fo) { std::perror("open"); return 1; }
  int32_t h[8];
  float scale;
  if (std::fread(h, 4, 8, fi) != 8 || h[0] != 0x4f524231 || std::fread(&scale, 4, 1, fi) != 1) { std::fprintf(stderr, "bad header\n"); return 1; }
  const int n = h[1], W = h[2], H = h[3];
  const bool plain_malloc = std::getenv("ORB_REF_MALLOC") != nullptr;   // the reference's stock allocator: F2 ties follow malloc's addresses
  g_arena_size = (size_t)1 << 31;
  g_arena = (char*)std::malloc(g_arena_size);
  {
    ORB_SLAM2::ORBextractor ex(h[4], scale, h[5], h[6], h[7]);
    const int32_t nl = ex.GetLevels();
    put(fo, &nl, 4);
    const std::vector<float> t0 = ex.GetScaleFactors(), t1 = ex.GetInverseScaleFactors(), t2 = ex.GetScaleSigmaSquares(),
                             t3 = ex.GetInverseScaleSigmaSquares();
    put(fo, t0.data(), 4 * nl); put(fo, t1.data(), 4 * nl); put(fo, t2.data(), 4 * nl); put(fo, t3.data(), 4 * nl);
    std::vector<unsigned char> frame((size_t)W * H);
    for (int i = 0; i < n; ++i) {
      if (std::fread(frame.data(), 1, frame.size(), fi) != frame.size()) { std::fprintf(stderr, "short input\n"); return 1; }
      ex.mvImagePyramid.clear();
      ex.mvImagePyramid.resize(nl);
      g_arena_pos = 0;
      g_arena_on = !plain_malloc;
      {
        cv::Mat img(H, W, CV_8UC1, frame.data()), desc;
        std::vector<cv::KeyPoint> kps;
        ex(img, cv::Mat(), kps, desc);
        const int32_t N = (int32_t)kps.size();
        put(fo, &N, 4);
        put(fo, kps.data(), (size_t)N * sizeof(cv::KeyPoint));
        for (int r = 0; r < N; ++r) put(fo, desc.ptr(r), 32);
        for (int l = 0; l < nl; ++l) {
          const cv::Mat& m = ex.mvImagePyramid[l];
          const int32_t d[2] = {m.cols, m.rows};
          put(fo, d, 8);
          for (int y = 0; y < m.rows; ++y) put(fo, m.ptr(y), m.cols);
        }
        ex.mvImagePyramid.clear();      // everything allocated from the arena dies before the reset
        ex.mvImagePyramid.resize(nl);
      }
      g_arena_on = false;
    }
  }
  std::fclose(fo);
  std::fclose(fi);
  return 0;
}
