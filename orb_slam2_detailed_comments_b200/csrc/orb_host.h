// Shared host-side pieces of the C-ABI translation units (orb_api.cu: extractor; orb_api_match.cu: matcher, vocabulary, frame
// geometry): error reporting, the CUDA-call guard, the grow-only device buffer and the kernel launchers of the other .cu files.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <chrono>
#include <vector>

#include "../../include/orb_b200.h"
#include "orb_common.cuh"

// ---- kernel launchers (other translation units) -------------------------------------------------------------
struct OrbKnn2;
void orb_launch_pyramid(const uint8_t*, int, long long, uint8_t*, const ResizeTab*, const int* rs_tiles, const void* rs_maps, int num_sms, const OrbGeom&,
                        int, cudaStream_t, long long*, cudaEvent_t);
void orb_launch_fast(const uint8_t*, const void* tiles, const void* maps, int*, uint32_t*, int*, const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_fast_v1(const uint8_t*, int*, uint32_t*, int*, const OrbGeom&, int, cudaStream_t, long long*);   // round-1 kernel (A/B only)
void orb_fast_build_tiles(const OrbGeom&, std::vector<int4>&);
// per-level TMA descriptors (tma_maps.cu): 0 = ok; maps_out: ORB_LEVEL_MAPS_BYTES, 64-byte aligned; box_h == 0 -> hCell + 6 of each level
int orb_build_level_maps(const uint8_t* base, int frames, const OrbGeom&, int box_w, int box_h, void* maps_out);
#define ORB_LEVEL_MAPS_BYTES (128 * ORB_MAX_LEVELS)
void orb_launch_quadtree(const int*, const uint32_t*, const QtScratch&, uint32_t*, int*, int*, const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_blur(const uint8_t*, uint8_t*, const void* maps, const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_describe(const uint8_t*, const uint8_t*, const void* ic_maps, const void* pt_maps, const uint32_t*, const int*, OrbKeypoint*, uint8_t*, int, int*, int*,
                         const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_stereo(const uint8_t*, const OrbKeypoint*, const uint8_t*, const int*, int, float, float, float*, float*, int*, int*, int*,
                       unsigned short*, const OrbGeom&, int, cudaStream_t, long long*);
int orb_knn_num_chunks(int Q, long long D, int num_sms);
void orb_launch_knn(const uint8_t*, int, const uint8_t*, long long, long long, int, unsigned long long*, int, OrbKnn2*, cudaStream_t);
void orb_launch_knn_merge_shards(const OrbKnn2*, int, int, int, OrbKnn2*, cudaStream_t);
// box: OrbPeerBox (knn.cu) = ORB_KNN_PEER_MAX_RANKS slot pointers followed by as many flag pointers
void orb_launch_knn_peer(const uint8_t*, int, const uint8_t*, long long, long long, int, unsigned long long*, int, const void* box, int my_rank, int nranks,
                         int max_q, unsigned epoch, unsigned* ticket, OrbKnn2*, int* err_flag, cudaStream_t);
void orb_launch_distance_batch(const uint8_t*, const uint8_t*, long long, int*, cudaStream_t);
long long orb_knn_launches();
void orb_knn_count_launches(long long n);
void orb_launch_search_by_bow(const uint8_t*, const OrbKeypoint*, const uint8_t*, const int*, const int*, const int*, const int*, int, const uint8_t*,
                              const OrbKeypoint*, const uint8_t*, const int*, const int*, const int*, const int*, int, int, int, float, int, void*, int*, uint2*, int*,
                              int*, int*, int*, cudaStream_t, long long*, const void* epi = nullptr, const float* kf_uright = nullptr,
                              const float* f_uright = nullptr);
void orb_launch_search_init(const OrbKeypoint*, const uint8_t*, const int*, int, const void*, const uint8_t*, const int*, int, int, float, float, float,
                            float, int, float, int, uint4*, int*, uint2*, int*, int*, int*, int*, cudaStream_t, long long*);
void orb_launch_undistort(const void*, const OrbKeypoint*, const int*, int, int, OrbKeypoint*, cudaStream_t);
struct BowVoc {   // == bow.cu
  const int* child_start; const int* child_id; const uint4* child_desc; const uint8_t* is_leaf; const int* word_of; const double* word_weight;
  int L;
};
void orb_launch_bow(const BowVoc&, int, const uint8_t*, const int*, int, int, int*, int*, int*, double*, int*, int*, int*, int*, int*,
                    cudaStream_t, long long*);

void orb_launch_distinctive(const uint8_t*, const long long*, int, int*, int*, cudaStream_t);
void orb_launch_guided(const OrbKeypoint*, const uint8_t*, const float*, const uint8_t*, const int*, int, const void*, const uint8_t*, const int*,
                       int, int, float, float, float, float, int, float, int, uint4*, int*, uint2*, int*, int*, int*, int*, cudaStream_t, long long*);
void orb_launch_fuse_search(const OrbKeypoint*, const uint8_t*, const float*, const int*, int, const void*, const uint8_t*, const int*, int, int, float,
                            float, float, float, const float*, int, uint4*, int*, int*, int*, cudaStream_t, long long*);
double orb_popc_peak(int num_sms);


static_assert(sizeof(orb_keypoint) == 28 && sizeof(OrbKeypoint) == 28, "cv::KeyPoint layout");
static_assert(sizeof(orb_knn2) == 24, "orb_knn2 layout");

// ---- errors ------------------------------------------------------------------------------------------
int fail(int code, const char* fmt, ...);   // stores the message for orb_last_error() (thread-local) and returns `code`
#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) return fail(ORB_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

// A handle lives on one device; an entry point works on that device and leaves the CALLER's current device as it found it
// (a drop-in library must not move the host thread to another GPU behind the application's back).
struct OrbDeviceGuard {
  int prev = -1;
  cudaError_t err;
  explicit OrbDeviceGuard(int dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    else if (err == cudaSuccess) prev = -1;   // already there: nothing to restore
  }
  ~OrbDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  OrbDeviceGuard(const OrbDeviceGuard&) = delete;
  OrbDeviceGuard& operator=(const OrbDeviceGuard&) = delete;
};
#define ORB_ON_DEVICE(dev)                                                                                            \
  OrbDeviceGuard orb_dev_guard_(dev);                                                                                 \
  if (orb_dev_guard_.err != cudaSuccess) return fail(ORB_E_CUDA, "cudaSetDevice(%d): %s", (int)(dev), cudaGetErrorString(orb_dev_guard_.err))

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int ensure(size_t need) {
    if (need <= bytes) return ORB_OK;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    cudaError_t e = cudaMalloc(&p, need);
    if (e != cudaSuccess) return fail(ORB_E_CUDA, "cudaMalloc(%zu): %s", need, cudaGetErrorString(e));
    bytes = need;
    return ORB_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

