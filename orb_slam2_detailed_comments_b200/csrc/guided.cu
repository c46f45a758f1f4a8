// Guided matching core: Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea
// (src/Frame.cc:432-460, 729-745, 633-717), the candidate loops of ORBmatcher::SearchByProjection
// (src/ORBmatcher.cc:69-184, 1702-1866) and ComputeThreeMaxima (src/ORBmatcher.cc:2026-2068).
//
// The reference walks the queries (projected map points) in order and a matched keypoint blocks later queries
// ("mvpMapPoints[idx]->Observations() > 0"), so the result is order-dependent. Three kernels keep that order exact while
// doing the Hamming work in parallel:
//   1. orb_guided_grid_kernel   (CTA per frame): counting sort of the keypoints into the 64 x 48 grid, cells numbered
//      column-major (ix * 48 + iy) and keypoints in index order inside a cell. GetFeaturesInArea's scan order
//      (ix, then iy, then insertion order) is then simply ascending position in the sorted array, and the cells
//      iy0..iy1 of one column are ONE contiguous range.
//   2. orb_guided_candidates_kernel (warp per query): window scan, static filters (level window, |dx|,|dy| < r, caller's
//      claimed flags, stereo column check) and DescriptorDistance; writes the query's candidate list, up to 32 words
//      (dist, sorted position | level, keypoint index), in any order — the position inside the word carries the scan order.
//   3. orb_guided_resolve_kernel (warp per frame): walks the queries in order; lane l holds candidate l of the current
//      query, drops it if its keypoint has been claimed since, and two REDUX.MIN give best and second best ("strict <,
//      first in scan order wins" == unsigned min over dist << 18 | position). Lists are staged one group of 32 queries
//      ahead with cp.async. A query with more than 32 candidates is rescanned by the warp with the claims applied.
//      Accept rule, assignment, claim bit, rotation histogram and the three-maxima filter follow.
#include <atomic>

#include "orb_common.cuh"

#define GRID_COLS 64
#define GRID_ROWS 48
#define GRID_CELLS (GRID_COLS * GRID_ROWS)
#define HISTO_LENGTH 30
#define GUIDED_WARPS 8

struct GuidedQuery {   // == orb_proj_query
  float u, v, radius, ur, angle;
  short min_level, max_level;
  int flags, reserved;
};
static_assert(sizeof(GuidedQuery) == 32, "orb_proj_query layout");

struct OrbEpi {   // == orb_epipolar: SearchForTriangulation's per-pair geometry
  float f12[9];             // F12 row-major
  float ex, ey;             // epipole of camera 1 in image 2 (src/ORBmatcher.cc:885-897)
  float level_sigma2[16];   // pKF2->mvLevelSigma2
  float scale_factor[16];   // pKF2->mvScaleFactors
  float reserved;
};
static_assert(sizeof(OrbEpi) == 176, "orb_epipolar layout");

struct GuidedFrame {
  const uint4* items;        // sorted keypoints: x bits, y bits, octave, index
  const int* cell_start;     // [GRID_CELLS + 1]
  const uint4* desc;         // [n][2]
  const float* uright;       // or nullptr
  const uint8_t* claimed;    // or nullptr (static claims)
  float min_x, min_y, inv_w, inv_h;
  // SearchByBoW: FeatureVector of the target frame as CSR (orb_bow_transform layout)
  const int* fv_start;
  const int* fv_feature;
  // SearchForTriangulation: undistorted keypoints of the target keyframe and the pair's epipolar geometry
  const OrbKeypoint* kps2;
  const OrbEpi* epi;
  // Fuse: mvInvLevelSigma2 of the keyframe (16 floats) for the reprojection chi-square test, or nullptr; fuse != 0 also switches the
  // SearchByProjection-style right-coordinate window test off (Fuse reads mvuRight only inside the chi-square test)
  const float* inv_sigma2;
  int fuse;
};

// A candidate is one 64-bit word {hi, lo}. hi = dist << 18 | sorted position: "strict <, first in scan order wins" is a plain
// unsigned min over it and it is unique per candidate. lo = level << 18 | keypoint index.
#define GUIDED_LIST 32   // candidates kept per query (one per lane of the resolving warp); longer lists fall back to a rescan
#define GUIDED_NONE 0xffffffffu

// GetFeaturesInArea + the static candidate filters + DescriptorDistance for one query, by one warp. Four groups of 8 lanes
// take the grid columns of the window in turn; the lanes of a group stride over the column's contiguous range of sorted
// keypoints. Every slot is handed to emit(hi, lo) by its lane, hi == GUIDED_NONE for "no candidate"; the trip count is
// warp-uniform so emit() may use warp collectives.
template <typename Emit>
__device__ __forceinline__ void guided_scan(const GuidedFrame& F, const GuidedQuery& q, const uint4 qd0, const uint4 qd1, int lane,
                                            Emit emit) {
  const float x = q.u, y = q.v, r = q.radius;
  // src/Frame.cc:649-668
  const int cx0 = max(0, (int)floorf((x - F.min_x - r) * F.inv_w));
  const int cx1 = min(GRID_COLS - 1, (int)ceilf((x - F.min_x + r) * F.inv_w));
  const int cy0 = max(0, (int)floorf((y - F.min_y - r) * F.inv_h));
  const int cy1 = min(GRID_ROWS - 1, (int)ceilf((y - F.min_y + r) * F.inv_h));
  if (cx0 >= GRID_COLS || cx1 < 0 || cy0 >= GRID_ROWS || cy1 < 0) return;
  const int min_level = q.min_level, max_level = q.max_level;
  const bool check_levels = (min_level > 0) || (max_level >= 0);
  const int grp = lane >> 3, sub = lane & 7;
  for (int cx = cx0; cx <= cx1; cx += 4) {   // warp-uniform
    const int ix = cx + grp;
    int e0 = 0, e1 = 0;
    if (ix <= cx1) { e0 = F.cell_start[ix * GRID_ROWS + cy0]; e1 = F.cell_start[ix * GRID_ROWS + cy1 + 1]; }
    int trips = (e1 - e0 + 7) >> 3;
    trips = max(trips, __shfl_xor_sync(0xffffffffu, trips, 8));
    trips = max(trips, __shfl_xor_sync(0xffffffffu, trips, 16));
    for (int t = 0; t < trips; ++t) {
      const int e = e0 + t * 8 + sub;
      unsigned hi = GUIDED_NONE, lo = 0;
      if (e < e1) {
        const uint4 it = F.items[e];
        const int oct = (int)it.z, idx = (int)it.w;
        bool ok = !(check_levels && (oct < min_level || (max_level >= 0 && oct > max_level)));
        const float dx = __uint_as_float(it.x) - x, dy = __uint_as_float(it.y) - y;
        ok = ok && fabsf(dx) < r && fabsf(dy) < r;
        if (ok) {   // the three per-keypoint loads are independent of each other: issue them together
          const uint8_t cl = F.claimed ? F.claimed[idx] : (uint8_t)0;
          const float ur = F.uright ? F.uright[idx] : -1.f;
          const uint4 a = F.desc[2 * idx], b = F.desc[2 * idx + 1];
          const int dist = __popc(a.x ^ qd0.x) + __popc(a.y ^ qd0.y) + __popc(a.z ^ qd0.z) + __popc(a.w ^ qd0.w) +
                           __popc(b.x ^ qd1.x) + __popc(b.y ^ qd1.y) + __popc(b.z ^ qd1.z) + __popc(b.w ^ qd1.w);
          bool keep = !cl;
          if (F.fuse) {
            if (F.inv_sigma2) {   // src/ORBmatcher.cc:1193-1218: float32 error, double compare against the chi-square bound
              const float ex = __fsub_rn(x, __uint_as_float(it.x)), ey = __fsub_rn(y, __uint_as_float(it.y));
              float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
              double bound = 5.99;
              if (ur >= 0.f) {
                const float er = __fsub_rn(q.ur, ur);
                e2 = __fadd_rn(e2, __fmul_rn(er, er));
                bound = 7.8;
              }
              if ((double)__fmul_rn(e2, F.inv_sigma2[min(max(oct, 0), 15)]) > bound) keep = false;
            }
          } else if (ur > 0.f && fabsf(q.ur - ur) > r) {
            keep = false;
          }
          if (keep) {
            hi = ((unsigned)dist << 18) | (unsigned)e;
            lo = ((unsigned)(oct & 0xff) << 18) | (unsigned)idx;
          }
        }
      }
      emit(hi, lo);
    }
  }
}

// SearchByBoW candidates (src/ORBmatcher.cc:296-322): the target frame's features in the query's vocabulary node, i.e. one
// contiguous range of the FeatureVector CSR (slot = q.reserved, resolved by orb_bow_queries_kernel); ascending position in
// the CSR is the reference's loop order. Same emit() contract as guided_scan.
//
// TRI = SearchForTriangulation (src/ORBmatcher.cc:879-1090): the same node ranges, but a candidate must also lie away from the
// epipole (both keypoints monocular, :1005-1013) and within the chi-square bound of the epipolar line (CheckDistEpipolarLine,
// :206-245), and "dist > bestDist -> continue" lets a LATER candidate of equal distance replace the earlier one: the position
// part of the key is reversed, so the unsigned min is "smallest distance, last in scan order". Both checks are pure functions
// of the pair, so testing them for every candidate instead of only for those at or below the running best changes nothing.
template <bool TRI, typename Emit>
__device__ __forceinline__ void bow_scan(const GuidedFrame& F, const GuidedQuery& q, const uint4 qd0, const uint4 qd1, int lane, Emit emit) {
  if (q.reserved < 0) return;
  const int e0 = F.fv_start[q.reserved], e1 = F.fv_start[q.reserved + 1];
  float la = 0.f, lb = 0.f, lc = 0.f;
  if (TRI) {   // epipolar line of kp1 in image 2 (:208-214), float32 without contraction
    const float* f = F.epi->f12;
    la = __fadd_rn(__fadd_rn(__fmul_rn(q.u, f[0]), __fmul_rn(q.v, f[3])), f[6]);
    lb = __fadd_rn(__fadd_rn(__fmul_rn(q.u, f[1]), __fmul_rn(q.v, f[4])), f[7]);
    lc = __fadd_rn(__fadd_rn(__fmul_rn(q.u, f[2]), __fmul_rn(q.v, f[5])), f[8]);
  }
  for (int base = e0; base < e1; base += 32) {
    const int e = base + lane;
    unsigned hi = GUIDED_NONE, lo = 0;
    if (e < e1) {
      const int idx = F.fv_feature[e];
      const uint8_t blocked = F.claimed ? F.claimed[idx] : (uint8_t)0;   // KeyFrame-KeyFrame form: "!pMP2 || pMP2->isBad()" (:771-776)
      const uint4 a = F.desc[2 * idx], b = F.desc[2 * idx + 1];
      const int dist = __popc(a.x ^ qd0.x) + __popc(a.y ^ qd0.y) + __popc(a.z ^ qd0.z) + __popc(a.w ^ qd0.w) +
                       __popc(b.x ^ qd1.x) + __popc(b.y ^ qd1.y) + __popc(b.z ^ qd1.z) + __popc(b.w ^ qd1.w);
      bool ok = !blocked;
      if (TRI && ok) {
        const OrbKeypoint k2 = F.kps2[idx];
        const int oct = min(max(k2.octave, 0), 15);
        const bool stereo1 = q.ur >= 0.f, stereo2 = F.uright ? F.uright[idx] >= 0.f : false;
        if (!stereo1 && !stereo2) {   // :1005-1013
          const float dex = __fsub_rn(F.epi->ex, k2.x), dey = __fsub_rn(F.epi->ey, k2.y);
          if (__fadd_rn(__fmul_rn(dex, dex), __fmul_rn(dey, dey)) < __fmul_rn(100.f, F.epi->scale_factor[oct])) ok = false;
        }
        if (ok) {                     // CheckDistEpipolarLine (:215-245)
          const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, k2.x), __fmul_rn(lb, k2.y)), lc);
          const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
          if (den == 0.f) ok = false;
          else {
            const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
            ok = (double)dsqr < __dmul_rn(3.84, (double)F.epi->level_sigma2[oct]);
          }
        }
      }
      if (ok) {
        hi = ((unsigned)dist << 18) | (unsigned)(TRI ? 0x3ffff - e : e);
        lo = (unsigned)idx;
      }
    }
    emit(hi, lo);
  }
}

// ---- 1. grid ----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) orb_guided_grid_kernel(const OrbKeypoint* __restrict__ kps, const int* __restrict__ n_kps, int cap,
                                                              float min_x, float min_y, float inv_w, float inv_h,
                                                              uint4* __restrict__ items, int* __restrict__ cell_start) {
  __shared__ int s_cnt[GRID_CELLS];
  __shared__ int s_warp[8];
  const int f = blockIdx.x, tid = threadIdx.x;
  const int n = min(n_kps[f], cap);
  kps += (size_t)f * cap;
  items += (size_t)f * cap;
  cell_start += (size_t)f * (GRID_CELLS + 1);
  for (int c = tid; c < GRID_CELLS; c += 256) s_cnt[c] = 0;
  __syncthreads();
  auto cell_of = [&](const OrbKeypoint& k) {   // PosInGrid (src/Frame.cc:734-744)
    const int px = (int)roundf((k.x - min_x) * inv_w), py = (int)roundf((k.y - min_y) * inv_h);
    return (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? -1 : px * GRID_ROWS + py;
  };
  for (int i = tid; i < n; i += 256) {
    const int c = cell_of(kps[i]);
    if (c >= 0) atomicAdd(&s_cnt[c], 1);
  }
  __syncthreads();
  // exclusive scan of the 3072 counts: 12 consecutive cells per thread
  int local[12], sum = 0;
#pragma unroll
  for (int j = 0; j < 12; ++j) { local[j] = s_cnt[tid * 12 + j]; sum += local[j]; }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((tid & 31) >= o) incl += v;
  }
  if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
  __syncthreads();
  int base = incl - sum;
  for (int w = 0; w < (tid >> 5); ++w) base += s_warp[w];
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    cell_start[tid * 12 + j] = base;
    s_cnt[tid * 12 + j] = base;   // becomes the fill cursor
    base += local[j];
  }
  if (tid == 255) cell_start[GRID_CELLS] = base;
  __syncthreads();
  for (int i = tid; i < n; i += 256) {
    const OrbKeypoint k = kps[i];
    const int c = cell_of(k);
    if (c >= 0) {
      const int pos = atomicAdd(&s_cnt[c], 1);
      items[pos] = make_uint4(__float_as_uint(k.x), __float_as_uint(k.y), (unsigned)k.octave, (unsigned)i);
    }
  }
  __syncthreads();
  // insertion order inside a cell = ascending keypoint index: sort each (short) cell segment
  for (int c = tid; c < GRID_CELLS; c += 256) {
    const int e0 = cell_start[c], e1 = s_cnt[c];
    for (int a = e0 + 1; a < e1; ++a) {
      const uint4 v = items[a];
      int b = a - 1;
      while (b >= e0 && items[b].w > v.w) { items[b + 1] = items[b]; --b; }
      items[b + 1] = v;
    }
  }
}

// ---- 2. candidate list per query -------------------------------------------------------------------------------------------
template <int MODE>   // 0 = grid window (SearchByProjection), 1 = vocabulary node (SearchByBoW), 2 = node + epipolar checks (SearchForTriangulation)
__global__ void __launch_bounds__(GUIDED_WARPS * 32)
orb_guided_candidates_kernel(const uint4* __restrict__ items, const int* __restrict__ cell_start, const int* __restrict__ fv_start,
                             const int* __restrict__ fv_feature, const OrbKeypoint* __restrict__ kps2, const OrbEpi* __restrict__ epi,
                             const uint8_t* __restrict__ desc,
                             const float* __restrict__ uright, const uint8_t* __restrict__ claimed, int cap,
                             const GuidedQuery* __restrict__ queries, const uint8_t* __restrict__ qdesc, const int* __restrict__ n_queries,
                             int qcap, float min_x, float min_y, float inv_w, float inv_h, uint2* __restrict__ lists,
                             int* __restrict__ list_cnt) {
  constexpr bool BOW = MODE >= 1;
  const int f = blockIdx.y, lane = threadIdx.x & 31;
  const int iq = blockIdx.x * GUIDED_WARPS + (threadIdx.x >> 5);
  if (iq >= min(n_queries[f], qcap)) return;
  GuidedFrame F;
  F.kps2 = MODE == 2 ? kps2 + (size_t)f * cap : nullptr;
  F.epi = MODE == 2 ? epi + f : nullptr;
  F.items = items + (size_t)f * cap;
  F.cell_start = cell_start + (size_t)f * (GRID_CELLS + 1);
  F.desc = reinterpret_cast<const uint4*>(desc + (size_t)f * cap * 32);
  F.uright = uright ? uright + (size_t)f * cap : nullptr;
  F.claimed = claimed ? claimed + (size_t)f * cap : nullptr;
  F.min_x = min_x; F.min_y = min_y; F.inv_w = inv_w; F.inv_h = inv_h;
  F.inv_sigma2 = nullptr; F.fuse = 0;
  F.fv_start = BOW ? fv_start + (size_t)f * (cap + 1) : nullptr;
  F.fv_feature = BOW ? fv_feature + (size_t)f * cap : nullptr;
  const size_t qi = (size_t)f * qcap + iq;
  const GuidedQuery q = queries[qi];
  uint2* out = lists + qi * GUIDED_LIST;
  int n = 0;
  if (!(q.flags & 1)) {
    const uint4* qd = reinterpret_cast<const uint4*>(qdesc + qi * 32);
    auto emit = [&](unsigned hi, unsigned lo) {
      const unsigned m = __ballot_sync(0xffffffffu, hi != GUIDED_NONE);
      const int pos = n + __popc(m & ((1u << lane) - 1u));
      if (hi != GUIDED_NONE && pos < GUIDED_LIST) out[pos] = make_uint2(lo, hi);
      n += __popc(m);
    };
    if (BOW) bow_scan<MODE == 2>(F, q, qd[0], qd[1], lane, emit);
    else guided_scan(F, q, qd[0], qd[1], lane, emit);
  }
  if (lane >= n && lane < GUIDED_LIST) out[lane] = make_uint2(0u, GUIDED_NONE);   // pad: the resolving warp reads all 32 slots
  if (lane == 0) list_cnt[qi] = n;   // > GUIDED_LIST = overflow, the list is incomplete
}

// ---- 3. sequential claims, accept rule, rotation filter --------------------------------------------------------------------
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}

template <int MODE>
__global__ void __launch_bounds__(32)
orb_guided_resolve_kernel(const uint4* __restrict__ items, const int* __restrict__ cell_start, const int* __restrict__ fv_start,
                          const int* __restrict__ fv_feature, const OrbEpi* __restrict__ epi, const OrbKeypoint* __restrict__ kps,
                          const uint8_t* __restrict__ desc, const float* __restrict__ uright, const uint8_t* __restrict__ claimed,
                          int cap, const GuidedQuery* __restrict__ queries, const uint8_t* __restrict__ qdesc,
                          const int* __restrict__ n_queries, int qcap, float min_x, float min_y, float inv_w, float inv_h,
                          int th_dist, float nn_ratio, int ratio_rule, int check_orientation, const uint2* __restrict__ lists,
                          const int* __restrict__ list_cnt, int* __restrict__ assigned, int* __restrict__ query_match,
                          int* __restrict__ nmatches) {
  extern __shared__ unsigned s_dyn[];   // claim bits, (cap + 31) / 32 words
  __shared__ uint2 s_list[2][32][GUIDED_LIST];   // candidate lists of two groups of 32 queries (16 KB)
  __shared__ int s_hist[HISTO_LENGTH];
  constexpr bool BOW = MODE >= 1;
  const int f = blockIdx.x, lane = threadIdx.x;
  const int m = min(n_queries[f], qcap);
  GuidedFrame F;
  F.kps2 = MODE == 2 ? kps + (size_t)f * cap : nullptr;
  F.epi = MODE == 2 ? epi + f : nullptr;
  F.items = items + (size_t)f * cap;
  F.cell_start = cell_start + (size_t)f * (GRID_CELLS + 1);
  F.desc = reinterpret_cast<const uint4*>(desc + (size_t)f * cap * 32);
  F.uright = uright ? uright + (size_t)f * cap : nullptr;
  F.claimed = claimed ? claimed + (size_t)f * cap : nullptr;
  F.min_x = min_x; F.min_y = min_y; F.inv_w = inv_w; F.inv_h = inv_h;
  F.inv_sigma2 = nullptr; F.fuse = 0;
  F.fv_start = BOW ? fv_start + (size_t)f * (cap + 1) : nullptr;
  F.fv_feature = BOW ? fv_feature + (size_t)f * cap : nullptr;
  kps += (size_t)f * cap;
  queries += (size_t)f * qcap;
  qdesc += (size_t)f * qcap * 32;
  lists += (size_t)f * qcap * GUIDED_LIST;
  list_cnt += (size_t)f * qcap;
  assigned += (size_t)f * cap;
  query_match += (size_t)f * qcap;
  auto stage = [&](int g0, int buf) {   // lists of queries [g0, g0 + 32) -> s_list[buf]; one 256-byte row per instruction
    const int cnt = min(32, m - g0);
    for (int j = 0; j < cnt; ++j) cp_async8(&s_list[buf][j][lane], lists + (size_t)(g0 + j) * GUIDED_LIST + lane);
    asm volatile("cp.async.commit_group;");
  };
  if (m > 0) stage(0, 0);
  for (int i = lane; i < (cap + 31) / 32; i += 32) s_dyn[i] = 0u;
  for (int i = lane; i < cap; i += 32) assigned[i] = -1;
  for (int i = lane; i < qcap; i += 32) query_match[i] = -1;
  if (lane < HISTO_LENGTH) s_hist[lane] = 0;
  __syncwarp();
  const bool use_second = nn_ratio > 0.f;
  int nm = 0;
  for (int g0 = 0, buf = 0; g0 < m; g0 += 32, buf ^= 1) {
    const int cnt = min(32, m - g0);
    int my_cnt = 0, my_flags = 0;
    if (lane < cnt) { my_cnt = list_cnt[g0 + lane]; my_flags = queries[g0 + lane].flags; }
    if (g0 + 32 < m) {
      stage(g0 + 32, buf ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncwarp();
    for (int j = 0; j < cnt; ++j) {
      const int n = __shfl_sync(0xffffffffu, my_cnt, j);
      if (n == 0) continue;
      const int flags = __shfl_sync(0xffffffffu, my_flags, j);
      const int iq = g0 + j;
      unsigned hi1 = GUIDED_NONE, lo1 = 0, hi2 = GUIDED_NONE, lo2 = 0;   // best / second best, identical in all lanes
      if (n <= GUIDED_LIST) {
        const uint2 c = s_list[buf][j][lane];
        const unsigned idx = c.x & 0x3ffffu;
        const bool live = c.y != GUIDED_NONE && !((s_dyn[idx >> 5] >> (idx & 31)) & 1u);
        const unsigned hi = live ? c.y : GUIDED_NONE;
        hi1 = __reduce_min_sync(0xffffffffu, hi);
        if (hi1 == GUIDED_NONE) continue;
        const int l1 = __ffs(__ballot_sync(0xffffffffu, hi == hi1)) - 1;
        lo1 = __shfl_sync(0xffffffffu, c.x, l1);
        if (use_second) {
          hi2 = __reduce_min_sync(0xffffffffu, lane == l1 ? GUIDED_NONE : hi);
          if (hi2 != GUIDED_NONE) lo2 = __shfl_sync(0xffffffffu, c.x, __ffs(__ballot_sync(0xffffffffu, hi == hi2)) - 1);
        }
      } else {   // more candidates than list slots: rescan the window with the claims applied
        const GuidedQuery q = queries[iq];
        const uint4* qd = reinterpret_cast<const uint4*>(qdesc + (size_t)iq * 32);
        unsigned long long k1 = ~0ull, k2 = ~0ull;
        auto emit = [&](unsigned hi, unsigned lo) {
          const unsigned idx = lo & 0x3ffffu;
          if (hi != GUIDED_NONE && !((s_dyn[idx >> 5] >> (idx & 31)) & 1u)) {
            const unsigned long long k = ((unsigned long long)hi << 32) | lo;
            if (k < k1) { k2 = k1; k1 = k; }
            else if (k < k2) k2 = k;
          }
        };
        if (BOW) bow_scan<MODE == 2>(F, q, qd[0], qd[1], lane, emit);
        else guided_scan(F, q, qd[0], qd[1], lane, emit);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
          const unsigned long long lo = min(k1, o1), hi = max(k1, o1);
          k2 = min(hi, min(k2, o2));
          k1 = lo;
        }
        hi1 = (unsigned)(k1 >> 32); lo1 = (unsigned)k1;
        hi2 = (unsigned)(k2 >> 32); lo2 = (unsigned)k2;
        if (hi1 == GUIDED_NONE) continue;
      }
      const int d1 = (int)(hi1 >> 18), idx1 = (int)(lo1 & 0x3ffffu), l1 = (int)(lo1 >> 18);
      bool accept = d1 <= th_dist;                                                          // src/ORBmatcher.cc:170, 1824
      if (use_second) {
        const int d2 = hi2 == GUIDED_NONE ? 256 : (int)(hi2 >> 18);                         // bestDist2 = 256, bestLevel2 = -1 (:114-115)
        const int l2 = hi2 == GUIDED_NONE ? -1 : (int)(lo2 >> 18);
        if (ratio_rule == 0) {
          if (l1 == l2 && (float)d1 > nn_ratio * (float)d2) accept = false;                 // :174 (level-conditional reject)
        } else if (!((float)d1 < nn_ratio * (float)d2)) {
          accept = false;                                                                  // :342 (SearchByBoW: accept iff best < r * second)
        }
      }
      if (accept) {
        ++nm;
        if (lane == 0) {
          assigned[idx1] = iq;          // F.mvpMapPoints[bestIdx] = pMP (:177, :1826)
          query_match[iq] = idx1;
          if (flags & 2) s_dyn[idx1 >> 5] |= 1u << (idx1 & 31);
        }
        __syncwarp();
      }
    }
  }
  if (check_orientation) {   // :1828-1863
    const float factor = HISTO_LENGTH / 360.0f;
    __syncwarp();
    auto bin_of = [&](int iq, int t) {
      float rot = queries[iq].angle - kps[t].angle;
      if (rot < 0.0f) rot += 360.0f;
      int bin = (int)roundf(rot * factor);
      if (bin == HISTO_LENGTH) bin = 0;
      return bin;
    };
    for (int iq = lane; iq < m; iq += 32) {
      const int t = query_match[iq];
      if (t >= 0) {
        const int bin = bin_of(iq, t);
        if (bin >= 0 && bin < HISTO_LENGTH) atomicAdd(&s_hist[bin], 1);
      }
    }
    __syncwarp();
    // ComputeThreeMaxima (:2026-2068), every lane redundantly
    int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      const int s = s_hist[i];
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
    int removed = 0;
    for (int iq = lane; iq < m; iq += 32) {
      const int t = query_match[iq];
      if (t >= 0) {
        const int bin = bin_of(iq, t);
        if (bin >= 0 && bin < HISTO_LENGTH && bin != ind1 && bin != ind2 && bin != ind3) {
          assigned[t] = -1;
          query_match[iq] = -1;
          ++removed;
        }
      }
    }
    nm -= __reduce_add_sync(0xffffffffu, removed);
  }
  if (lane == 0) nmatches[f] = nm;
}

// ---- Fuse / SearchBySim3 core: the best candidate of every query on its own ------------------------------------------------
// ORBmatcher::Fuse (src/ORBmatcher.cc:1100-1280, 1282-1430) and the two projection passes of SearchBySim3 (:1433-1700) pick, for
// every projected map point independently (no claims between queries), the keypoint of GetFeaturesInArea(u, v, radius) inside the
// level window with the smallest DescriptorDistance, first in scan order on ties (strict <); Fuse(pKF, vpMapPoints, th)
// additionally drops candidates whose reprojection error fails the chi-square test. One warp per query.
__global__ void __launch_bounds__(GUIDED_WARPS * 32)
orb_fuse_search_kernel(const uint4* __restrict__ items, const int* __restrict__ cell_start, const uint8_t* __restrict__ desc,
                       const float* __restrict__ uright, int cap, const GuidedQuery* __restrict__ queries, const uint8_t* __restrict__ qdesc,
                       const int* __restrict__ n_queries, int qcap, float min_x, float min_y, float inv_w, float inv_h,
                       const float* __restrict__ inv_sigma2, int th_dist, int* __restrict__ best_idx, int* __restrict__ best_dist) {
  const int f = blockIdx.y, lane = threadIdx.x & 31;
  const int iq = blockIdx.x * GUIDED_WARPS + (threadIdx.x >> 5);
  if (iq >= qcap) return;
  const size_t qi = (size_t)f * qcap + iq;
  int bi = -1, bd = 256;
  if (iq < min(n_queries[f], qcap)) {
    GuidedFrame F;
    F.items = items + (size_t)f * cap;
    F.cell_start = cell_start + (size_t)f * (GRID_CELLS + 1);
    F.desc = reinterpret_cast<const uint4*>(desc + (size_t)f * cap * 32);
    F.uright = uright ? uright + (size_t)f * cap : nullptr;
    F.claimed = nullptr; F.fv_start = nullptr; F.fv_feature = nullptr; F.kps2 = nullptr; F.epi = nullptr;
    F.min_x = min_x; F.min_y = min_y; F.inv_w = inv_w; F.inv_h = inv_h;
    F.inv_sigma2 = inv_sigma2; F.fuse = 1;
    const GuidedQuery q = queries[qi];
    if (!(q.flags & 1)) {
      const uint4* qd = reinterpret_cast<const uint4*>(qdesc + qi * 32);
      unsigned long long k = ~0ull;
      auto emit = [&](unsigned hi, unsigned lo) {
        if (hi != GUIDED_NONE) k = min(k, ((unsigned long long)hi << 32) | lo);
      };
      guided_scan(F, q, qd[0], qd[1], lane, emit);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) k = min(k, __shfl_xor_sync(0xffffffffu, k, o));
      if (k != ~0ull) { bd = (int)(k >> 50); bi = (int)(k & 0x3ffffull); }
    }
  }
  if (lane == 0) {
    best_dist[qi] = bd;                        // bestDist (256 = no candidate), reported even above the threshold
    best_idx[qi] = bd <= th_dist ? bi : -1;    // "if (bestDist <= TH_LOW)" (:1232), TH_HIGH in SearchBySim3 (:1575)
  }
}

void orb_launch_fuse_search(const OrbKeypoint* kps, const uint8_t* desc, const float* uright, const int* n_kps, int cap, const void* queries,
                            const uint8_t* qdesc, const int* n_queries, int qcap, int frames, float min_x, float min_y, float max_x, float max_y,
                            const float* inv_sigma2, int th_dist, uint4* items, int* cell_start, int* best_idx, int* best_dist, cudaStream_t st,
                            long long* launches) {
  const float inv_w = static_cast<float>(GRID_COLS) / static_cast<float>(max_x - min_x);
  const float inv_h = static_cast<float>(GRID_ROWS) / static_cast<float>(max_y - min_y);
  orb_guided_grid_kernel<<<frames, 256, 0, st>>>(kps, n_kps, cap, min_x, min_y, inv_w, inv_h, items, cell_start);
  orb_fuse_search_kernel<<<dim3((qcap + GUIDED_WARPS - 1) / GUIDED_WARPS, frames), GUIDED_WARPS * 32, 0, st>>>(
      items, cell_start, desc, uright, cap, static_cast<const GuidedQuery*>(queries), qdesc, n_queries, qcap, min_x, min_y, inv_w, inv_h, inv_sigma2,
      th_dist, best_idx, best_dist);
  *launches += 2;
}

void orb_launch_guided(const OrbKeypoint* kps, const uint8_t* desc, const float* uright, const uint8_t* claimed, const int* n_kps, int cap,
                       const void* queries, const uint8_t* qdesc, const int* n_queries, int qcap, int frames, float min_x, float min_y,
                       float max_x, float max_y, int th_dist, float nn_ratio, int check_orientation, uint4* items, int* cell_start,
                       uint2* lists, int* list_cnt, int* assigned, int* query_match, int* nmatches, cudaStream_t st, long long* launches) {
  // mfGridElementWidthInv / HeightInv (src/Frame.cc:200-202)
  const float inv_w = static_cast<float>(GRID_COLS) / static_cast<float>(max_x - min_x);
  const float inv_h = static_cast<float>(GRID_ROWS) / static_cast<float>(max_y - min_y);
  const GuidedQuery* q = static_cast<const GuidedQuery*>(queries);
  orb_guided_grid_kernel<<<frames, 256, 0, st>>>(kps, n_kps, cap, min_x, min_y, inv_w, inv_h, items, cell_start);
  orb_guided_candidates_kernel<0><<<dim3((qcap + GUIDED_WARPS - 1) / GUIDED_WARPS, frames), GUIDED_WARPS * 32, 0, st>>>(
      items, cell_start, nullptr, nullptr, nullptr, nullptr, desc, uright, claimed, cap, q, qdesc, n_queries, qcap, min_x, min_y, inv_w, inv_h, lists, list_cnt);
  const size_t smem = (size_t)((cap + 31) / 32) * sizeof(unsigned);
  orb_guided_resolve_kernel<0><<<frames, 32, smem, st>>>(items, cell_start, nullptr, nullptr, nullptr, kps, desc, uright, claimed, cap, q, qdesc,
                                                             n_queries, qcap, min_x, min_y, inv_w, inv_h, th_dist, nn_ratio, 0,
                                                             check_orientation, lists, list_cnt, assigned, query_match, nmatches);
  *launches += 3;
}

// ---- SearchForInitialization (src/ORBmatcher.cc:570-712) -------------------------------------------------------------------------
// Same window scan, different bookkeeping: a keypoint matched at distance d stays available to later queries that are strictly
// closer ("if (vMatchedDistance[i2] <= dist) continue", :624); a better match evicts the previous owner (:647-651). The rotation
// histogram counts every accept, evicted ones included (:673), but only standing matches are removed (:696-700).
// One warp per frame pair; shared memory holds vMatchedDistance (u16, 0xffff = INT_MAX) and vnMatches21 per keypoint.
__global__ void __launch_bounds__(32)
orb_guided_resolve_init_kernel(const uint4* __restrict__ items, const int* __restrict__ cell_start, const OrbKeypoint* __restrict__ kps,
                               const uint8_t* __restrict__ desc, int cap, const GuidedQuery* __restrict__ queries,
                               const uint8_t* __restrict__ qdesc, const int* __restrict__ n_queries, int qcap, float min_x, float min_y,
                               float inv_w, float inv_h, int th_dist, float nn_ratio, int check_orientation,
                               const uint2* __restrict__ lists, const int* __restrict__ list_cnt, int* __restrict__ assigned,
                               int* __restrict__ query_match, int* __restrict__ nmatches) {
  extern __shared__ int s_owner[];                                            // [cap] vnMatches21
  unsigned short* s_mdist = reinterpret_cast<unsigned short*>(s_owner + cap);   // [cap] vMatchedDistance
  __shared__ int s_hist[HISTO_LENGTH];
  const int f = blockIdx.x, lane = threadIdx.x;
  const int m = min(n_queries[f], qcap);
  GuidedFrame F;
  F.kps2 = nullptr; F.epi = nullptr;
  F.items = items + (size_t)f * cap;
  F.cell_start = cell_start + (size_t)f * (GRID_CELLS + 1);
  F.desc = reinterpret_cast<const uint4*>(desc + (size_t)f * cap * 32);
  F.uright = nullptr; F.claimed = nullptr; F.fv_start = nullptr; F.fv_feature = nullptr;
  F.min_x = min_x; F.min_y = min_y; F.inv_w = inv_w; F.inv_h = inv_h;
  F.inv_sigma2 = nullptr; F.fuse = 0;
  kps += (size_t)f * cap;
  queries += (size_t)f * qcap;
  qdesc += (size_t)f * qcap * 32;
  lists += (size_t)f * qcap * GUIDED_LIST;
  list_cnt += (size_t)f * qcap;
  assigned += (size_t)f * cap;
  query_match += (size_t)f * qcap;
  for (int i = lane; i < cap; i += 32) { s_owner[i] = -1; s_mdist[i] = 0xffffu; assigned[i] = -1; }
  for (int i = lane; i < qcap; i += 32) query_match[i] = -1;
  if (lane < HISTO_LENGTH) s_hist[lane] = 0;
  __syncwarp();
  const float factor = HISTO_LENGTH / 360.0f;
  int nm = 0;
  for (int iq = 0; iq < m; ++iq) {
    const int n = list_cnt[iq];
    if (n == 0) continue;
    unsigned hi1, lo1 = 0, hi2;
    if (n <= GUIDED_LIST) {
      const uint2 c = lists[(size_t)iq * GUIDED_LIST + lane];
      const unsigned idx = c.x & 0x3ffffu;
      const bool live = c.y != GUIDED_NONE && (c.y >> 18) < (unsigned)s_mdist[idx];   // vMatchedDistance[i2] <= dist -> skip (:624)
      const unsigned hi = live ? c.y : GUIDED_NONE;
      hi1 = __reduce_min_sync(0xffffffffu, hi);
      if (hi1 == GUIDED_NONE) continue;
      const int l1 = __ffs(__ballot_sync(0xffffffffu, hi == hi1)) - 1;
      lo1 = __shfl_sync(0xffffffffu, c.x, l1);
      hi2 = __reduce_min_sync(0xffffffffu, lane == l1 ? GUIDED_NONE : hi);
    } else {
      const GuidedQuery q = queries[iq];
      const uint4* qd = reinterpret_cast<const uint4*>(qdesc + (size_t)iq * 32);
      unsigned long long k1 = ~0ull, k2 = ~0ull;
      guided_scan(F, q, qd[0], qd[1], lane, [&](unsigned hi, unsigned lo) {
        if (hi != GUIDED_NONE && (hi >> 18) < (unsigned)s_mdist[lo & 0x3ffffu]) {
          const unsigned long long k = ((unsigned long long)hi << 32) | lo;
          if (k < k1) { k2 = k1; k1 = k; }
          else if (k < k2) k2 = k;
        }
      });
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
        const unsigned long long lo = min(k1, o1), hi = max(k1, o1);
        k2 = min(hi, min(k2, o2));
        k1 = lo;
      }
      hi1 = (unsigned)(k1 >> 32); lo1 = (unsigned)k1; hi2 = (unsigned)(k2 >> 32);
      if (hi1 == GUIDED_NONE) continue;
    }
    const int d1 = (int)(hi1 >> 18), idx1 = (int)(lo1 & 0x3ffffu);
    const float second = hi2 == GUIDED_NONE ? (float)2147483647 : (float)(int)(hi2 >> 18);   // bestDist2 = INT_MAX (:612)
    if (d1 <= th_dist && (float)d1 < second * nn_ratio) {                                   // :641-644
      const int prev = s_owner[idx1];
      if (prev >= 0) --nm;                                                                   // :647-651
      ++nm;
      if (lane == 0) {
        if (prev >= 0) query_match[prev] = -1;
        query_match[iq] = idx1;
        s_owner[idx1] = iq;
        s_mdist[idx1] = (unsigned short)d1;
        if (check_orientation) {
          float rot = queries[iq].angle - kps[idx1].angle;
          if (rot < 0.0f) rot += 360.0f;
          int bin = (int)roundf(rot * factor);
          if (bin == HISTO_LENGTH) bin = 0;
          if (bin >= 0 && bin < HISTO_LENGTH) ++s_hist[bin];
        }
      }
      __syncwarp();
    }
  }
  __syncwarp();
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      const int s = s_hist[i];
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < 0.1f * (float)max1) { ind3 = -1; }
    int removed = 0;
    for (int iq = lane; iq < m; iq += 32) {
      const int t = query_match[iq];
      if (t >= 0) {
        float rot = queries[iq].angle - kps[t].angle;
        if (rot < 0.0f) rot += 360.0f;
        int bin = (int)roundf(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        if (bin >= 0 && bin < HISTO_LENGTH && bin != ind1 && bin != ind2 && bin != ind3) { query_match[iq] = -1; ++removed; }   // :696-700
      }
    }
    nm -= __reduce_add_sync(0xffffffffu, removed);
    __syncwarp();
  }
  for (int iq = lane; iq < m; iq += 32) {   // inverse map of the standing matches
    const int t = query_match[iq];
    if (t >= 0) assigned[t] = iq;
  }
  if (lane == 0) nmatches[f] = nm;
}

void orb_launch_search_init(const OrbKeypoint* kps, const uint8_t* desc, const int* n_kps, int cap, const void* queries, const uint8_t* qdesc,
                            const int* n_queries, int qcap, int frames, float min_x, float min_y, float max_x, float max_y, int th_dist,
                            float nn_ratio, int check_orientation, uint4* items, int* cell_start, uint2* lists, int* list_cnt, int* assigned,
                            int* query_match, int* nmatches, cudaStream_t st, long long* launches) {
  const float inv_w = static_cast<float>(GRID_COLS) / static_cast<float>(max_x - min_x);
  const float inv_h = static_cast<float>(GRID_ROWS) / static_cast<float>(max_y - min_y);
  const GuidedQuery* q = static_cast<const GuidedQuery*>(queries);
  orb_guided_grid_kernel<<<frames, 256, 0, st>>>(kps, n_kps, cap, min_x, min_y, inv_w, inv_h, items, cell_start);
  orb_guided_candidates_kernel<0><<<dim3((qcap + GUIDED_WARPS - 1) / GUIDED_WARPS, frames), GUIDED_WARPS * 32, 0, st>>>(
      items, cell_start, nullptr, nullptr, nullptr, nullptr, desc, nullptr, nullptr, cap, q, qdesc, n_queries, qcap, min_x, min_y, inv_w, inv_h, lists, list_cnt);
  const size_t smem = (size_t)cap * (sizeof(int) + sizeof(unsigned short)) + 16;
  static std::atomic<bool> configured[64];
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (smem > 48 * 1024 && !configured[dev].load(std::memory_order_acquire)) {
    cudaFuncSetAttribute(orb_guided_resolve_init_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    configured[dev].store(true, std::memory_order_release);
  }
  orb_guided_resolve_init_kernel<<<frames, 32, smem, st>>>(items, cell_start, kps, desc, cap, q, qdesc, n_queries, qcap, min_x, min_y, inv_w,
                                                           inv_h, th_dist, nn_ratio, check_orientation, lists, list_cnt, assigned, query_match,
                                                           nmatches);
  *launches += 3;
}

// ---- SearchByBoW (src/ORBmatcher.cc:248-410) -----------------------------------------------------------------------------------
// Builds the query records from the keyframe's FeatureVector: feature i of node N becomes a query with the slot of N in the
// target frame's FeatureVector (-1 = node not shared: skipped, the reference's lower_bound jumps :376-385), its own angle and the
// CLAIMS flag (every match blocks its keypoint, :300-301). Features outside the FeatureVector, or without a valid map point
// (:283-289), stay SKIP. Queries are resolved in feature-index order, which is the reference's order inside every node.
__global__ void __launch_bounds__(256)
orb_bow_queries_kernel(const OrbKeypoint* __restrict__ kf_kps, const uint8_t* __restrict__ kf_valid, const float* __restrict__ kf_uright,
                       const int* __restrict__ kf_node,
                       const int* __restrict__ kf_start, const int* __restrict__ kf_feature, const int* __restrict__ kf_count, int qcap,
                       const int* __restrict__ f_node, const int* __restrict__ f_count, int cap, GuidedQuery* __restrict__ q,
                       int* __restrict__ n_queries) {
  const int f = blockIdx.x, tid = threadIdx.x;
  kf_kps += (size_t)f * qcap; kf_valid += (size_t)f * qcap;
  kf_node += (size_t)f * qcap; kf_start += (size_t)f * (qcap + 1); kf_feature += (size_t)f * qcap;
  f_node += (size_t)f * cap;
  q += (size_t)f * qcap;
  for (int i = tid; i < qcap; i += 256) {
    GuidedQuery e;
    e.u = e.v = e.radius = e.ur = e.angle = 0.f;
    e.min_level = e.max_level = -1;
    e.flags = 1;   // SKIP
    e.reserved = -1;
    q[i] = e;
  }
  __syncthreads();
  const int nk = min(kf_count[f], qcap), nf = min(f_count[f], cap);
  const int total = nk > 0 ? kf_start[nk] : 0;
  for (int e = tid; e < total; e += 256) {
    int lo = 0, hi = nk;   // node slot of CSR entry e: last j with kf_start[j] <= e
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (kf_start[mid] <= e) lo = mid; else hi = mid; }
    const int node = kf_node[lo];
    int a = 0, b = nf;     // slot of the same node id in the target frame's (ascending) node list
    while (a < b) { const int mid = (a + b) >> 1; if (f_node[mid] < node) a = mid + 1; else b = mid; }
    const int slot = (a < nf && f_node[a] == node) ? a : -1;
    const int i = kf_feature[e];
    if (i >= 0 && i < qcap && kf_valid[i] && slot >= 0) {
      q[i].angle = kf_kps[i].angle;
      q[i].u = kf_kps[i].x; q[i].v = kf_kps[i].y;                               // SearchForTriangulation: kp1 and its stereo flag
      q[i].ur = kf_uright ? kf_uright[(size_t)f * qcap + i] : -1.f;
      q[i].reserved = slot;
      q[i].flags = 2;   // CLAIMS
    }
  }
  if (tid == 0) n_queries[f] = qcap;
}

void orb_launch_search_by_bow(const uint8_t* kf_desc, const OrbKeypoint* kf_kps, const uint8_t* kf_valid, const int* kf_node, const int* kf_start,
                              const int* kf_feature, const int* kf_count, int qcap, const uint8_t* f_desc, const OrbKeypoint* f_kps,
                              const uint8_t* f_blocked, const int* f_node, const int* f_start, const int* f_feature, const int* f_count, int cap,
                              int frames, int th_dist,
                              float nn_ratio, int check_orientation, void* queries, int* n_queries, uint2* lists, int* list_cnt, int* assigned,
                              int* query_match, int* nmatches, cudaStream_t st, long long* launches, const void* epi, const float* kf_uright,
                              const float* f_uright) {
  GuidedQuery* q = static_cast<GuidedQuery*>(queries);
  const OrbEpi* ep = static_cast<const OrbEpi*>(epi);   // non-null = SearchForTriangulation
  orb_bow_queries_kernel<<<frames, 256, 0, st>>>(kf_kps, kf_valid, ep ? kf_uright : nullptr, kf_node, kf_start, kf_feature, kf_count, qcap, f_node,
                                                  f_count, cap, q, n_queries);
  const dim3 grid((qcap + GUIDED_WARPS - 1) / GUIDED_WARPS, frames);
  const size_t smem = (size_t)((cap + 31) / 32) * sizeof(unsigned);
  if (ep) {
    orb_guided_candidates_kernel<2><<<grid, GUIDED_WARPS * 32, 0, st>>>(nullptr, nullptr, f_start, f_feature, f_kps, ep, f_desc, f_uright, f_blocked,
                                                                          cap, q, kf_desc, n_queries, qcap, 0.f, 0.f, 0.f, 0.f, lists, list_cnt);
    orb_guided_resolve_kernel<2><<<frames, 32, smem, st>>>(nullptr, nullptr, f_start, f_feature, ep, f_kps, f_desc, f_uright, f_blocked, cap, q, kf_desc,
                                                           n_queries, qcap, 0.f, 0.f, 0.f, 0.f, th_dist, nn_ratio, 1, check_orientation, lists,
                                                           list_cnt, assigned, query_match, nmatches);
  } else {
    orb_guided_candidates_kernel<1><<<grid, GUIDED_WARPS * 32, 0, st>>>(nullptr, nullptr, f_start, f_feature, nullptr, nullptr, f_desc, nullptr,
                                                                          f_blocked, cap, q, kf_desc, n_queries, qcap, 0.f, 0.f, 0.f, 0.f, lists, list_cnt);
    orb_guided_resolve_kernel<1><<<frames, 32, smem, st>>>(nullptr, nullptr, f_start, f_feature, nullptr, f_kps, f_desc, nullptr, f_blocked, cap, q,
                                                           kf_desc, n_queries, qcap, 0.f, 0.f, 0.f, 0.f, th_dist, nn_ratio, 1, check_orientation, lists,
                                                           list_cnt, assigned, query_match, nmatches);
  }
  *launches += 3;
}
