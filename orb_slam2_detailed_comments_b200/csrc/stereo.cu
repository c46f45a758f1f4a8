// Frame::ComputeStereoMatches (src/Frame.cc:880-1176) on the device, for the rectified pairs of one pass: frames (2p, 2p+1) of the
// pass are the left / right image of pair p. It consumes what the extraction kernels left in HBM (keypoints, descriptors, padded
// pyramids), so the reference's D2H copy of both mvImagePyramid vectors (src/Frame.cc:910, 1040) disappears.
//
//   orb_stereo_match_kernel  one warp per left keypoint:
//     1. coarse match: the reference bins the right keypoints by image row (rows floor(y-r)..ceil(y+r), r = 2*scale[octave]) and
//        scans the bin of row (int)vL in ascending iR with `dist < bestDist` (bestDist starts at TH_HIGH). Here every lane tests
//        the bin membership of its right keypoints directly, keeps its best (dist, iR) and the warp takes the minimum of
//        (dist << 16 | iR): the same winner (smallest distance, earliest index).
//     2. refinement: 11 x 11 SAD of centre-subtracted patches over 11 horizontal offsets on level kpL.octave, lanes stride the 121
//        pixels, 11 REDUX adds; the arithmetic is integer-valued in the reference's floats, so integer sums are exact.
//     3. parabola fit, disparity checks and depth in float32 with explicit round-to-nearest ops (no contraction).
//   orb_stereo_filter_kernel one CTA per pair: median of the kept SAD values by a 16-step bitwise selection (the reference sorts
//     (dist, iL) pairs and takes element size/2), then removes matches with dist >= 1.5f * 1.4f * median.
// Canonicalisations shared with the oracle: image rows outside [0, nRows) hold no candidates; an empty match list skips the filter.
#include "orb_common.cuh"

#define ST_WARPS 8
#define ST_TH_HIGH 100
#define ST_TH_ORB 75   // (TH_HIGH + TH_LOW) / 2

// Counting sort of the right keypoints of every pair by image row (int)y (clamped to the image): row_start[pair][nRows + 1] and
// sorted[pair][cap]. The matcher then only visits the rows within the largest search radius of (int)vL; the order inside a row
// is irrelevant because the winner is the minimum of (dist << 16 | iR).
__global__ void __launch_bounds__(256)
orb_stereo_bin_kernel(const OrbKeypoint* __restrict__ kps, const int* __restrict__ counts, int cap, int nRows,
                      int* __restrict__ row_start, unsigned short* __restrict__ sorted) {
  extern __shared__ int s_hist[];   // [nRows + 1] counts -> starts -> cursors
  __shared__ int s_part[256];
  const int pair = blockIdx.x;
  const int nR = min(counts[2 * pair + 1], cap);
  const OrbKeypoint* kr = kps + (long long)(2 * pair + 1) * cap;
  for (int i = threadIdx.x; i <= nRows; i += 256) s_hist[i] = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < nR; i += 256) atomicAdd(&s_hist[min(max((int)kr[i].y, 0), nRows - 1)], 1);
  __syncthreads();
  // exclusive scan: every thread owns a contiguous slice
  const int per = (nRows + 256) / 256;
  const int lo = threadIdx.x * per, hi = min(lo + per, nRows + 1);
  int sum = 0;
  for (int i = lo; i < hi; ++i) sum += s_hist[i];
  s_part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int t = 0; t < 256; ++t) { const int v = s_part[t]; s_part[t] = acc; acc += v; }
  }
  __syncthreads();
  int acc = s_part[threadIdx.x];
  int* rs = row_start + (long long)pair * (nRows + 1);
  for (int i = lo; i < hi; ++i) { const int v = s_hist[i]; s_hist[i] = acc; rs[i] = acc; acc += v; }
  __syncthreads();
  unsigned short* so = sorted + (long long)pair * cap;
  for (int i = threadIdx.x; i < nR; i += 256) so[atomicAdd(&s_hist[min(max((int)kr[i].y, 0), nRows - 1)], 1)] = (unsigned short)i;
}

__global__ void __launch_bounds__(ST_WARPS * 32)
orb_stereo_match_kernel(const uint8_t* __restrict__ pyr, const OrbKeypoint* __restrict__ kps, const uint8_t* __restrict__ desc,
                        const int* __restrict__ counts, int cap, float bf, float b, const int* __restrict__ row_start,
                        const unsigned short* __restrict__ sorted, int margin, float* __restrict__ uright,
                        float* __restrict__ depth, int* __restrict__ sad_out, const __grid_constant__ OrbGeom g) {
  const int pair = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int iL = blockIdx.x * ST_WARPS + (threadIdx.x >> 5);
  const int fL = 2 * pair, fR = 2 * pair + 1;
  const int nL = min(counts[fL], cap);
  if (iL >= cap) return;
  const long long o = (long long)pair * cap + iL;
  if (iL >= nL) {
    if (lane == 0) { uright[o] = -1.f; depth[o] = -1.f; sad_out[o] = -1; }
    return;
  }
  const OrbKeypoint* kl = kps + (long long)fL * cap;
  const OrbKeypoint* kr = kps + (long long)fR * cap;
  const float uL = kl[iL].x, vL = kl[iL].y;
  const int levelL = kl[iL].octave;
  const int nRows = g.lv[0].rows;
  const int row = (int)vL;
  const float maxD = __fdiv_rn(bf, b);          // mbf / minZ, minZ = mb
  const float minU = __fsub_rn(uL, maxD), maxU = uL;   // minD = 0
  unsigned best = ((unsigned)ST_TH_HIGH << 16);  // (dist << 16) | iR; only dist < TH_HIGH can beat it
  if (row >= 0 && row < nRows && !(maxU < 0.f)) {
    const uint4* dl4 = reinterpret_cast<const uint4*>(desc + ((long long)fL * cap + iL) * 32);
    const uint4 a0 = __ldg(dl4), a1 = __ldg(dl4 + 1);
    const int* rs = row_start + (long long)pair * (nRows + 1);
    const unsigned short* so = sorted + (long long)pair * cap;
    const int j0 = rs[max(row - margin, 0)], j1 = rs[min(row + margin + 1, nRows)];
    for (int j = j0 + lane; j < j1; j += 32) {
      const int iR = so[j];
      const OrbKeypoint k = kr[iR];
      const float r = __fmul_rn(2.0f, g.lv[k.octave].scale);
      const int maxr = (int)ceilf(__fadd_rn(k.y, r));
      const int minr = (int)floorf(__fsub_rn(k.y, r));
      if (row < minr || row > maxr) continue;
      if (k.octave < levelL - 1 || k.octave > levelL + 1) continue;
      if (!(k.x >= minU && k.x <= maxU)) continue;
      const uint4* dr4 = reinterpret_cast<const uint4*>(desc + ((long long)fR * cap + iR) * 32);
      const uint4 b0 = __ldg(dr4), b1 = __ldg(dr4 + 1);
      const int dist = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                       __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      const unsigned key = ((unsigned)dist << 16) | (unsigned)iR;
      best = min(best, key);   // smallest distance, then earliest index: the winner of the reference's ascending strict-< scan
    }
  }
  best = __reduce_min_sync(0xffffffffu, best);
  const int bestDist = (int)(best >> 16);
  float out_u = -1.f, out_d = -1.f;
  int out_sad = -1;
  if (bestDist < ST_TH_ORB) {   // uniform over the warp
    const int bestIdxR = (int)(best & 0xffffu);
    const float uR0 = kr[bestIdxR].x;
    const OrbLevel& L = g.lv[levelL];
    const float sf = L.inv_scale;
    const float scaleduL = roundf(__fmul_rn(uL, sf)), scaledvL = roundf(__fmul_rn(vL, sf)), scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const float iniu = __fsub_rn(__fsub_rn(scaleduR0, 5.f), 5.f);
    const float endu = __fadd_rn(__fadd_rn(__fadd_rn(scaleduR0, 5.f), 5.f), 1.f);
    if (!(iniu < 0.f || endu >= (float)L.cols)) {
      const int cy = (int)scaledvL, cxl = (int)scaleduL, cxr = (int)scaleduR0;
      const uint8_t* pl = pyr + (unsigned long long)fL * g.frame_stride + L.plane_off + (long long)cy * L.pitch + cxl;
      const uint8_t* pr = pyr + (unsigned long long)fR * g.frame_stride + L.plane_off + (long long)cy * L.pitch + cxr;
      const int cl = pl[0];
      int sad[11];
#pragma unroll
      for (int i = 0; i < 11; ++i) sad[i] = 0;
      for (int p = lane; p < 121; p += 32) {
        const int dy = p / 11 - 5, dx = p % 11 - 5;
        const int a = (int)pl[dy * L.pitch + dx] - cl;
        const uint8_t* rr = pr + dy * L.pitch + dx;
#pragma unroll
        for (int i = 0; i < 11; ++i) sad[i] += abs(a - ((int)rr[i - 5] - (int)pr[i - 5]));
      }
#pragma unroll
      for (int i = 0; i < 11; ++i) sad[i] = __reduce_add_sync(0xffffffffu, sad[i]);
      int bestSad = 0x7fffffff, bestinc = 0;
#pragma unroll
      for (int i = 0; i < 11; ++i)
        if ((float)sad[i] < (float)bestSad) { bestSad = sad[i]; bestinc = i - 5; }   // float compare like the reference (exact here)
      if (bestinc != -5 && bestinc != 5) {
        float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
        for (int i = 1; i < 10; ++i)
          if (i - 5 == bestinc) { d1 = (float)sad[i - 1]; d2 = (float)sad[i]; d3 = (float)sad[i + 1]; }
        const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2)));
        const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), den);
        if (!(deltaR < -1.f || deltaR > 1.f)) {
          float bestuR = __fmul_rn(L.scale, __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
          float disparity = __fsub_rn(uL, bestuR);
          if (disparity >= 0.f && disparity < maxD) {
            if (disparity <= 0.f) {
              disparity = 0.01f;
              bestuR = (float)((double)uL - 0.01);
            }
            out_d = __fdiv_rn(bf, disparity);
            out_u = bestuR;
            out_sad = bestSad;
          }
        }
      }
    }
  }
  if (lane == 0) { uright[o] = out_u; depth[o] = out_d; sad_out[o] = out_sad; }
}

__global__ void __launch_bounds__(256)
orb_stereo_filter_kernel(const int* __restrict__ counts, int cap, float* __restrict__ uright, float* __restrict__ depth,
                         const int* __restrict__ sad, int* __restrict__ matched) {
  __shared__ int s_cnt[8];
  __shared__ int s_total;
  const int pair = blockIdx.x;
  const int n = min(counts[2 * pair], cap);
  const int* sd = sad + (long long)pair * cap;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  auto block_count = [&](int limit) {   // number of kept matches with sad < limit
    int c = 0;
    for (int i = threadIdx.x; i < n; i += 256) {
      const int v = sd[i];
      c += (v >= 0 && v < limit);
    }
    c = __reduce_add_sync(0xffffffffu, c);
    __syncthreads();
    if (lane == 0) s_cnt[warp] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < 8; ++w) t += s_cnt[w];
      s_total = t;
    }
    __syncthreads();
    return s_total;
  };
  const int m = block_count(0x7fffffff);
  if (m == 0) {
    if (threadIdx.x == 0 && matched) matched[pair] = 0;
    return;
  }
  const int k = m / 2;   // sorted (dist, iL) pairs: element size/2 is the k-th smallest dist
  int prefix = 0;
  for (int bit = 15; bit >= 0; --bit) {
    const int cand = prefix | (1 << bit);
    if (block_count(cand) <= k) prefix = cand;   // the k-th smallest value is >= cand
  }
  const float thDist = (1.5f * 1.4f) * (float)prefix;
  int kept = 0;
  for (int i = threadIdx.x; i < n; i += 256) {
    const int v = sd[i];
    if (v >= 0) {
      if ((float)v < thDist) ++kept;
      else { uright[(long long)pair * cap + i] = -1.f; depth[(long long)pair * cap + i] = -1.f; }
    }
  }
  if (matched) {
    kept = __reduce_add_sync(0xffffffffu, kept);
    __syncthreads();
    if (lane == 0) s_cnt[warp] = kept;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < 8; ++w) t += s_cnt[w];
      matched[pair] = t;
    }
  }
}

void orb_launch_stereo(const uint8_t* pyr, const OrbKeypoint* kps, const uint8_t* desc, const int* counts, int cap, float bf, float b,
                       float* uright, float* depth, int* sad, int* matched, int* row_start, unsigned short* sorted, const OrbGeom& g,
                       int pairs, cudaStream_t st, long long* launches) {
  const int nRows = g.lv[0].rows;
  float smax = 1.f;
  for (int l = 0; l < g.nlevels; ++l) smax = fmaxf(smax, g.lv[l].scale);
  const int margin = (int)ceilf(2.f * smax) + 2;   // rows a right keypoint can be binned away from (int)y: ceil(y + r) - (int)y <= ceil(r) + 1
  const size_t smem = (size_t)(nRows + 1) * sizeof(int);
  orb_stereo_bin_kernel<<<pairs, 256, smem, st>>>(kps, counts, cap, nRows, row_start, sorted);
  orb_stereo_match_kernel<<<dim3((cap + ST_WARPS - 1) / ST_WARPS, pairs), ST_WARPS * 32, 0, st>>>(pyr, kps, desc, counts, cap, bf, b, row_start,
                                                                                                 sorted, margin, uright, depth, sad, g);
  orb_stereo_filter_kernel<<<pairs, 256, 0, st>>>(counts, cap, uright, depth, sad, matched);
  *launches += 3;
}
