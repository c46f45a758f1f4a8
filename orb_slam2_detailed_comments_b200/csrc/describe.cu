// Gaussian blur (src/ORBextractor.cc:2039-2047), IC_Angle (:104-161), computeOrbDescriptor (:177-254) and the
// output assembly of ORBextractor::operator() (:2020-2081).
#include <cstdlib>

#include "describe_common.cuh"
#include "orb_tma.cuh"

// ---- 7x7 sigma-2 Gaussian, OpenCV 8-bit fixed-point path (taps 8.8): out = (sum_ij k_i k_j p_ij + 32768) >> 16,
// BORDER_REFLECT_101 at the border of the PADDED level (this fork blurs the padded image; the pyramid kernels have already
// written that reflection into the plane's apron, so no index is remapped here). All intermediate sums are exact integers
// (row sums <= 257*255 fit 16 bits), so the order of the two passes is free. CTA tile = 128 x 32 output pixels, 320 threads:
//   H pass: one thread = 4 adjacent pixels of two adjacent rows: three 32-bit loads per row, the seven taps of a pixel are
//           two DP4A on byte windows (PRMT); the two rows' sums are stored as vertical 16-bit pairs [h(r), h(r+1)], r even;
//   V pass: one thread = 4 pixels x 4 output rows from 6 staged row pairs; the seven taps are four DP2A on the pairs
//           (tap pairs differ for even and odd output rows), accumulator preloaded with the rounding constant.
#define BL_TW 128
#define BL_TH 32
#define BL_THREADS 320
#define BL_RAW_W 160      // TMA box: columns x0 - 16 .. x0 + 143 (16-byte aligned start; the taps need x0 - 4 .. x0 + 131)
#define BL_RAW_H 40       // rows y0 - 4 .. y0 + 35 (row y0 - 4 and rows past y0 + 34 only ever meet a zero tap)
// V pass shared by both blur kernels: one thread = 4 pixels x 4 output rows from 6 staged row pairs.
template <bool CLAMP, int TH = BL_TH>
__device__ __forceinline__ void blur_v_pass(const uint4 (*s_p)[BL_TW / 4], uint8_t* __restrict__ out, const OrbLevel& L, int x0, int y0,
                                            const OrbGeom& g) {
  if (threadIdx.x >= 256) return;
  const int q = threadIdx.x & 31;
  const int x = x0 + 4 * q;
  if (x >= L.cols) return;
  const int pitch = L.pitch;
#pragma unroll
 for (int seg = threadIdx.x >> 5; seg < TH / 4; seg += 8) {   // output rows y0 + 4*seg .. + 3 = local rows 4*seg + 4 .. + 7
  uint4 P[6];                                                // pairs 2*seg .. 2*seg + 5 = local rows 4*seg .. 4*seg + 11
#pragma unroll
  for (int i = 0; i < 6; ++i) P[i] = s_p[2 * seg + i][q];
  // even local row t = 2m: [0,k0].P(m-2) + [k1,k2].P(m-1) + [k3,k2].P(m) + [k1,k0].P(m+1)
  // odd  local row t = 2m+1: [k0,k1].P(m-1) + [k2,k3].P(m) + [k2,k1].P(m+1) + [k0,0].P(m+2)
  const unsigned e0 = g.blur_k[2], e1 = g.blur_k[3], e2 = g.blur_k[4], e3 = g.blur_k[5];
  const unsigned o0 = g.blur_k[6], o1 = g.blur_k[7], o2 = g.blur_k[8], o3 = g.blur_k[9];
  uint8_t* dst = out + (unsigned)((y0 + 4 * seg) * pitch + x);   // 32-bit offset inside the frame's plane
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = y0 + 4 * seg + r;
    if (y < L.rows) {
      const int m = 2 + (r >> 1);                           // pair index (relative to P[0]) holding local row t
      unsigned c[4];
#pragma unroll
      for (int p = 0; p < 4; ++p) {
#define PW(i) (p == 0 ? P[i].x : p == 1 ? P[i].y : p == 2 ? P[i].z : P[i].w)
        if ((r & 1) == 0)
          c[p] = __dp2a_lo(PW(m - 2), e0, __dp2a_lo(PW(m - 1), e1, __dp2a_lo(PW(m), e2, __dp2a_lo(PW(m + 1), e3, 32768u))));
        else
          c[p] = __dp2a_lo(PW(m - 1), o0, __dp2a_lo(PW(m), o1, __dp2a_lo(PW(m + 1), o2, __dp2a_lo(PW(m + 2), o3, 32768u))));
#undef PW
        if (CLAMP) c[p] = min(c[p], 0xffffffu);
      }
      // byte 2 of each accumulator is the output pixel
      const unsigned lo = __byte_perm(c[0], c[1], 0x0062), hi = __byte_perm(c[2], c[3], 0x0062);
      *reinterpret_cast<unsigned*>(dst + r * pitch) = __byte_perm(lo, hi, 0x5410);
    }
  }
 }
}

// Default blur kernel, tile = 128 x TH pixels with TH = 64 by default (orb_blur_kernel below: 128 x 32 only). Same shared layout and V pass; the H
// pass runs in rounds of 320 items ((TH + 8) / 2 row pairs x 16 groups: two rounds at TH = 64, unrolled, all loads of both rounds issued up front).
// The taller tile is worth +3.4 % on the whole step (118.8 k -> 122.8 k images/s at KITTI, 173.7 k -> 180.0 k at TUM, same box): the kernel is bound
// by load latency, a 32-row CTA lives ~4 us and the launch / drain of such short CTAs held the achieved occupancy at 58 % of a 94 % limit.
// H-pass item:
//   one thread = 8 adjacent pixels of two adjacent rows: one aligned 64-bit load per row plus the 32-bit words left and right of it, i.e.
//   6 loads per 16 pixel-rows instead of 12, and one tile / row / address computation per 16 pixel-rows instead of per 8.
template <bool CLAMP, int TH>
__global__ void __launch_bounds__(BL_THREADS) orb_blur8_kernel(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blr,
                                                               const __grid_constant__ OrbGeom g) {
  __shared__ uint4 s_p[(TH + 8) / 2][BL_TW / 4];   // [row pair][pixel quad]: word j = [h(2k, x+j), h(2k+1, x+j)]
  const int level = orb_level_of(g.blur_base, (int)blockIdx.x);
  const OrbLevel& L = g.lv[level];
  const unsigned t = blockIdx.x - L.blur_tile_base;
  const unsigned ty_ = (t * L.blur_inv_tx) >> 20;            // t / blur_tiles_x (host-checked range)
  const int x0 = (int)(t - ty_ * L.blur_tiles_x) * BL_TW, y0 = (int)ty_ * TH;
  const unsigned long long off = (unsigned long long)blockIdx.y * g.frame_stride + L.plane_off;
  const unsigned tapA = g.blur_k[0], tapB = g.blur_k[1];
#pragma unroll
  for (int item = threadIdx.x; item < (TH + 8) / 2 * 16; item += BL_THREADS) {   // TH = 32: exactly one round
    const int k = item >> 4, q = item & 15;                  // row pair (local rows 2k, 2k+1 <-> plane rows y0 - 4 + ..), 8-pixel group
    const bool need = y0 - 4 + 2 * k <= L.rows + 3;          // pairs past the last row a stored pixel reads (bottom tiles of a level) are not summed;
                                                             // no `break`: the rounds stay unrolled with all their loads issued up front
    const int pitch = L.pitch;
    const uint8_t* src = pyr + off;                          // uniform; the per-thread part is a 32-bit offset
    const int x = x0 + 8 * q;
    // words past the level's apron are never used by a stored pixel: clamp the column so that every load stays inside the row
    const int xl = min(x, (L.cols + 4) & ~7);
    const int ry0 = min(y0 - 4 + 2 * k, L.rows + 3), ry1 = min(y0 - 3 + 2 * k, L.rows + 3);
    const int o0 = ry0 * pitch + xl, o1 = ry1 * pitch + xl;   // ry0 = -4 for the first tile row: the plane's apron rows
    const uint2 w0 = __ldg(reinterpret_cast<const uint2*>(src + o0));
    const uint2 w1 = __ldg(reinterpret_cast<const uint2*>(src + o1));
    // the words left and right of them: loaded as well (independent loads, all six in flight; fetching them from the neighbouring lanes by
    // shuffle measured 1.3 % slower on the whole step: the shuffle can only issue once the load has landed)
    const unsigned l0 = __ldg(reinterpret_cast<const unsigned*>(src + o0) - 1), l1 = __ldg(reinterpret_cast<const unsigned*>(src + o1) - 1);
    const unsigned r0 = __ldg(reinterpret_cast<const unsigned*>(src + o0) + 2), r1 = __ldg(reinterpret_cast<const unsigned*>(src + o1) + 2);
    if (x < L.cols && need) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        unsigned h[2][4];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const uint2 w = rr ? w1 : w0;
          const unsigned c0 = half ? w.x : (rr ? l1 : l0), c1 = half ? w.y : w.x, c2 = half ? (rr ? r1 : r0) : w.y;
          // W(a) = bytes B[a..a+3] of the 12 bytes B[0..11] = pixels x-4 .. x+7 of this half; pixel j: W(j+1).tapA + W(j+5).tapB
          const unsigned w1_ = __byte_perm(c0, c1, 0x4321), w2_ = __byte_perm(c0, c1, 0x5432), w3_ = __byte_perm(c0, c1, 0x6543),
                         w5_ = __byte_perm(c1, c2, 0x4321), w6_ = __byte_perm(c1, c2, 0x5432), w7_ = __byte_perm(c1, c2, 0x6543);
          h[rr][0] = __dp4a(w1_, tapA, __dp4a(w5_, tapB, 0u));
          h[rr][1] = __dp4a(w2_, tapA, __dp4a(w6_, tapB, 0u));
          h[rr][2] = __dp4a(w3_, tapA, __dp4a(w7_, tapB, 0u));
          h[rr][3] = __dp4a(c1, tapA, __dp4a(c2, tapB, 0u));
        }
        s_p[k][2 * q + half] = make_uint4(__byte_perm(h[0][0], h[1][0], 0x5410), __byte_perm(h[0][1], h[1][1], 0x5410),
                                          __byte_perm(h[0][2], h[1][2], 0x5410), __byte_perm(h[0][3], h[1][3], 0x5410));
      }
    }
  }
  __syncthreads();
  blur_v_pass<CLAMP, TH>(s_p, blr + off, L, x0, y0, g);
}

// TMA == true (ORB_B200_BLUR_TMA=1; measured alternative, 7 % slower: the CTA idles for the box's latency, 41 % of the stall samples on the
// mbarrier wait, while the instruction count only drops from 89.1 M to 87.6 M per pass): ONE thread fetches the tile's 160 x 40-byte source window with a cp.async.bulk.tensor box (per-level tensor map of the
// un-blurred planes; bytes outside the plane allocation arrive as zeros and only feed outputs that are not written) and the H pass reads it
// from shared memory. TMA == false (default): every thread loads its 12 source words from global memory, all issued before any arithmetic.
template <bool CLAMP, bool TMA>   // CLAMP: tap sum > 256 (legacy taps sum to 257): saturate like OpenCV's ufixedpoint -> u8 cast
__global__ void __launch_bounds__(BL_THREADS) orb_blur_kernel(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blr,
                                                              const __grid_constant__ OrbGeom g, const __grid_constant__ OrbLevelMaps maps) {
  __shared__ uint4 s_p[(BL_TH + 8) / 2][BL_TW / 4];   // [row pair][pixel quad]: word j = [h(2k, x+j), h(2k+1, x+j)]
  __shared__ __align__(128) unsigned s_raw[TMA ? BL_RAW_H * BL_RAW_W / 4 : 4];
  __shared__ __align__(8) unsigned long long s_bar;
  const int level = orb_level_of(g.blur_base, (int)blockIdx.x);
  const OrbLevel& L = g.lv[level];
  const unsigned t = blockIdx.x - L.blur_tile_base;
  const unsigned ty_ = (t * L.blur_inv_tx) >> 20;            // t / blur_tiles_x (host-checked range)
  const int x0 = (int)(t - ty_ * L.blur_tiles_x) * BL_TW, y0 = (int)ty_ * BL_TH;
  const unsigned long long off = (unsigned long long)blockIdx.y * g.frame_stride + L.plane_off;
  const unsigned tapA = g.blur_k[0], tapB = g.blur_k[1];
  unsigned cw[2][2][3];
  bool act[2];
  if (TMA) {
    const unsigned bar = orb_smem_u32(&s_bar);
    if (threadIdx.x == 0) {
      orb_mbar_init(bar, 1);
      orb_mbar_expect_tx(bar, BL_RAW_H * BL_RAW_W);
      // tensor coordinates = plane coordinates + (16-byte row lead, apron rows): column (x0 - 16) + 16, row (y0 - 4) + 4
      orb_tma_load_3d(orb_smem_u32(s_raw), &maps.m[level], x0, y0, (int)blockIdx.y, bar);
    }
    __syncthreads();                                          // the barrier's initialisation is visible to every waiter
    orb_mbar_wait(bar, 0);
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int i = threadIdx.x + it * BL_THREADS;   // 20 row pairs x 32 quads = 640 items
      const int k = i >> 5, q = i & 31;
      act[it] = x0 + 4 * q < L.cols;
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const unsigned* row = s_raw + (2 * k + rr) * (BL_RAW_W / 4) + 3 + q;   // word holding pixels x - 4 .. x - 1
        cw[it][rr][0] = row[0]; cw[it][rr][1] = row[1]; cw[it][rr][2] = row[2];
      }
    }
  } else {
    const uint8_t* src = pyr + off;
    const int pitch = L.pitch;
    // local row r <-> plane row y0 - 4 + r (row -4 and rows past rows + 3 only ever meet a zero tap)
    // all 12 loads of this thread's two H-pass items (item = 4 pixels x 2 rows) are issued before any arithmetic
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int i = threadIdx.x + it * BL_THREADS;   // 20 row pairs x 32 quads = 640 items
      const int k = i >> 5, q = i & 31;
      const int x = x0 + 4 * q;
      act[it] = x < L.cols;
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const int ry = min(y0 - 4 + 2 * k + rr, L.rows + 3);
        const unsigned* row = reinterpret_cast<const unsigned*>(src + (long long)ry * pitch + x);
        if (act[it]) { cw[it][rr][0] = __ldg(row - 1); cw[it][rr][1] = __ldg(row); cw[it][rr][2] = __ldg(row + 1); }
      }
    }
  }
  const int pitch = L.pitch;
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int i = threadIdx.x + it * BL_THREADS;
    const int k = i >> 5, q = i & 31;
    if (act[it]) {
      unsigned h[2][4];
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const unsigned c0 = cw[it][rr][0], c1 = cw[it][rr][1], c2 = cw[it][rr][2];
        // W(a) = bytes B[a..a+3] of the 12 loaded bytes B[0..11] = pixels x-4 .. x+7; pixel j: W(j+1).tapA + W(j+5).tapB
        const unsigned w1 = __byte_perm(c0, c1, 0x4321), w2 = __byte_perm(c0, c1, 0x5432), w3 = __byte_perm(c0, c1, 0x6543),
                       w5 = __byte_perm(c1, c2, 0x4321), w6 = __byte_perm(c1, c2, 0x5432), w7 = __byte_perm(c1, c2, 0x6543);
        h[rr][0] = __dp4a(w1, tapA, __dp4a(w5, tapB, 0u));
        h[rr][1] = __dp4a(w2, tapA, __dp4a(w6, tapB, 0u));
        h[rr][2] = __dp4a(w3, tapA, __dp4a(w7, tapB, 0u));
        h[rr][3] = __dp4a(c1, tapA, __dp4a(c2, tapB, 0u));
      }
      s_p[k][q] = make_uint4(__byte_perm(h[0][0], h[1][0], 0x5410), __byte_perm(h[0][1], h[1][1], 0x5410),
                             __byte_perm(h[0][2], h[1][2], 0x5410), __byte_perm(h[0][3], h[1][3], 0x5410));
    }
  }
  __syncthreads();
  blur_v_pass<CLAMP>(s_p, blr + off, L, x0, y0, g);
}

// Orientation (un-blurred level), descriptor (blurred level) and the final cv::KeyPoint (pt *= scale for level != 0).
// One warp handles DESC_KPW keypoints of one frame back to back, so the per-lane tables stay in registers; lane i < DESC_KPW owns keypoint i
// of the warp's batch (level lookup and the packed record are fetched once, up front; it writes that keypoint's 28-byte cv::KeyPoint at the end).
//   * staging, MODE 2 (default): the IC_Angle window (un-blurred, 31 rows) and the descriptor patch (blurred, 37 rows) are fetched with 16-byte
//     cp.async copies (global -> shared without a register round trip) from the 16-byte boundary below the window into 48- / 64-byte rows:
//     3 + 5 copy instructions per lane and keypoint. Both windows are issued up front as two groups: the orientation starts when the first
//     has landed. MODE 0 (ORB_B200_DESC_NARROW=1, rounds 1-2): 4-byte copies into tight 36- / 40-byte rows, 11 + 13 instructions per lane -
//     a quarter of the kernel's instructions; the wide copies gave +6.6 % on the whole step (109.1 k -> 116.3 k images/s, same box).
//   * staging, MODE 1 (ORB_B200_DESC_TMA=1, kept as the measured alternative): ONE lane issues two TMA box loads per keypoint
//     (cp.async.bulk.tensor.3d, per-level tensor maps), 31 x 48 bytes and 37 x 64 bytes: the innermost box coordinate has to be 16-byte
//     aligned (anything else raises an illegal-instruction fault, tests/cuda/tma_probe.cu), so a box starts at the 16-byte boundary below
//     the window. 22 % fewer instructions, but 2 x 136 k boxes of ~34 short rows per pass saturate the TMA unit: 1.71 ms against 0.96 ms
//     per 512 frames with 73 % of the stall samples on the mbarrier wait (profiles/r2_tma_ab.md).
//   * IC_Angle: lane = patch row v (31 rows): the row's words are funnel-shifted to the window start; m10 += dp4a(word & disc mask, u),
//     rowsum += dp4a(word & disc mask, 1), m01 = v * rowsum; two REDUX adds finish the moments (exact int32).
//   * rBRIEF: lane = descriptor byte; its 8 test pairs (x0, y0, x1, y1 as floats) come from shared memory, one LDS.128 each. cvRound is
//     the float add of 1.5 * 2^23 (round-half-even happens in the FADD; F2I would go through the quarter-rate conversion pipe); the
//     integer bias of both coordinates is folded into the 32-bit shared-memory base address of the patch.
template <bool WIDE> struct DescLayout;
template <> struct DescLayout<false> { enum { IC_PITCH = 36, PT_PITCH = 40, IC_BYTES = 31 * 36 + 4, PT_BYTES = 37 * 40 + 8 }; };
template <> struct DescLayout<true> { enum { IC_PITCH = 48, PT_PITCH = 64, IC_BYTES = 1536, PT_BYTES = 2432 }; };   // rounded to the 128-byte TMA alignment

__device__ __forceinline__ void cp_async4(unsigned smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem), "l"(gmem));
}
__device__ __forceinline__ void cp_async16(unsigned smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem), "l"(gmem));
}

// MODE 0: 4-byte cp.async into tight rows; 1: TMA boxes; 2: 16-byte cp.async into the 16-byte-aligned rows of the TMA layout
template <int MODE>
__global__ void __launch_bounds__(DESC_WARPS * 32, 40 / DESC_WARPS)
orb_describe_kernel(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blr, const uint32_t* __restrict__ kept,
                    const int* __restrict__ kept_cnt, OrbKeypoint* __restrict__ kps, uint8_t* __restrict__ desc, int cap,
                    int* __restrict__ counts, int* __restrict__ err_flag, const __grid_constant__ OrbGeom g,
                    const __grid_constant__ OrbLevelMaps ic_maps, const __grid_constant__ OrbLevelMaps pt_maps) {
  constexpr bool TMA = MODE == 1, WIDE = MODE != 0;
  typedef DescLayout<WIDE> Lay;
  __shared__ __align__(128) uint8_t s_win[DESC_WARPS][Lay::IC_BYTES + Lay::PT_BYTES];
  __shared__ float4 s_pat[8 * 32];   // rBRIEF pattern as floats, [test bit][lane]: one conflict-free LDS.128 per test (256 test pairs)
  __shared__ __align__(8) unsigned long long s_bar[DESC_WARPS][2];
  const int frame = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // cumulative kept counts: lane l holds the inclusive prefix of level l
  int cinc = lane < g.nlevels ? kept_cnt[frame * g.nlevels + lane] : 0;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, cinc, o);
    if (lane >= o) cinc += y;
  }
  const int total = __shfl_sync(0xffffffffu, cinc, g.nlevels - 1);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    counts[frame] = min(total, cap);
    if (total > cap) atomicOr(err_flag, 8);
  }
  const int w_first = (blockIdx.x * DESC_WARPS + warp) * DESC_KPW;
  const int n = min(min(w_first + DESC_KPW, total), cap) - w_first;   // keypoints of this warp
#pragma unroll
  for (int b = warp; b < 8; b += DESC_WARPS) {
    const char4 p = reinterpret_cast<const char4*>(d_pattern)[lane * 8 + b];
    s_pat[b * 32 + lane] = make_float4((float)p.x, (float)p.y, (float)p.z, (float)p.w);
  }
  const unsigned bar_ic = orb_smem_u32(&s_bar[warp][0]), bar_pt = orb_smem_u32(&s_bar[warp][1]);
  if (TMA && lane == 0) { orb_mbar_init(bar_ic, 1); orb_mbar_init(bar_pt, 1); }
  __syncthreads();
  if (n <= 0) return;

  // lane i < n: level and packed record of keypoint w_first + i (level = number of levels whose inclusive prefix is <= w)
  int my_level = 0;
  uint32_t my_q = 0;
  {
    const int w = w_first + lane;
    int base = 0;
    for (int l = 0; l < g.nlevels - 1; ++l) {
      const int c = __shfl_sync(0xffffffffu, cinc, l);
      if (c <= w) { my_level = l + 1; base = c; }
    }
    if (lane < n) my_q = kept[(long long)frame * g.node_total + g.lv[my_level].node_off + (w - base)];
  }
  float my_angle = 0.f;

  // per-lane constants
  const int v = lane - ORB_HALF_PATCH;                   // patch row of this lane (lane 31: none)
  unsigned dm[8];                                        // byte mask of the disc in this lane's row, words u = -16 + 4i .. + 3
  {
    const int hw = lane < 31 ? c_umax[v < 0 ? -v : v] : -1;   // half width of the disc in this row (:866-910, symmetric)
    // one bit per byte u = -16..15: bits [16 - hw, 16 + hw]; each nibble is then widened to a byte mask
    const unsigned bits = hw >= 0 ? ((2u << (2 * hw)) - 1u) << (16 - hw) : 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) dm[i] = ((((bits >> (4 * i)) & 15u) * 0x00204081u) & 0x01010101u) * 0xffu;
  }
  const uint8_t* ic = s_win[warp];
  const unsigned ic_s = orb_smem_u32(ic), pt_s = ic_s + Lay::IC_BYTES;
  // patch byte (r, c) relative to the keypoint sits at pt_s + (r + 18) * pitch + (c + 18) + (window misalignment); r and c arrive as float
  // bit patterns 0x4B400000 + value
  const unsigned pt_bias = pt_s + 18u * Lay::PT_PITCH + 18u - ((unsigned)Lay::PT_PITCH + 1u) * 0x4B400000u;
  unsigned failed = 0;
  // cp.async staging: three window rows per copy instruction, lane = (row % 3, word)
  const int ic_sub = lane / 9, ic_w = lane - 9 * ic_sub;      // lanes 27..31 idle
  const int pt_sub = lane / 10, pt_w = lane - 10 * pt_sub;    // lanes 30, 31 idle
  const bool ic_act = lane < 27, pt_act = lane < 30;
  const unsigned ic_dst = ic_s + 4u * (unsigned)(9 * ic_sub + ic_w), pt_dst = pt_s + 4u * (unsigned)(10 * pt_sub + pt_w);

  for (int i = 0; i < n; ++i) {
    const int level = __shfl_sync(0xffffffffu, my_level, i);
    const uint32_t q = __shfl_sync(0xffffffffu, my_q, i);
    const int x = (int)(q & 0xfffu) + ORB_MIN_BORDER, y = (int)((q >> 12) & 0xfffu) + ORB_MIN_BORDER;   // :1573-1589
    __syncwarp();                                              // the previous keypoint's gathers are done
    const unsigned parity = (unsigned)i & 1u;
    unsigned s_ic, s_pt;                                       // byte offset of window column 0 inside the staged row
    if (TMA) {
      if (lane == 0) {
        // tensor coordinates = level coordinates + (16-byte row lead, apron rows); planes are 16-byte aligned, so column alignment == address alignment
        orb_mbar_expect_tx(bar_ic, 31 * Lay::IC_PITCH);
        orb_tma_load_3d(ic_s, &ic_maps.m[level], ((x - 16) & ~15) + 16, y - 15 + ORB_APRON_ROWS, frame, bar_ic);
        orb_mbar_expect_tx(bar_pt, 37 * Lay::PT_PITCH);
        orb_tma_load_3d(pt_s, &pt_maps.m[level], ((x - 18) & ~15) + 16, y - 18 + ORB_APRON_ROWS, frame, bar_pt);
      }
      s_ic = (unsigned)(x - 16) & 15u;
      s_pt = (unsigned)(x - 18) & 15u;
      failed |= orb_mbar_wait(bar_ic, parity) ^ 1u;
    } else if (MODE == 2) {
      // 31 x 3 and 37 x 4 sixteen-byte chunks: lane + 32 k -> (row, chunk); 8 copy instructions per lane instead of 24
      const OrbLevel& L = g.lv[level];
      const int pitch = L.pitch;
      const unsigned long long off = (unsigned long long)frame * g.frame_stride + L.plane_off;
      const uint8_t* icb = pyr + off + (long long)(y - 15) * pitch + ((x - 16) & ~15);
      const uint8_t* ptb = blr + off + (long long)(y - 18) * pitch + ((x - 18) & ~15) + (lane >> 2) * pitch + 16 * (lane & 3);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int idx = lane + 32 * k, row = idx / 3, ch = idx - 3 * row;
        if (idx < 93) cp_async16(ic_s + (unsigned)(row * Lay::IC_PITCH + 16 * ch), icb + row * pitch + 16 * ch);
      }
      asm volatile("cp.async.commit_group;");
      const unsigned pt_d = pt_s + (unsigned)((lane >> 2) * Lay::PT_PITCH + 16 * (lane & 3));
#pragma unroll
      for (int k = 0; k < 5; ++k)
        if (k < 4 || lane < 20) cp_async16(pt_d + (unsigned)(8 * k * Lay::PT_PITCH), ptb + 8 * k * pitch);
      asm volatile("cp.async.commit_group;");
      s_ic = (unsigned)(x - 16) & 15u;
      s_pt = (unsigned)(x - 18) & 15u;
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncwarp();
    } else {
      const OrbLevel& L = g.lv[level];
      const int pitch = L.pitch;
      const unsigned long long off = (unsigned long long)frame * g.frame_stride + L.plane_off;
      const int xi = (x - 16) & ~3, xp = (x - 18) & ~3;      // planes are 16-byte aligned: column alignment == address alignment
      const uint8_t* icl = pyr + off + (long long)(y - 15 + ic_sub) * pitch + xi + 4 * ic_w;
      const uint8_t* ptl = blr + off + (long long)(y - 18 + pt_sub) * pitch + xp + 4 * pt_w;
      const long long p3 = 3LL * pitch;
#pragma unroll
      for (int k = 0; k < 11; ++k)
        if (ic_act && 3 * k + ic_sub < 31) cp_async4(ic_dst + 108u * k, icl + k * p3);
      asm volatile("cp.async.commit_group;");
#pragma unroll
      for (int k = 0; k < 13; ++k)
        if (pt_act && 3 * k + pt_sub < 37) cp_async4(pt_dst + 120u * k, ptl + k * p3);
      asm volatile("cp.async.commit_group;");
      s_ic = (unsigned)(x - 16) & 3u;
      s_pt = (unsigned)(x - 18) & 3u;
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncwarp();
    }
    // IC_Angle (:104-161): lane = row v, window column c <-> u = c - 16
    int m10 = 0, rs = 0;
    if (lane < 31) {
      unsigned w[12];
      if (WIDE) {
        const uint4* rowp = reinterpret_cast<const uint4*>(ic + lane * Lay::IC_PITCH);
        const uint4 q0 = rowp[0], q1 = rowp[1], q2 = rowp[2];
        w[0] = q0.x; w[1] = q0.y; w[2] = q0.z; w[3] = q0.w; w[4] = q1.x; w[5] = q1.y; w[6] = q1.z; w[7] = q1.w;
        w[8] = q2.x; w[9] = q2.y; w[10] = q2.z; w[11] = q2.w;
        if (s_ic & 8u) {
#pragma unroll
          for (int k = 0; k < 10; ++k) w[k] = w[k + 2];
        }
        if (s_ic & 4u) {
#pragma unroll
          for (int k = 0; k < 9; ++k) w[k] = w[k + 1];
        }
      } else {
        const unsigned* rowp = reinterpret_cast<const unsigned*>(ic + lane * Lay::IC_PITCH);
#pragma unroll
        for (int k = 0; k < 9; ++k) w[k] = rowp[k];
      }
      const unsigned sh = 8u * (s_ic & 3u);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const unsigned wv = __funnelshift_r(w[k], w[k + 1], sh) & dm[k];
        const int u0 = -16 + 4 * k;                       // signed byte coefficients u0 .. u0 + 3 (compile-time constant)
        const unsigned cu = ((unsigned)u0 & 0xffu) | (((unsigned)(u0 + 1) & 0xffu) << 8) | (((unsigned)(u0 + 2) & 0xffu) << 16) |
                            (((unsigned)(u0 + 3) & 0xffu) << 24);
        m10 = dp4a_u8_s8(wv, cu, m10);
        rs = (int)__dp4a(wv, 0x01010101u, (unsigned)rs);
      }
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    const int m01 = __reduce_add_sync(0xffffffffu, v * rs);
    const float angle = orb_fast_atan2((float)m01, (float)m10);
    if (lane == i) my_angle = angle;
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    float a, b;
    orb_sincos(__fmul_rn(angle, factorPI), &b, &a);
    // computeOrbDescriptor (:177-254) on the staged patch
    if (TMA) {
      failed |= orb_mbar_wait(bar_pt, parity) ^ 1u;
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
    }
    const unsigned pt_off = pt_bias + s_pt;
    const float magic = 12582912.f;   // 1.5 * 2^23: (v + magic) has round-half-even(v) in its low mantissa bits, bit pattern 0x4B400000 + round(v)
    int val = 0;
#pragma unroll
    for (int bit = 0; bit < 8; ++bit) {
      const float4 p = s_pat[bit * 32 + lane];   // x0, y0, x1, y1
      const unsigned r0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)), magic));
      const unsigned c0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)), magic));
      const unsigned r1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)), magic));
      const unsigned c1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)), magic));
      unsigned t0, t1;
      asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t0) : "r"(r0 * (unsigned)Lay::PT_PITCH + c0 + pt_off));
      asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t1) : "r"(r1 * (unsigned)Lay::PT_PITCH + c1 + pt_off));
      val |= (t0 < t1) << bit;
    }
    desc[((long long)frame * cap + w_first + i) * 32 + lane] = (uint8_t)val;
  }
  if (failed && lane == 0) atomicOr(err_flag, 16);
  if (lane < n) {   // the 28-byte cv::KeyPoint of the lane's keypoint
    const OrbLevel& L = g.lv[my_level];
    const int x = (int)(my_q & 0xfffu) + ORB_MIN_BORDER, y = (int)((my_q >> 12) & 0xfffu) + ORB_MIN_BORDER;
    float* o = reinterpret_cast<float*>(kps + (long long)frame * cap + w_first + lane);
    o[0] = my_level ? __fmul_rn((float)x, L.scale) : (float)x;
    o[1] = my_level ? __fmul_rn((float)y, L.scale) : (float)y;
    o[2] = L.patch_size;
    o[3] = my_angle;
    o[4] = (float)(my_q >> 24);
    o[5] = __int_as_float(my_level);
    o[6] = __int_as_float(-1);
  }
}

// maps: per-level TMA descriptors of pyr with a 160 x 40-byte box, used with ORB_B200_BLUR_TMA=1 only
void orb_launch_blur(const uint8_t* pyr, uint8_t* blr, const void* maps, const OrbGeom& g, int batch, cudaStream_t st, long long* launches) {
  static const bool want_tma = getenv("ORB_B200_BLUR_TMA") != nullptr;   // measured alternative, off by default (profiles/r2_tma_ab.md)
  static const OrbLevelMaps no_maps = {};
  int sum = g.blur_taps[3];
  for (int i = 0; i < 3; ++i) sum += 2 * g.blur_taps[i];
  const dim3 grid(g.blur_tiles_total, batch);
  const bool tma = maps != nullptr && want_tma;
  const OrbLevelMaps& M = tma ? *reinterpret_cast<const OrbLevelMaps*>(maps) : no_maps;
  static const bool v1 = getenv("ORB_B200_BLUR_V1") != nullptr;   // the 4-pixel H pass of rounds 1-2 (A/B; needs ORB_B200_BLUR_TH=32 as well)
  if (g.blur_th == 128) {
    if (sum > 256) orb_blur8_kernel<true, 128><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g);
    else orb_blur8_kernel<false, 128><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g);
  } else if (g.blur_th == 64) {   // the tile table was built for this tile height, only this kernel reads it that way
    if (sum > 256) orb_blur8_kernel<true, 64><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g);
    else orb_blur8_kernel<false, 64><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g);
  } else if (!tma && !v1) {   // ORB_B200_BLUR_TH=32
    if (sum > 256) orb_blur8_kernel<true, 32><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g);
    else orb_blur8_kernel<false, 32><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g);
  } else if (sum > 256) {
    if (tma) orb_blur_kernel<true, true><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g, M);
    else orb_blur_kernel<true, false><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g, M);
  } else {
    if (tma) orb_blur_kernel<false, true><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g, M);
    else orb_blur_kernel<false, false><<<grid, BL_THREADS, 0, st>>>(pyr, blr, g, M);
  }
  ++*launches;
}

// ic_maps / pt_maps: per-level TMA descriptors of pyr (48 B x 31 rows) and blr (64 B x 37 rows), used only with ORB_B200_DESC_TMA=1 (the measured
// alternative, see the kernel comment); the default stages with cp.async.
void orb_launch_describe(const uint8_t* pyr, const uint8_t* blr, const void* ic_maps, const void* pt_maps, const uint32_t* kept, const int* kept_cnt,
                         OrbKeypoint* kps, uint8_t* desc, int cap, int* counts, int* err_flag, const OrbGeom& g, int batch, cudaStream_t st,
                         long long* launches) {
  static const bool want_tma = getenv("ORB_B200_DESC_TMA") != nullptr;
  static const OrbLevelMaps no_maps = {};
  ++*launches;
  const int per_cta = DESC_WARPS * DESC_KPW;
  const dim3 grid((cap + per_cta - 1) / per_cta, batch);
  static const bool want_wide = getenv("ORB_B200_DESC_NARROW") == nullptr;   // default; ORB_B200_DESC_NARROW=1: the 4-byte copies of rounds 1-2 (A/B)
  if (want_tma && ic_maps && pt_maps)
    orb_describe_kernel<1><<<grid, DESC_WARPS * 32, 0, st>>>(pyr, blr, kept, kept_cnt, kps, desc, cap, counts, err_flag, g,
                                                             *reinterpret_cast<const OrbLevelMaps*>(ic_maps),
                                                             *reinterpret_cast<const OrbLevelMaps*>(pt_maps));
  else if (want_wide)
    orb_describe_kernel<2><<<grid, DESC_WARPS * 32, 0, st>>>(pyr, blr, kept, kept_cnt, kps, desc, cap, counts, err_flag, g, no_maps, no_maps);
  else
    orb_describe_kernel<0><<<grid, DESC_WARPS * 32, 0, st>>>(pyr, blr, kept, kept_cnt, kps, desc, cap, counts, err_flag, g, no_maps, no_maps);
}
