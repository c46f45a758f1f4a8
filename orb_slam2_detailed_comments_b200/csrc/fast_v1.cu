// ROUND-1 KERNEL, kept for one same-box A/B (ORB_B200_FAST_V1=1) against the round-2 kernel in fast.cu; not the default path.
// Per-cell FAST-9/16 with the iniThFAST / minThFAST retry: the cell loop of
// ORBextractor::ComputeKeyPointsOctTree (src/ORBextractor.cc:1465-1548) and cv::FAST(cell, th, true).
//
// cornerScore is threshold-free (S = max(A,B) - 1) and "corner at T" <=> S >= T, so ONE score pass serves both
// thresholds. The detection domains of the reference's 30-px cells ([ini+3, ini+cell+3) in x and y) tile the
// level without overlap, so S is computed once per pixel; only the 3x3 NMS is per cell (neighbours outside the
// cell's detection window count as 0, because the reference runs cv::FAST on the cell sub-image). A cell emits
// its NMS survivors with S >= ini, or, if there are none, those with S >= min, row-major (cv::FAST's order).
//
// One CTA = one cell row x up to fast_G cells (a band of <= 248 x hCell detection pixels):
//   1. the band window (+3 px halo) is staged in shared memory with 16-byte loads;
//   2. quick reject, 4 pixels per thread with byte-SIMD (VABSDIFF4 + carry-free byte compares): a 9-arc always
//      holds one of {up, down} and one of {left, right}, so min(max|d_ud|, max|d_lr|) > minTh is necessary for
//      S >= minTh; a thread keeps its survivors as bit masks and one warp scan compacts them into the warp's candidate list;
//   3. candidates are scored 32 at a time with full lanes. The dark and the bright arc tests share one instruction
//      stream: ring pixel r_k becomes the 16-bit lane pair [256 + v - r_k, 256 - v + r_k] (one IMAD) and
//      max_arcs(min_9) of both lanes is 16 + 16 VIMNMX3.U16x2 + an 8-op max tree. Corners (S >= minTh) are
//      compacted in place to the front of the warp's list;
//   4. each corner does the 3x3 NMS on the shared score plane (cell-window masked) and sets its bit in the
//      min / ini survivor bit planes;
//   5. per cell (one warp each, lane = row): 64-bit row masks from the bit planes, popc + warp scan give every
//      survivor its row-major rank; ini/min selection by the cell's ini count.
#include <algorithm>
#include <atomic>
#include "orb_common.cuh"

#define FAST_THREADS 256
#define FAST_WARPS 8
#define FAST_WP 256      // window row pitch in smem: 16 B front pad + 15 alignment + ORB_FAST_TILE_W (219) + 6 halo = 256
#define FAST_SP 240      // score plane row pitch: 15 + 219 + 4 <= 240
#define FAST_BP 11       // survivor bit plane row pitch in words (window columns < 288 -> 9 words, +2 for the 64-bit extract)

// cornerScore (OpenCV fast_score.cpp, 16-pixel ring, K = 8): S = max(max_arcs min_9(v - r), max_arcs min_9(r - v)) - 1,
// clamped at 0. Lanes of every 32-bit value: lo = 256 + (v - r), hi = 256 + (r - v), both in [1, 511].
__device__ __forceinline__ int fast1_score(const uint8_t* __restrict__ p) {
  constexpr int wp = FAST_WP;
  const unsigned v = p[0];
  const unsigned base = (256u + v) | ((256u - v) << 16);
  unsigned e[16];
  // ring offsets (dx,dy), k = 0..15: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
  e[0] = p[3 * wp];       e[1] = p[3 * wp + 1];   e[2] = p[2 * wp + 2];   e[3] = p[wp + 3];
  e[4] = p[3];            e[5] = p[-wp + 3];      e[6] = p[-2 * wp + 2];  e[7] = p[-3 * wp + 1];
  e[8] = p[-3 * wp];      e[9] = p[-3 * wp - 1];  e[10] = p[-2 * wp - 2]; e[11] = p[-wp - 3];
  e[12] = p[-3];          e[13] = p[wp - 3];      e[14] = p[2 * wp - 2];  e[15] = p[3 * wp - 1];
#pragma unroll
  for (int k = 0; k < 16; ++k) e[k] = base + e[k] * 0xffffu;   // (r << 16) - r: no borrow, lo >= 1
  unsigned m3[16], m9[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) m3[k] = __vimin3_u16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
#pragma unroll
  for (int k = 0; k < 16; ++k) m9[k] = __vimin3_u16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
  const unsigned a0 = __vimax3_u16x2(m9[0], m9[1], m9[2]), a1 = __vimax3_u16x2(m9[3], m9[4], m9[5]),
                 a2 = __vimax3_u16x2(m9[6], m9[7], m9[8]), a3 = __vimax3_u16x2(m9[9], m9[10], m9[11]),
                 a4 = __vimax3_u16x2(m9[12], m9[13], m9[14]);
  const unsigned b = __vimax3_u16x2(__vimax3_u16x2(a0, a1, a2), __vimax3_u16x2(a3, a4, m9[15]), 0u);
  const int s = (int)max(b & 0xffffu, b >> 16) - 257;
  return max(s, 0);
}

// per-byte a > t (bit 7 of each result byte), t <= 126: ((a & 0x7f) | 0x80) - (t + 1) never borrows across bytes
__device__ __forceinline__ unsigned fast1_gt4(unsigned a, unsigned t1) {
  return (((a & 0x7f7f7f7fu) | 0x80808080u) - t1) | a;
}

// Quick reject + compaction + scoring for the rows row0, row0 + FAST_WARPS, row0 + 2 * FAST_WARPS, ... < row1 of one warp (rows are
// dealt round-robin to the warps: a textured blob spans adjacent rows, so every warp gets a share of its candidates and the
// barrier after the scoring does not wait for one loaded warp).
//   SECOND == false: candidates = pixels passing the compass test at iniTh (superset of all S >= iniTh);
//   SECOND == true : candidates = pixels passing at minTh but not at iniTh (not scored yet), in the columns marked in `need`.
// Candidate codes go to list[start ..], corners (S >= minTh) are compacted in place to list[start ..] as (y << 9) | x and
// their score is written to the score plane. Returns the number of corners appended.
template <bool SMALL_TH, bool SECOND>
__device__ __forceinline__ int fast1_detect(const uint8_t* __restrict__ win, uint8_t* __restrict__ sc, const uint8_t* __restrict__ need,
                                           uint16_t* __restrict__ list, int start, int row0, int row1, int qbase, int nq, int lane,
                                           int ini_th, int min_th) {
  const unsigned thi4 = (unsigned)ini_th * 0x01010101u, ti1 = thi4 + 0x01010101u;
  const unsigned thm4 = (unsigned)min_th * 0x01010101u, tm1 = thm4 + 0x01010101u;
  const uint32_t lt = (1u << lane) - 1u;
  // candidate bits of this thread: byte (ri & 3) of cand[ri >> 2] = the 8 pixels of group (lane + 5*ri) & 31 in row row0 + 8*ri
  // (a tile row is <= 29 aligned groups of 8 pixels: one pass per row). The per-row rotation spreads a textured blob
  // (candidates cluster in 2-D) over many lanes.
  unsigned cand[2] = {0u, 0u};
  for (int ri = 0; row0 + FAST_WARPS * ri < row1; ++ri) {
    const int y = row0 + FAST_WARPS * ri;
    const int o = (lane + 5 * ri) & 31;
    if (o < nq) {
      const int x = qbase + 8 * o;
      const uint2 colmask = *reinterpret_cast<const uint2*>(need + x);   // bit 7 of byte k: column x + k takes part
      if (colmask.x | colmask.y) {
        const uint8_t* row = win + (y + 3) * FAST_WP + 16 + x;
        const unsigned c0 = *reinterpret_cast<const unsigned*>(row - 4), c3 = *reinterpret_cast<const unsigned*>(row + 8);
        const uint2 c = *reinterpret_cast<const uint2*>(row);
        const uint2 up = *reinterpret_cast<const uint2*>(row - 3 * FAST_WP), dn = *reinterpret_cast<const uint2*>(row + 3 * FAST_WP);
        unsigned pass[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const unsigned c1 = h ? c.y : c.x;
          const unsigned lf = h ? __byte_perm(c.x, c.y, 0x4321) : __byte_perm(c0, c.x, 0x4321);
          const unsigned rt = h ? __byte_perm(c.y, c3, 0x6543) : __byte_perm(c.x, c.y, 0x6543);
          const unsigned au = __vabsdiffu4(c1, h ? up.y : up.x), ad = __vabsdiffu4(c1, h ? dn.y : dn.x), al = __vabsdiffu4(c1, lf),
                         ar = __vabsdiffu4(c1, rt);
          unsigned ps;
          if (SMALL_TH) {
            ps = (fast1_gt4(au, SECOND ? tm1 : ti1) | fast1_gt4(ad, SECOND ? tm1 : ti1)) &
                 (fast1_gt4(al, SECOND ? tm1 : ti1) | fast1_gt4(ar, SECOND ? tm1 : ti1));
            if (SECOND) ps &= ~((fast1_gt4(au, ti1) | fast1_gt4(ad, ti1)) & (fast1_gt4(al, ti1) | fast1_gt4(ar, ti1)));
          } else {
            ps = (__vcmpgtu4(au, SECOND ? thm4 : thi4) | __vcmpgtu4(ad, SECOND ? thm4 : thi4)) &
                 (__vcmpgtu4(al, SECOND ? thm4 : thi4) | __vcmpgtu4(ar, SECOND ? thm4 : thi4));
            if (SECOND) ps &= ~((__vcmpgtu4(au, thi4) | __vcmpgtu4(ad, thi4)) & (__vcmpgtu4(al, thi4) | __vcmpgtu4(ar, thi4)));
          }
          pass[h] = ps & (h ? colmask.y : colmask.x) & 0x80808080u;
        }
        // bits 7, 15, 23, 31 of each half -> one nibble each (the partial products never carry)
        const unsigned bits8 = ((pass[0] * 0x204081u) >> 28) | (((pass[1] * 0x204081u) >> 28) << 4);
        const unsigned b = bits8 << ((8 * ri) & 31);
        if (ri < 4) cand[0] |= b; else cand[1] |= b;
      }
    }
  }
  // one warp scan places every thread's candidates in the warp's list (order is irrelevant). Entry = lane:5 | ri:3 | k:3,
  // decoded by full warps in the scoring loop
  int cnt;
  {
    const int c = __popc(cand[0]) + __popc(cand[1]);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int yv = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += yv;
    }
    cnt = __shfl_sync(0xffffffffu, incl, 31);
    uint16_t* out = list + start + (incl - c);
#pragma unroll
    for (int hi = 0; hi < 2; ++hi) {
      unsigned m = cand[hi];
      const unsigned code = (unsigned)((lane << 6) | (hi << 5));
      while (m) {
        *out++ = (uint16_t)(code | (unsigned)(__ffs(m) - 1));
        m &= m - 1;
      }
    }
  }
  __syncwarp();
  // score the candidates, 32 per step
  int ncorner = 0;
  for (int c0 = 0; c0 < cnt; c0 += 32) {
    const int i = c0 + lane;
    int e = 0, s = 0;
    if (i < cnt) {
      const int code = list[start + i];
      const int ri = (code >> 3) & 7;
      const int o = ((code >> 6) + 5 * ri) & 31;
      e = ((row0 + FAST_WARPS * ri) << 9) | (qbase + 8 * o + (code & 7));
      s = fast1_score(win + ((e >> 9) + 3) * FAST_WP + 16 + (e & 511));
    }
    const bool corner = s >= min_th;
    __syncwarp();                                            // the reads of this chunk precede the in-place writes below
    const uint32_t b = __ballot_sync(0xffffffffu, corner);
    if (corner) {
      sc[((e >> 9) + 1) * FAST_SP + (e & 511)] = (uint8_t)s;
      list[start + ncorner + __popc(b & lt)] = (uint16_t)e;
    }
    ncorner += __popc(b);
  }
  return ncorner;
}

// 3x3 NMS of the corners list[0 .. n) against their neighbours inside the same cell's detection window.
//   SECOND == false: corners with S >= iniTh -> plane `pl` (ini survivors);
//   SECOND == true : corners in the columns marked in `need` (cells that found no ini survivor) -> plane `pl` (min survivors).
template <bool SECOND>
__device__ __forceinline__ void fast1_nms(const uint8_t* __restrict__ sc, const uint8_t* __restrict__ need, const uint16_t* __restrict__ list,
                                         int n, int lane, uint32_t* __restrict__ pl, int wCell, unsigned inv, int xa, int xlim, int ini_th) {
  for (int i = lane; i < n; i += 32) {
    const int e = list[i];
    const int y = e >> 9, x = e & 511;
    const uint8_t* q = sc + (y + 1) * FAST_SP + x;
    const int s = q[0];
    if (SECOND ? (need[x] == 0) : (s < ini_th)) continue;
    const unsigned p = (unsigned)(x + xa - ORB_MIN_BORDER - 3);
    const unsigned r = p - __umulhi(p, inv) * (unsigned)wCell;   // column inside the cell's detection window (inv = 2^32 / wCell + 1)
    bool ok = s > q[-FAST_SP] && s > q[FAST_SP];
    if (r > 0) ok = ok && s > q[-1] && s > q[-FAST_SP - 1] && s > q[FAST_SP - 1];
    if ((int)r < wCell - 1 && x + 1 < xlim) ok = ok && s > q[1] && s > q[-FAST_SP + 1] && s > q[FAST_SP + 1];
    if (ok) atomicOr(&pl[y * FAST_BP + (x >> 5)], 1u << (x & 31));
  }
}

template <bool SMALL_TH>   // SMALL_TH: iniTh <= 126 (carry-free byte compares); otherwise the generic video intrinsic
__global__ void __launch_bounds__(FAST_THREADS)
orb_fast_band_v1_kernel(const uint8_t* __restrict__ pyr, int* __restrict__ cell_cnt, uint32_t* __restrict__ cell_kp,
                     int* __restrict__ err_flag, int win_rows, int sc_rows, int row_cap, const __grid_constant__ OrbGeom g) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* win = smem;                                    // [win_rows][FAST_WP], pixel x at column 16 + x
  uint8_t* sc = win + win_rows * FAST_WP;                 // [sc_rows][FAST_SP], score of domain row y at row y + 1
  uint32_t* pl_min = reinterpret_cast<uint32_t*>(sc + sc_rows * FAST_SP);   // [sc_rows - 2][FAST_BP] NMS survivors, S >= minTh
  uint32_t* pl_ini = pl_min + (sc_rows - 2) * FAST_BP;                       //                         S >= iniTh
  uint16_t* lists = reinterpret_cast<uint16_t*>(pl_ini + (sc_rows - 2) * FAST_BP);   // [max_h][row_cap], a warp's rows are contiguous
  uint8_t* need = reinterpret_cast<uint8_t*>(lists + (sc_rows - 2) * row_cap);        // [FAST_WP] 0x80 = column takes part in the pass
  __shared__ int s_retry;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int frame = blockIdx.y;
  const int level = orb_level_of(g.fast_base, (int)blockIdx.x);
  const OrbLevel& L = g.lv[level];
  const int t = blockIdx.x - L.fast_tile_base;
  const int ci = (int)(((unsigned)t * L.fast_inv_nseg) >> 20), sg = t - ci * L.fast_nseg;   // t / fast_nseg
  const int j0 = sg * L.fast_G, j1 = min(j0 + L.fast_G, L.nCols);
  const int wCell = L.wCell;
  int* cnt_row = cell_cnt + (long long)frame * g.ncells_total + L.cell_base + ci * L.nCols;
  uint32_t* kp_row = cell_kp + ((long long)frame * g.ncells_total + L.cell_base + ci * L.nCols) * g.cell_cap;

  const int iniY = ORB_MIN_BORDER + ci * L.hCell;
  const int wh = min(iniY + L.hCell + 6, L.maxBY) - iniY, dh = wh - 6;   // window rows, detection rows
  const int X0 = ORB_MIN_BORDER + j0 * wCell;                            // window [X0, X1)
  const int X1 = min(ORB_MIN_BORDER + j1 * wCell + 6, L.maxBX);
  if (iniY >= L.maxBY - 3 || dh <= 0 || X1 - X0 <= 6) {                  // :1479-1503 skip rules / empty domain
    for (int j = j0 + tid; j < j1; j += FAST_THREADS) cnt_row[j] = 0;
    return;
  }
  const int xa = X0 & ~15;                         // 16-byte aligned global column of window column 0
  const int nvec = (X1 - xa + 15) >> 4;
  const uint8_t* src = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off + (long long)iniY * L.pitch + xa;
  // 1. stage the window; clear the score plane and the survivor bit planes
  // i / nvec == (i * inv) >> 16 for i < 65 * 18 holds for any inv in [65536 / nvec, 65536 / nvec + 2]: one MUFU.RCP instead of a division
  const unsigned inv_nvec = (unsigned)(__frcp_rn((float)nvec) * 65536.f) + 1u;
  for (int i = tid; i < wh * nvec; i += FAST_THREADS) {
    const int y = (int)(((unsigned)i * inv_nvec) >> 16), v = i - y * nvec;
    // global -> shared directly (LDGSTS.128); the copies land while the planes below are cleared
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(win + y * FAST_WP + 16 + 16 * v)),
                 "l"(src + (long long)y * L.pitch + 16 * v));
  }
  asm volatile("cp.async.commit_group;");
  for (int i = tid; i < (dh + 2) * (FAST_SP / 16); i += FAST_THREADS) reinterpret_cast<uint4*>(sc)[i] = make_uint4(0, 0, 0, 0);
  // both bit planes, 16 bytes per thread (the rounding may run a few words into the still empty lists)
  for (int i = tid; i < (2 * (sc_rows - 2) * FAST_BP + 3) / 4; i += FAST_THREADS) reinterpret_cast<uint4*>(pl_min)[i] = make_uint4(0, 0, 0, 0);
  const int dx0 = X0 + 3 - xa, dx1 = X1 - 3 - xa;       // detection columns in window coordinates
  for (int i = tid; i < FAST_WP; i += FAST_THREADS) need[i] = (i >= dx0 && i < dx1) ? 0x80 : 0;   // pass 1: every detection column
  if (tid == 0) s_retry = 0;
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  // 2. first detection at iniTh (the reference's first cv::FAST call per cell, :1510)
  const int qbase = dx0 & ~7;                            // groups of 8 aligned pixels
  const int nq = (dx1 - qbase + 7) >> 3;                 // <= 29
  // warp w owns max_h / 8 (+ 1 if w < max_h % 8) rows of the largest level: its list starts after the rows of the warps before it
  uint16_t* list = lists + (warp * ((sc_rows - 2) >> 3) + min(warp, (sc_rows - 2) & 7)) * row_cap;
  // warp w owns the detection rows w, w + 8, w + 16, ... (<= 8 rows, dh <= 64)
  const int row0 = warp, row1 = dh;
  const int xlim = L.maxBX - 3 - xa;                     // detection columns end here (last cell may be clipped)
  int ncorner = fast1_detect<SMALL_TH, false>(win, sc, need, list, 0, row0, row1, qbase, nq, lane, g.ini_th, g.min_th);
  __syncthreads();
  // 3. NMS of the corners with S >= iniTh -> ini survivor plane. A suppressor of such a corner has S >= iniTh itself, so it was
  //    scored above. `need` is re-initialised meanwhile (nobody reads it until the next barrier).
  fast1_nms<false>(sc, need, list, ncorner, lane, pl_ini, wCell, L.inv_wcell, xa, xlim, g.ini_th);
  for (int i = tid; i < FAST_WP / 4; i += FAST_THREADS) reinterpret_cast<unsigned*>(need)[i] = 0u;
  __syncthreads();
  // 4. cells without an ini survivor retry at minTh (:1521): mark their detection columns
  for (int cj = j0 + warp; cj < j1; cj += FAST_WARPS) {
    const int cx0 = ORB_MIN_BORDER + cj * wCell + 3 - xa;
    const int cx1 = min(ORB_MIN_BORDER + (cj + 1) * wCell + 3, L.maxBX - 3) - xa;
    const int dw = cx1 - cx0;
    if (dw <= 0) continue;
    const unsigned long long wmask = dw >= 64 ? ~0ull : ((1ull << dw) - 1ull);
    const int w0 = cx0 >> 5, sh = cx0 & 31;
    bool any = false;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int y = lane + 32 * r;
      if (y < dh) {
        const uint32_t* b = pl_ini + y * FAST_BP + w0;
        const unsigned long long m = (((unsigned long long)__funnelshift_r(b[1], b[2], sh) << 32) | __funnelshift_r(b[0], b[1], sh)) & wmask;
        any = any || m != 0ull;
      }
    }
    if (!__any_sync(0xffffffffu, any)) {
      if (lane < dw) need[cx0 + lane] = 0x80;
      if (lane + 32 < dw) need[cx0 + lane + 32] = 0x80;
      if (lane == 0) s_retry = 1;
    }
  }
  __syncthreads();
  if (s_retry) {   // uniform over the CTA
    // 4a. the pixels that pass the compass test at minTh but not at iniTh, in the retry cells only
    const int n2 = fast1_detect<SMALL_TH, true>(win, sc, need, list, ncorner, row0, row1, qbase, nq, lane, g.ini_th, g.min_th);
    ncorner += n2;
    __syncthreads();
    // 4b. NMS of every corner (S >= minTh) of the retry cells -> min survivor plane
    fast1_nms<true>(sc, need, list, ncorner, lane, pl_min, wCell, L.inv_wcell, xa, xlim, g.ini_th);
    __syncthreads();
  }

  // 5. per cell: row-major ranks from the survivor bit planes, ini/min selection, emission. Lane = detection row
  //    (two row slots cover dh <= 64); a cell's detection window is < 64 columns wide.
  for (int cj = j0 + warp; cj < j1; cj += FAST_WARPS) {
    const int cx0 = ORB_MIN_BORDER + cj * wCell + 3 - xa;
    const int cx1 = min(ORB_MIN_BORDER + (cj + 1) * wCell + 3, L.maxBX - 3) - xa;
    const int dw = cx1 - cx0;
    if (dw <= 0) {
      if (lane == 0) cnt_row[cj] = 0;
      continue;
    }
    const unsigned long long wmask = dw >= 64 ? ~0ull : ((1ull << dw) - 1ull);
    const int w0 = cx0 >> 5, sh = cx0 & 31;
    unsigned long long m_min[2], m_ini[2];
    int n_ini = 0;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int y = lane + 32 * r;
      m_min[r] = 0ull; m_ini[r] = 0ull;
      if (y < dh) {
        const uint32_t* a = pl_min + y * FAST_BP + w0;
        const uint32_t* b = pl_ini + y * FAST_BP + w0;
        m_min[r] = (((unsigned long long)__funnelshift_r(a[1], a[2], sh) << 32) | __funnelshift_r(a[0], a[1], sh)) & wmask;
        m_ini[r] = (((unsigned long long)__funnelshift_r(b[1], b[2], sh) << 32) | __funnelshift_r(b[0], b[1], sh)) & wmask;
        n_ini += __popcll(m_ini[r]);
      }
    }
    n_ini = __reduce_add_sync(0xffffffffu, n_ini);
    const unsigned long long m0 = n_ini > 0 ? m_ini[0] : m_min[0], m1 = n_ini > 0 ? m_ini[1] : m_min[1];
    const int c0 = __popcll(m0), c1 = __popcll(m1);
    // exclusive scan over rows 0..63 (slot 0 of all lanes, then slot 1)
    int x0 = c0, x1 = c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y0 = __shfl_up_sync(0xffffffffu, x0, o), y1 = __shfl_up_sync(0xffffffffu, x1, o);
      if (lane >= o) { x0 += y0; x1 += y1; }
    }
    const int tot0 = __shfl_sync(0xffffffffu, x0, 31), tot1 = __shfl_sync(0xffffffffu, x1, 31);
    const int total = tot0 + tot1;
    uint32_t* kp_out = kp_row + (long long)cj * g.cell_cap;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      unsigned long long m = r ? m1 : m0;
      int rank = r ? tot0 + x1 - c1 : x0 - c0;
      const int y = lane + 32 * r;
      while (m) {
        const int bx = __ffsll((long long)m) - 1;
        m &= m - 1;
        const int x = cx0 + bx;
        if (rank < g.cell_cap) {
          const uint32_t s = sc[(y + 1) * FAST_SP + x];
          // cell-local FAST coordinates shifted by (j*wCell, i*hCell) (:1540-1541) == level pixel - 16
          kp_out[rank] = (uint32_t)(x + xa - ORB_MIN_BORDER) | ((uint32_t)(y + 3 + ci * L.hCell) << 12) | (s << 24);
        }
        ++rank;
      }
    }
    if (lane == 0) {
      cnt_row[cj] = min(total, g.cell_cap);
      if (total > g.cell_cap) atomicOr(err_flag, 1);
    }
  }
}

void orb_launch_fast_v1(const uint8_t* pyr, int* cell_cnt, uint32_t* cell_kp, int* err_flag, const OrbGeom& g, int batch,
                     cudaStream_t st, long long* launches) {
  int max_h = 0;
  for (int l = 0; l < g.nlevels; ++l) max_h = max(max_h, g.lv[l].hCell);
  const int win_rows = max_h + 6, sc_rows = max_h + 2;
  // a warp's list holds one entry per detection pixel of its rows at most: rows x (aligned quads of the widest tile) x 4
  int row_cap = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    const int dw = std::min(g.lv[l].fast_G * g.lv[l].wCell, g.lv[l].maxBX - ORB_MIN_BORDER - 6);   // detection columns of a tile
    row_cap = std::max(row_cap, 4 * ((dw + 3 + 3) / 4));                                           // + quad alignment at the front
  }
  row_cap = std::min(row_cap, 256);
  const size_t smem = (size_t)win_rows * FAST_WP + (size_t)sc_rows * FAST_SP + 2 * (size_t)max_h * FAST_BP * 4 +
                      (size_t)max_h * row_cap * 2 + FAST_WP;   // KITTI: 40 KB -> 5 resident CTAs per SM
  // The opt-in shared memory size is a per-context function attribute: raise it once per device to the sm_100 maximum
  // (227 KB minus the kernel's static shared memory), so that concurrent callers with different level sizes never lower each other's limit.
  typedef void (*Kern)(const uint8_t*, int*, uint32_t*, int*, int, int, int, const OrbGeom);
  static const Kern kerns[2] = {orb_fast_band_v1_kernel<false>, orb_fast_band_v1_kernel<true>};
  static std::atomic<bool> configured[64][2];
  const int small = g.ini_th <= 126;      // both thresholds (min <= ini) in the carry-free byte-compare range
  const Kern kern = kerns[small];
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!configured[dev][small].load(std::memory_order_acquire)) {
    int optin = 0;
    cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);   // the limit covers static + dynamic
    configured[dev][small].store(true, std::memory_order_release);
  }
  const dim3 grid(g.fast_tiles_total, batch);
  kern<<<grid, FAST_THREADS, smem, st>>>(pyr, cell_cnt, cell_kp, err_flag, win_rows, sc_rows, row_cap, g);
  ++*launches;
}
