// Per-cell FAST-9/16 with the iniThFAST / minThFAST retry: the cell loop of
// ORBextractor::ComputeKeyPointsOctTree (src/ORBextractor.cc:1465-1548) and cv::FAST(cell, th, true).
//
// cornerScore is threshold-free (S = max(A,B) - 1) and "corner at T" <=> S >= T, so ONE score serves both thresholds.
// The detection domains of the reference's 30-px cells ([ini+3, ini+cell+3) in x and y) tile the level without
// overlap, so S is computed once per pixel; only the 3x3 NMS is per cell (neighbours outside the cell's detection
// window count as 0, because the reference runs cv::FAST on the cell sub-image). A cell emits its NMS survivors with
// S >= ini, or, if there are none, those with S >= min, row-major (cv::FAST's order).
//
// One CTA = one cell row x up to fast_G cells (a band of <= 219 x hCell detection pixels), 256 threads:
//   1. ONE thread issues a TMA box load (cp.async.bulk.tensor.3d, per-level tensor map over [frame][row][byte]) of the
//      band window (+3 px halo, 256-byte rows) into shared memory and arms an mbarrier; the others meanwhile clear the
//      score plane / survivor planes and build the column tables (cell index, cell-edge flags, detection mask);
//   2. ONE sweep over the window, 8 pixels per thread with byte-SIMD (VABSDIFF4 + carry-free byte compares): a 9-arc
//      always holds one of {up, down} and one of {left, right}, so min(max|d_ud|, max|d_lr|) > T is necessary for
//      S >= T. The four |d| words are compared against BOTH thresholds while they are in registers: a thread keeps
//      "passes at iniTh" and "passes at minTh only" as two bit masks (<= 64 bits each: its <= 8 rows x 8 pixels);
//   3. the iniTh candidates are compacted with one warp scan and scored 32 at a time with full lanes. The dark and the
//      bright arc tests share one instruction stream: ring pixel r_k becomes the 16-bit lane pair
//      [256 + v - r_k, 256 - v + r_k] (one IMAD) and max_arcs(min_9) of both lanes is 16 + 16 VIMNMX3.U16x2 + an
//      8-op max tree. Corners (S >= minTh) are compacted in place to the front of the warp's list;
//   4. 3x3 NMS of the corners with S >= iniTh (cell-edge flags from the column table) -> per-cell 64-bit survivor rows
//      + a per-cell "has an ini survivor" flag;
//   5. if a cell of the band has no ini survivor (:1521), the minTh-only bits of step 2 inside those cells are
//      compacted and scored (no second sweep), and the NMS runs over every corner of those cells;
//   6. per cell (one warp each, lane = row pair): popc + warp scan give every survivor its row-major rank.
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "orb_common.cuh"
#include "orb_tma.cuh"

#define FAST_THREADS 256
#define FAST_WARPS 8
#define FAST_WP 256      // window row pitch in smem: 16 B front pad + 15 alignment + ORB_FAST_TILE_W (219) + 6 halo = 256
#define FAST_SP 256      // score plane row pitch = FAST_WP: the score of the pixel at window byte offset a lives at sc[a - FAST_SC_OFF]
#define FAST_SC_OFF (2 * FAST_WP + 16)   // window offset (y + 3) * 256 + 16 + x  ->  score plane offset (y + 1) * 256 + x
#define FAST_TAB_BYTES (2 * FAST_WP + 32)   // host-built per-band tables: coltab[256] u16 + det8[32] u8
#define FAST_MAXC 8      // cells per band (219 / 30 = 7) rounded up

// cornerScore (OpenCV fast_score.cpp, 16-pixel ring, K = 8): S = max(max_arcs min_9(v - r), max_arcs min_9(r - v)) - 1,
// clamped at 0. Lanes of every 32-bit value: lo = 256 + (v - r), hi = 256 + (r - v), both in [1, 511].
__device__ __forceinline__ int fast_score(unsigned p) {   // p = 32-bit shared-memory address of the candidate pixel in the window
  constexpr int wp = FAST_WP;
  unsigned v, e[16];
  asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(p));
  const unsigned base = (256u + v) | ((256u - v) << 16);
  // ring offsets (dx,dy), k = 0..15: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
  asm("ld.shared.u8 %0, [%1+768];" : "=r"(e[0]) : "r"(p));   asm("ld.shared.u8 %0, [%1+769];" : "=r"(e[1]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+514];" : "=r"(e[2]) : "r"(p));   asm("ld.shared.u8 %0, [%1+259];" : "=r"(e[3]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+3];" : "=r"(e[4]) : "r"(p));     asm("ld.shared.u8 %0, [%1+-253];" : "=r"(e[5]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+-510];" : "=r"(e[6]) : "r"(p));  asm("ld.shared.u8 %0, [%1+-767];" : "=r"(e[7]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+-768];" : "=r"(e[8]) : "r"(p));  asm("ld.shared.u8 %0, [%1+-769];" : "=r"(e[9]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+-514];" : "=r"(e[10]) : "r"(p)); asm("ld.shared.u8 %0, [%1+-259];" : "=r"(e[11]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+-3];" : "=r"(e[12]) : "r"(p));   asm("ld.shared.u8 %0, [%1+253];" : "=r"(e[13]) : "r"(p));
  asm("ld.shared.u8 %0, [%1+510];" : "=r"(e[14]) : "r"(p));  asm("ld.shared.u8 %0, [%1+767];" : "=r"(e[15]) : "r"(p));
  static_assert(wp == 256, "the immediates above are ring offsets for a 256-byte window pitch");
#pragma unroll
  for (int k = 0; k < 16; ++k) e[k] = base + e[k] * 0xffffu;   // (r << 16) - r: no borrow, lo >= 1
  // max over the 16 arcs of min over 9 = max over even k of min(min(e[k+1 .. k+8]), max(e[k], e[k+9])):
  // arcs k and k+1 share e[k+1 .. k+8]
  unsigned p2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) p2[j] = __vimin3_u16x2(e[2 * j + 1], e[(2 * j + 2) & 15], 0xffffffffu);   // min(e[2j+1], e[2j+2])
  unsigned M[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {   // even k = 2j: shared part = p2[j], p2[j+1], p2[j+2], p2[j+3]
    const unsigned t = __vimin3_u16x2(p2[j], p2[(j + 1) & 7], p2[(j + 2) & 7]);
    const unsigned u = __vimax3_u16x2(e[2 * j], e[(2 * j + 9) & 15], 0u);
    M[j] = __vimin3_u16x2(t, p2[(j + 3) & 7], u);
  }
  const unsigned a0 = __vimax3_u16x2(M[0], M[1], M[2]), a1 = __vimax3_u16x2(M[3], M[4], M[5]);
  const unsigned b = __vimax3_u16x2(__vimax3_u16x2(a0, a1, M[6]), M[7], 0u);
  const int s = (int)max(b & 0xffffu, b >> 16) - 257;
  return max(s, 0);
}

// bit 7 of each byte: a > t. SMALL (t <= 126): A = (a & 0x7f) | 0x80 is shared by both thresholds, (A - (t + 1)) | a never borrows
// across bytes. Otherwise the generic video intrinsic.
template <bool SMALL_TH>
__device__ __forceinline__ void fast_gt2(unsigned a, unsigned ti, unsigned tm, unsigned& gi, unsigned& gm) {
  if (SMALL_TH) {
    const unsigned A = (a & 0x7f7f7f7fu) | 0x80808080u;
    gi = A - ti;   // ti, tm = (t + 1) * 0x01010101
    gm = A - tm;
  } else {
    gi = __vcmpgtu4(a, ti);   // ti, tm = t * 0x01010101
    gm = __vcmpgtu4(a, tm);
  }
}

// The candidates of one warp (bit masks cand[2] of every lane: byte (ri & 3) of cand[ri >> 2] = the 8 pixels of group
// (lane + 5 * ri) & 31 in row warp + 8 * ri) go to list[start ..] as codes lane:5 | bit:6, are scored 32 at a time, and the corners
// (S >= minTh) are compacted in place to list[start ..] as their byte offset in the window, a = (y + 3) * 256 + 16 + x, with their score
// written to the score plane at a - FAST_SC_OFF (same pitch: NMS and emission address both planes from a). Returns the number of corners appended.
__device__ __forceinline__ int fast_score_candidates(const unsigned (&cand)[2], const uint8_t* __restrict__ win, uint8_t* __restrict__ sc,
                                                     uint16_t* __restrict__ list, int start, int warp, int qbase, int lane, int min_th) {
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
        *out++ = (uint16_t)(code | (unsigned)(__ffs(m) - 1));   // lane:5 | bit:6 (this loop is divergent: the decoding is left to the full-lane loop below)
        m &= m - 1;
      }
    }
  }
  // pad the list to a multiple of 32 with copies of its first entry (the lists have 32 spare entries per warp): the score loop then needs no
  // bounds branch around fast_score, only `i < cnt` in the corner predicate
  if (cnt & 31) {
    const int i = (cnt & ~31) + lane;
    if (i >= cnt) list[start + i] = list[start];
  }
  __syncwarp();
  // offset of pixel 0 of group 0 in row `warp`; row warp + 8 * ri is 8 * 256 bytes further per ri, group o 8 bytes per o
  const unsigned a0 = (unsigned)((warp + 3) * FAST_WP + 16 + qbase);
  const unsigned win_s = orb_smem_u32(win);
  unsigned lt;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(lt));
  int ncorner = 0;
  for (int c0 = 0; c0 < cnt; c0 += 32) {   // branch-free body: the compiler unrolls it by four, two chunks' loads in flight (+0.3 % over `unroll 1`)
    const int i = c0 + lane;
    const unsigned code = list[start + i];
    const unsigned ri = (code >> 3) & 7u;
    const unsigned a = a0 + ri * (8u * FAST_WP) + 8u * (((code >> 6) + 5u * ri) & 31u) + (code & 7u);
    const int s = fast_score(win_s + a);
    const bool corner = s >= min_th && i < cnt;
    __syncwarp();                                            // the reads of this chunk precede the in-place writes below
    const uint32_t b = __ballot_sync(0xffffffffu, corner);
    if (corner) {
      sc[a - FAST_SC_OFF] = (uint8_t)s;
      list[start + ncorner + __popc(b & lt)] = (uint16_t)a;
    }
    ncorner += __popc(b);
  }
  return ncorner;
}

// 3x3 NMS of the corners list[0 .. n) against their neighbours inside the same cell's detection window (coltab: cell:3 | left edge:1 |
// right edge:1 | column inside the cell:6 per window column). Survivors set their bit in the cell's 64-bit row of plane `pl`.
//   SECOND == false: corners with S >= iniTh, and the cell is flagged in hasini;
//   SECOND == true : corners of the cells without an ini survivor.
template <bool SECOND>
__device__ __forceinline__ void fast_nms(const uint8_t* __restrict__ sc, const uint16_t* __restrict__ coltab, uint8_t* __restrict__ hasini,
                                         const uint16_t* __restrict__ list, int n, int lane, unsigned long long* __restrict__ pl, int prow,
                                         int ini_th) {
  for (int i = lane; i < n; i += 32) {
    const unsigned a = list[i];                      // window offset (y + 3) * 256 + 16 + x
    const uint8_t* q = sc + a - FAST_SC_OFF;
    const int s = q[0];
    const unsigned ct = coltab[a & (FAST_WP - 1)];
    const int cl = ct & 7;
    if (SECOND ? (hasini[cl] != 0) : (s < ini_th)) continue;
    int m = max((int)q[-FAST_SP], (int)q[FAST_SP]);
    if (!(ct & 8u)) m = max(m, __vimax3_s32((int)q[-1], (int)q[-FAST_SP - 1], (int)q[FAST_SP - 1]));
    if (!(ct & 16u)) m = max(m, __vimax3_s32((int)q[1], (int)q[-FAST_SP + 1], (int)q[FAST_SP + 1]));
    if (s > m) {
      const unsigned r = ct >> 5;
      const int y = (int)(a >> 8) - 3;
      atomicOr(reinterpret_cast<unsigned*>(pl + cl * prow + y) + (r >> 5), 1u << (r & 31));
      if (!SECOND) hasini[cl] = 1;
    }
  }
}

template <bool SMALL_TH, bool TMA>   // SMALL_TH: iniTh <= 126 (carry-free byte compares); TMA: window by one bulk tensor copy, else per-thread cp.async
__global__ void __launch_bounds__(FAST_THREADS)
orb_fast_band_kernel(const uint8_t* __restrict__ pyr, const int4* __restrict__ tiles, int* __restrict__ cell_cnt,
                     uint32_t* __restrict__ cell_kp, int* __restrict__ err_flag, int win_rows, int sc_rows, int row_cap,
                     const __grid_constant__ OrbGeom g, const __grid_constant__ OrbLevelMaps maps) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int prow = (sc_rows - 1) & ~1;                    // survivor rows per cell, even (sc_rows - 2 rounded up)
  uint8_t* win = smem;                                    // [win_rows][FAST_WP], pixel x at column 16 + x
  uint8_t* sc = win + win_rows * FAST_WP;                 // [sc_rows][FAST_SP], score of domain row y at row y + 1
  // [FAST_MAXC][prow] NMS survivors, one 64-bit row per cell row. ONE plane serves both thresholds: a cell that retries at minTh has no
  // iniTh survivor, i.e. its rows are still all zero when step 5 writes them
  unsigned long long* pl_ini = reinterpret_cast<unsigned long long*>(sc + sc_rows * FAST_SP);
  unsigned long long* pl_min = pl_ini;
  uint16_t* coltab = reinterpret_cast<uint16_t*>(pl_ini + FAST_MAXC * prow);   // [256] per window column: cell:3 | left:1 | right:1 | column in cell:6; 0xffff outside
  uint8_t* det8 = reinterpret_cast<uint8_t*>(coltab + FAST_WP);   // [32] per 8-pixel group: detection columns (coltab + det8 = the band's host-built table)
  uint8_t* need8 = det8 + 32;                             // [32] per 8-pixel group: columns of the cells that retry at minTh
  uint8_t* hasini = need8 + 32;                           // [FAST_MAXC] cell has an NMS survivor with S >= iniTh
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(hasini + FAST_MAXC);
  uint16_t* lists = reinterpret_cast<uint16_t*>(mbar + 1);   // [max_h][row_cap] window offsets, a warp's rows are contiguous
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int frame = blockIdx.y;
  // the band's geometry depends on (level, cell row, first cell) only: two int4 per band, built once per shape on the host (orb_fast_build_tiles)
  const int4 T = __ldg(tiles + 2 * blockIdx.x), U = __ldg(tiles + 2 * blockIdx.x + 1);
  const int level = T.x & 255, j0 = T.y, j1 = T.y + ((T.x >> 8) & 255);
  // the column tables depend on (level, first cell) only, stored behind the band table
  const uint8_t* tab = reinterpret_cast<const uint8_t*>(tiles + 2 * g.fast_tiles_total) + (size_t)((unsigned)T.x >> 16) * FAST_TAB_BYTES;
  const OrbLevel& L = g.lv[level];
  const int wCell = L.wCell;
  const long long cell0 = (long long)frame * g.ncells_total + T.z;      // + L.cell_base + cell row * L.nCols
  int* cnt_row = cell_cnt + cell0;
  uint32_t* kp_row = cell_kp + cell0 * g.cell_cap;

  const int iniY = T.w;                                // ORB_MIN_BORDER + cell row * hCell
  const int wh = U.z & 255, dh = wh - 6;               // window rows, detection rows
  if (U.z >> 16) {                                     // :1479-1503 skip rules / empty domain
    for (int j = j0 + tid; j < j1; j += FAST_THREADS) cnt_row[j] = 0;
    return;
  }
  const int xa = U.x;                              // 16-byte aligned level column of window column 16
  // 1. stage the window
  if (TMA) {
    if (tid == 0) {
      const unsigned bar = orb_smem_u32(mbar);
      if (orb_smem_u32(win) & 127u) atomicOr(err_flag, 16);   // TMA destination alignment
      orb_mbar_init(bar, 1);
      // the box is the level's full window height (hCell + 6 rows x 256 bytes); rows / bytes outside the plane arrive as zeros
      orb_mbar_expect_tx(bar, (unsigned)((((U.z >> 8) & 255) + 6) * FAST_WP + FAST_TAB_BYTES));
      // tensor coordinates: byte column of the plane allocation (16-byte lead + level column xa - 16), row incl. the apron rows, frame
      orb_tma_load_3d(orb_smem_u32(win), &maps.m[level], xa, iniY + ORB_APRON_ROWS, frame, bar);
      orb_bulk_load_1d(orb_smem_u32(coltab), tab, FAST_TAB_BYTES, bar);   // column tables: same barrier, no thread instructions
    }
  } else {
    const int nvec = ((U.y >> 16) + 3 + 15) >> 4;   // window columns [xa, X1), X1 - xa = dx1 + 3
    const uint8_t* src = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off + (long long)iniY * L.pitch + xa;
    // i / nvec == (i * inv) >> 16 for i < 65 * 18 holds for any inv in [65536 / nvec, 65536 / nvec + 2]: one MUFU.RCP instead of a division
    const unsigned inv_nvec = (unsigned)(__frcp_rn((float)nvec) * 65536.f) + 1u;
    for (int i = tid; i < wh * nvec; i += FAST_THREADS) {
      const int y = (int)(((unsigned)i * inv_nvec) >> 16), v = i - y * nvec;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(orb_smem_u32(win + y * FAST_WP + 16 + 16 * v)),
                   "l"(src + (long long)y * L.pitch + 16 * v));
    }
    if (tid < FAST_TAB_BYTES / 16)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(orb_smem_u32(coltab) + 16 * tid), "l"(tab + 16 * tid));
    asm volatile("cp.async.commit_group;");
  }
  // ... meanwhile: clear the score plane and the survivor plane (contiguous)
  {
    uint4* z = reinterpret_cast<uint4*>(sc);
    const int nz = (sc_rows * FAST_SP + FAST_MAXC * prow * 8) >> 4;
    for (int i = tid; i < nz; i += FAST_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    if (tid < FAST_MAXC) hasini[tid] = 0;
  }
  const int dx0 = U.y & 0xffff, dx1 = U.y >> 16;        // detection columns in window coordinates (window column 16 + x)
  const int qbase = dx0 & ~7;                            // groups of 8 aligned pixels
  const int nq = (dx1 - qbase + 7) >> 3;                 // <= 29
  if (!TMA) asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (TMA) {
    const unsigned bar = orb_smem_u32(mbar);
    const unsigned done = orb_mbar_wait(bar, 0);   // bounded: a lost copy becomes ORB_E_CAPACITY flag 16 instead of a hang
    if (!done && tid == 0) atomicOr(err_flag, 16);
  }

  // 2. one sweep, both thresholds (the reference's two cv::FAST calls per cell, :1510 and :1521, differ only in the threshold).
  //    Warp w owns the detection rows w, w + 8, w + 16, ... (<= 8 rows, dh <= 64): a textured blob spans adjacent rows, so every warp
  //    gets a share of its candidates. The per-row lane rotation spreads it over the lanes as well.
  unsigned cand_i[2] = {0u, 0u}, cand_m[2] = {0u, 0u};
  {
    const unsigned ti = SMALL_TH ? ((unsigned)g.ini_th + 1u) * 0x01010101u : (unsigned)g.ini_th * 0x01010101u;
    const unsigned tm = SMALL_TH ? ((unsigned)g.min_th + 1u) * 0x01010101u : (unsigned)g.min_th * 0x01010101u;
#pragma unroll
    for (int ri = 0; ri < 8; ++ri) {   // unrolled: ri, the lane rotation, the mask word and its shift are compile-time constants
      if (warp + FAST_WARPS * ri >= dh) break;
      const int y = warp + FAST_WARPS * ri;
      const int o = (lane + 5 * ri) & 31;
      if (o < nq) {
        const uint8_t* row = win + (y + 3) * FAST_WP + 16 + qbase + 8 * o;
        const unsigned c0 = *reinterpret_cast<const unsigned*>(row - 4), c3 = *reinterpret_cast<const unsigned*>(row + 8);
        const uint2 c = *reinterpret_cast<const uint2*>(row);
        const uint2 up = *reinterpret_cast<const uint2*>(row - 3 * FAST_WP), dn = *reinterpret_cast<const uint2*>(row + 3 * FAST_WP);
        unsigned bi = 0u, bm = 0u;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const unsigned c1 = h ? c.y : c.x;
          const unsigned lf = h ? __byte_perm(c.x, c.y, 0x4321) : __byte_perm(c0, c.x, 0x4321);
          const unsigned rt = h ? __byte_perm(c.y, c3, 0x6543) : __byte_perm(c.x, c.y, 0x6543);
          const unsigned au = __vabsdiffu4(c1, h ? up.y : up.x), ad = __vabsdiffu4(c1, h ? dn.y : dn.x), al = __vabsdiffu4(c1, lf),
                         ar = __vabsdiffu4(c1, rt);
          unsigned ui, um, di, dm, li, lm, ri_, rm;
          fast_gt2<SMALL_TH>(au, ti, tm, ui, um);
          fast_gt2<SMALL_TH>(ad, ti, tm, di, dm);
          fast_gt2<SMALL_TH>(al, ti, tm, li, lm);
          fast_gt2<SMALL_TH>(ar, ti, tm, ri_, rm);
          unsigned gi, gm;
          if (SMALL_TH) {
            const unsigned pud = au | ad, plr = al | ar;
            gi = (ui | di | pud) & (li | ri_ | plr) & 0x80808080u;
            gm = (um | dm | pud) & (lm | rm | plr) & 0x80808080u;
          } else {
            gi = (ui | di) & (li | ri_) & 0x80808080u;
            gm = (um | dm) & (lm | rm) & 0x80808080u;
          }
          // bits 7, 15, 23, 31 -> one nibble (the partial products never carry)
          bi |= ((gi * 0x204081u) >> 28) << (4 * h);
          bm |= ((gm * 0x204081u) >> 28) << (4 * h);
        }
        const unsigned d8 = det8[o];
        bi &= d8;
        bm = (bm ^ bi) & d8;                              // passes at minTh but not at iniTh (ini >= min: bi is a subset of bm)
        const int sh = (8 * ri) & 31;
        if (ri < 4) { cand_i[0] |= bi << sh; cand_m[0] |= bm << sh; }
        else        { cand_i[1] |= bi << sh; cand_m[1] |= bm << sh; }
      }
    }
  }
  // warp w owns max_h / 8 (+ 1 if w < max_h % 8) rows of the largest level: its list starts after the rows of the warps before it
  uint16_t* list = lists + (warp * ((sc_rows - 2) >> 3) + min(warp, (sc_rows - 2) & 7)) * row_cap + 32 * warp;   // + 32 spare entries per warp (padding)
  // 3. the iniTh candidates (superset of all S >= iniTh): exact scores, corners = S >= minTh
  int ncorner = fast_score_candidates(cand_i, win, sc, list, 0, warp, qbase, lane, g.min_th);
  __syncthreads();
  // 4. NMS of the corners with S >= iniTh. A suppressor of such a corner has S >= iniTh itself, so it was scored above.
  fast_nms<false>(sc, coltab, hasini, list, ncorner, lane, pl_ini, prow, g.ini_th);
  __syncthreads();
  // 5. cells without an ini survivor retry at minTh (:1521)
  int retry;
  {
    const unsigned ct = coltab[tid];
    const bool nd = ct != 0xffffu && hasini[ct & 7] == 0;
    const unsigned b = __ballot_sync(0xffffffffu, nd);
    const int o = 4 * warp + lane - 2 - (qbase >> 3);
    if (lane < 4 && o >= 0 && o < 32) need8[o] = (uint8_t)(b >> (8 * lane));
    retry = __syncthreads_or(nd);
  }
  if (retry) {   // uniform over the CTA
    // 5a. the pixels that passed the compass test at minTh but not at iniTh, in the retry cells only
    for (int ri = 0; warp + FAST_WARPS * ri < dh; ++ri) {
      const unsigned keep = ~((0xffu ^ (unsigned)need8[(lane + 5 * ri) & 31]) << ((8 * ri) & 31));
      if (ri < 4) cand_m[0] &= keep; else cand_m[1] &= keep;
    }
    ncorner += fast_score_candidates(cand_m, win, sc, list, ncorner, warp, qbase, lane, g.min_th);
    __syncthreads();
    // 5b. NMS of every corner (S >= minTh) of the retry cells
    fast_nms<true>(sc, coltab, hasini, list, ncorner, lane, pl_min, prow, g.ini_th);
    __syncthreads();
  }

  // 6. per cell: row-major ranks from the cell's survivor rows (ini plane if the cell has an ini survivor), emission.
  //    Lane = detection rows 2 * lane, 2 * lane + 1 (dh <= 64); rows >= dh hold no bits.
  for (int cj = j0 + warp; cj < j1; cj += FAST_WARPS) {
    const int cl = cj - j0;
    const int cx0 = ORB_MIN_BORDER + cj * wCell + 3 - xa;
    if (min(ORB_MIN_BORDER + (cj + 1) * wCell + 3, L.maxBX - 3) - xa - cx0 <= 0) {   // clipped away: no detection column
      if (lane == 0) cnt_row[cj] = 0;
      continue;
    }
    const unsigned long long* pc = pl_ini + cl * prow;
    ulonglong2 mm = make_ulonglong2(0ull, 0ull);
    if (2 * lane < prow) mm = *reinterpret_cast<const ulonglong2*>(pc + 2 * lane);
    const int c0 = __popcll(mm.x), c1 = __popcll(mm.y);
    int incl = c0 + c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int yv = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += yv;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t* kp_out = kp_row + (long long)cj * g.cell_cap;
    int rank = incl - c0 - c1;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      unsigned long long m = r ? mm.y : mm.x;
      const int y = 2 * lane + r;
      while (m) {
        const int bx = __ffsll((long long)m) - 1;
        m &= m - 1;
        const int x = cx0 + bx;
        if (rank < g.cell_cap) {
          const uint32_t s = sc[(y + 1) * FAST_SP + x];
          // cell-local FAST coordinates shifted by (j*wCell, i*hCell) (:1540-1541) == level pixel - 16
          kp_out[rank] = (uint32_t)(x + xa - ORB_MIN_BORDER) | ((uint32_t)(y + 3 + iniY - ORB_MIN_BORDER) << 12) | (s << 24);
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

// ---- host side ------------------------------------------------------------------------------------------------
// Band table: two int4 per CTA of a frame, {level | cells << 8 | table << 16, first cell, first cell's index in the frame's cell arrays, window row iniY}
// and {xa, dx0 | dx1 << 16, window rows | hCell << 8 | skip << 16, 0}, levels in ascending order, followed by the column
// tables (FAST_TAB_BYTES each, one per (level, segment)): per window column cell:3 | left edge:1 | right edge:1 | column inside the cell:6
// (0xffff outside the detection columns), then per 8-pixel group the mask of detection columns.
void orb_fast_build_tiles(const OrbGeom& g, std::vector<int4>& tiles) {
  tiles.clear();
  std::vector<uint8_t> tabs;
  int ntab = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    const OrbLevel& L = g.lv[l];
    for (int sg = 0; sg < L.fast_nseg; ++sg) {
      const int j0 = sg * L.fast_G, j1 = std::min(j0 + L.fast_G, L.nCols);
      tabs.resize((size_t)(ntab + sg + 1) * FAST_TAB_BYTES);
      uint16_t* coltab = reinterpret_cast<uint16_t*>(tabs.data() + (size_t)(ntab + sg) * FAST_TAB_BYTES);
      uint8_t* det8 = reinterpret_cast<uint8_t*>(coltab + FAST_WP);
      const int X0 = ORB_MIN_BORDER + j0 * L.wCell, X1 = std::min(ORB_MIN_BORDER + j1 * L.wCell + 6, L.maxBX);
      const int xa = X0 & ~15, dx0 = X0 + 3 - xa, dx1 = X1 - 3 - xa, qbase = dx0 & ~7;
      memset(det8, 0, 32);
      for (int t = 0; t < FAST_WP; ++t) {
        const int x = t - 16;
        unsigned ct = 0xffffu;
        if (x >= dx0 && x < dx1) {
          const int p = x + xa - ORB_MIN_BORDER - 3;   // detection column of the level
          const int cell = p / L.wCell, r = p - cell * L.wCell;
          ct = (unsigned)(cell - j0) | (r == 0 ? 8u : 0u) | ((r == L.wCell - 1 || x == dx1 - 1) ? 16u : 0u) | ((unsigned)r << 5);
          const int o = (x - qbase) >> 3;
          if (o >= 0 && o < 32) det8[o] |= (uint8_t)(1u << ((x - qbase) & 7));
        }
        coltab[t] = (uint16_t)ct;
      }
    }
    for (int ci = 0; ci < L.nRows; ++ci)
      for (int sg = 0; sg < L.fast_nseg; ++sg) {
        const int j0 = sg * L.fast_G, j1 = std::min(j0 + L.fast_G, L.nCols);
        const int iniY = ORB_MIN_BORDER + ci * L.hCell;
        const int wh = std::min(iniY + L.hCell + 6, L.maxBY) - iniY;                                    // window rows; detection rows = wh - 6
        const int X0 = ORB_MIN_BORDER + j0 * L.wCell, X1 = std::min(ORB_MIN_BORDER + j1 * L.wCell + 6, L.maxBX);   // window [X0, X1)
        const int xa = X0 & ~15;
        const bool skip = iniY >= L.maxBY - 3 || wh - 6 <= 0 || X1 - X0 <= 6;
        tiles.push_back(make_int4(l | (std::max(j1 - j0, 0) << 8) | ((ntab + sg) << 16), j0, L.cell_base + ci * L.nCols, iniY));
        tiles.push_back(make_int4(xa, skip ? 0 : (X0 + 3 - xa) | ((X1 - 3 - xa) << 16), (skip ? 0 : wh) | (L.hCell << 8) | ((skip ? 1 : 0) << 16), 0));
      }
    ntab += L.fast_nseg;
  }
  const size_t nt = tiles.size();
  tiles.resize(nt + (tabs.size() + 15) / 16);
  if (!tabs.empty()) memcpy(reinterpret_cast<uint8_t*>(tiles.data() + nt), tabs.data(), tabs.size());
}

void orb_launch_fast(const uint8_t* pyr, const void* tiles, const void* maps, int* cell_cnt, uint32_t* cell_kp, int* err_flag, const OrbGeom& g,
                     int batch, cudaStream_t st, long long* launches) {
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
  const int prow = (sc_rows - 1) & ~1;
  const size_t smem = (size_t)win_rows * FAST_WP + (size_t)sc_rows * FAST_SP + (size_t)FAST_MAXC * prow * 8 + ((size_t)max_h * row_cap + 32 * FAST_WARPS) * 2 +
                      FAST_TAB_BYTES + 32 + FAST_MAXC + 8;   // KITTI: 38 KB
  // The opt-in shared memory size is a per-context function attribute: raise it once per device to the sm_100 maximum
  // (227 KB minus the kernel's static shared memory), so that concurrent callers with different level sizes never lower each other's limit.
  typedef void (*Kern)(const uint8_t*, const int4*, int*, uint32_t*, int*, int, int, int, const OrbGeom, const OrbLevelMaps);
  static const Kern kerns[2][2] = {{orb_fast_band_kernel<false, false>, orb_fast_band_kernel<false, true>},
                                   {orb_fast_band_kernel<true, false>, orb_fast_band_kernel<true, true>}};
  static std::atomic<bool> configured[64][2][2];
  static const bool no_tma = getenv("ORB_B200_NO_TMA") != nullptr;   // A/B switch: per-thread cp.async staging instead of the TMA box load
  const int small = g.ini_th <= 126;      // both thresholds (min <= ini) in the carry-free byte-compare range
  const int tma = (maps != nullptr && !no_tma) ? 1 : 0;
  const Kern kern = kerns[small][tma];
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!configured[dev][small][tma].load(std::memory_order_acquire)) {
    int optin = 0;
    cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);   // the limit covers static + dynamic
    configured[dev][small][tma].store(true, std::memory_order_release);
  }
  static const OrbLevelMaps no_maps = {};
  const dim3 grid(g.fast_tiles_total, batch);
  kern<<<grid, FAST_THREADS, smem, st>>>(pyr, reinterpret_cast<const int4*>(tiles), cell_cnt, cell_kp, err_flag, win_rows, sc_rows, row_cap, g,
                                         tma ? *reinterpret_cast<const OrbLevelMaps*>(maps) : no_maps);
  ++*launches;
}
