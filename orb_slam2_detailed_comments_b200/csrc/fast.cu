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
//   2. quick reject, 4 pixels per thread with byte-SIMD (VABSDIFF4): a 9-arc always holds one of {up, down}
//      and one of {left, right}, so min(max|d_ud|, max|d_lr|) > minTh is necessary for S >= minTh;
//   3. surviving candidates are compacted per warp (ballot) and scored 32 at a time with full lanes;
//   4. per cell (one warp each): NMS on the shared score plane, ballot/popc emission in row-major order.
#include "orb_common.cuh"

#define FAST_THREADS 256
#define FAST_WARPS 8
#define FAST_WP 288      // window row pitch in smem: 16 B front pad + <= 272 B
#define FAST_SP 272      // score plane row pitch
#define FAST_LIST 160    // per-warp candidate list capacity (< 32 + 128)

__device__ __forceinline__ int fast_score(const uint8_t* __restrict__ p, int wp, int min_th) {
  // ring offsets (dx,dy), k = 0..15: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
  const int v = p[0];
  int d[16];
  d[0] = v - p[3 * wp];      d[8] = v - p[-3 * wp];
  d[4] = v - p[3];           d[12] = v - p[-3];
  d[2] = v - p[2 * wp + 2];  d[10] = v - p[-2 * wp - 2];
  d[6] = v - p[-2 * wp + 2]; d[14] = v - p[2 * wp - 2];
  // every 9-arc holds one pixel of each opposite pair: sign-consistent reject on the four even pairs
  {
    const bool dark = (max(d[0], d[8]) > min_th) && (max(d[4], d[12]) > min_th) && (max(d[2], d[10]) > min_th) &&
                      (max(d[6], d[14]) > min_th);
    const bool bright = (min(d[0], d[8]) < -min_th) && (min(d[4], d[12]) < -min_th) && (min(d[2], d[10]) < -min_th) &&
                        (min(d[6], d[14]) < -min_th);
    if (!dark && !bright) return 0;
  }
  d[1] = v - p[3 * wp + 1];  d[3] = v - p[wp + 3];
  d[5] = v - p[-wp + 3];     d[7] = v - p[-3 * wp + 1];
  d[9] = v - p[-3 * wp - 1]; d[11] = v - p[-wp - 3];
  d[13] = v - p[wp - 3];     d[15] = v - p[3 * wp - 1];
  int lo2[16], hi2[16], lo4[16], hi4[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) { lo2[k] = min(d[k], d[(k + 1) & 15]); hi2[k] = max(d[k], d[(k + 1) & 15]); }
#pragma unroll
  for (int k = 0; k < 16; ++k) { lo4[k] = min(lo2[k], lo2[(k + 2) & 15]); hi4[k] = max(hi2[k], hi2[(k + 2) & 15]); }
  int A = -256, Bn = 256;  // A = max_k min(arc_k), Bn = min_k max(arc_k)  (B = -Bn)
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int lo9 = min(min(lo4[k], lo4[(k + 4) & 15]), d[(k + 8) & 15]);
    const int hi9 = max(max(hi4[k], hi4[(k + 4) & 15]), d[(k + 8) & 15]);
    A = max(A, lo9);
    Bn = min(Bn, hi9);
  }
  const int s = max(A, -Bn) - 1;
  return s < 0 ? 0 : s;
}

__global__ void __launch_bounds__(FAST_THREADS)
orb_fast_band_kernel(const uint8_t* __restrict__ pyr, int* __restrict__ cell_cnt, uint32_t* __restrict__ cell_kp,
                     int* __restrict__ err_flag, int win_rows, int sc_rows, const __grid_constant__ OrbGeom g) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* win = smem;                                    // [win_rows][FAST_WP], pixel x at column 16 + x
  uint8_t* sc = win + win_rows * FAST_WP;                 // [sc_rows][FAST_SP], score of domain row y at row y + 1
  uint16_t* lists = reinterpret_cast<uint16_t*>(sc + sc_rows * FAST_SP);        // [FAST_WARPS][FAST_LIST]
  uint32_t* ballots = reinterpret_cast<uint32_t*>(lists + FAST_WARPS * FAST_LIST);  // [FAST_WARPS][4 * (sc_rows - 2)]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int frame = blockIdx.y;
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < g.nlevels; ++l)
    if ((int)blockIdx.x >= g.lv[l].fast_tile_base) level = l;
  const OrbLevel& L = g.lv[level];
  const int t = blockIdx.x - L.fast_tile_base;
  const int ci = t / L.fast_nseg, sg = t - ci * L.fast_nseg;
  const int j0 = sg * L.fast_G, j1 = min(j0 + L.fast_G, L.nCols);
  int* cnt_row = cell_cnt + (long long)frame * g.ncells_total + L.cell_base + ci * L.nCols;
  uint32_t* kp_row = cell_kp + ((long long)frame * g.ncells_total + L.cell_base + ci * L.nCols) * ORB_CELL_CAP;

  const int iniY = ORB_MIN_BORDER + ci * L.hCell;
  const int wh = min(iniY + L.hCell + 6, L.maxBY) - iniY, dh = wh - 6;   // window rows, detection rows
  const int X0 = ORB_MIN_BORDER + j0 * L.wCell;                          // window [X0, X1)
  const int X1 = min(ORB_MIN_BORDER + j1 * L.wCell + 6, L.maxBX);
  if (iniY >= L.maxBY - 3 || dh <= 0 || X1 - X0 <= 6) {                  // :1479-1503 skip rules / empty domain
    for (int j = j0 + tid; j < j1; j += FAST_THREADS) cnt_row[j] = 0;
    return;
  }
  const int xa = X0 & ~15;                         // 16-byte aligned global column of window column 0
  const int nvec = (X1 - xa + 15) >> 4;
  const uint8_t* src = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off + (long long)iniY * L.pitch + xa;
  // 1. stage the window
  for (int i = tid; i < wh * nvec; i += FAST_THREADS) {
    const int y = i / nvec, v = i - y * nvec;
    *reinterpret_cast<uint4*>(win + y * FAST_WP + 16 + 16 * v) = *reinterpret_cast<const uint4*>(src + (long long)y * L.pitch + 16 * v);
  }
  for (int i = tid; i < (dh + 2) * (FAST_SP / 16); i += FAST_THREADS) reinterpret_cast<uint4*>(sc)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();

  // 2 + 3. quick reject by pixel quads, per-warp compaction, full score of candidates
  const int dx0 = X0 + 3 - xa, dx1 = X1 - 3 - xa;       // detection columns in window coordinates
  const int qbase = dx0 & ~3;
  const int nq = (dx1 - qbase + 3) >> 2;                 // <= 64
  const unsigned th4 = (unsigned)g.min_th * 0x01010101u;
  uint16_t* list = lists + warp * FAST_LIST;
  const uint32_t lt = (1u << lane) - 1u;
  int cnt = 0;
  for (int y = warp; y < dh; y += FAST_WARPS) {
    const uint8_t* row = win + (y + 3) * FAST_WP + 16;
    for (int h = 0; h < nq; h += 32) {
      const int q = h + lane;
      const int x = qbase + 4 * q;
      unsigned pass = 0;
      if (q < nq) {
        const unsigned c0 = *reinterpret_cast<const unsigned*>(row + x - 4), c1 = *reinterpret_cast<const unsigned*>(row + x),
                       c2 = *reinterpret_cast<const unsigned*>(row + x + 4);
        const unsigned up = *reinterpret_cast<const unsigned*>(row + x - 3 * FAST_WP), dn = *reinterpret_cast<const unsigned*>(row + x + 3 * FAST_WP);
        const unsigned lf = __byte_perm(c0, c1, 0x4321), rt = __byte_perm(c1, c2, 0x6543);
        const unsigned a = __vminu4(__vmaxu4(__vabsdiffu4(c1, up), __vabsdiffu4(c1, dn)),
                                    __vmaxu4(__vabsdiffu4(c1, lf), __vabsdiffu4(c1, rt)));
        pass = __vcmpgtu4(a, th4);
        if (pass) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (x + k < dx0 || x + k >= dx1) pass &= ~(0xffu << (8 * k));
        }
      }
      if (__any_sync(0xffffffffu, pass != 0)) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const bool f = (pass >> (8 * k)) & 1u;
          const uint32_t b = __ballot_sync(0xffffffffu, f);
          if (f) list[cnt + __popc(b & lt)] = (uint16_t)((y << 9) | (x + k));
          cnt += __popc(b);
        }
        __syncwarp();
        while (cnt >= 32) {
          cnt -= 32;
          const int e = list[cnt + lane];
          const int ey = e >> 9, ex = e & 511;
          const int s = fast_score(win + (ey + 3) * FAST_WP + 16 + ex, FAST_WP, g.min_th);
          if (s) sc[(ey + 1) * FAST_SP + ex] = (uint8_t)s;
          __syncwarp();
        }
      }
    }
  }
  if (lane < cnt) {
    const int e = list[lane];
    const int ey = e >> 9, ex = e & 511;
    const int s = fast_score(win + (ey + 3) * FAST_WP + 16 + ex, FAST_WP, g.min_th);
    if (s) sc[(ey + 1) * FAST_SP + ex] = (uint8_t)s;
  }
  __syncthreads();

  // 4. per cell: NMS inside the cell's detection window, ini/min selection, ordered emission
  uint32_t* bal = ballots + warp * 4 * (sc_rows - 2);   // [row][chunk][min, ini]
  for (int cj = j0 + warp; cj < j1; cj += FAST_WARPS) {
    const int cx0 = ORB_MIN_BORDER + cj * L.wCell + 3 - xa;
    const int cx1 = min(ORB_MIN_BORDER + (cj + 1) * L.wCell + 3, L.maxBX - 3) - xa;
    const int dw = cx1 - cx0;
    if (dw <= 0) {
      if (lane == 0) cnt_row[cj] = 0;
      continue;
    }
    const int nch = (dw + 31) >> 5;   // 1 or 2
    int n_ini = 0, n_min = 0;
    for (int y = 0; y < dh; ++y) {
      for (int c = 0; c < nch; ++c) {
        const int x = cx0 + (c << 5) + lane;
        bool f_min = false, f_ini = false;
        if (x < cx1) {
          const uint8_t* q = sc + (y + 1) * FAST_SP + x;
          const int s = q[0];
          if (s >= g.min_th) {
            const bool hl = x > cx0, hr = x + 1 < cx1;
            bool ok = s > q[-FAST_SP] && s > q[FAST_SP];
            if (hl) ok = ok && s > q[-1] && s > q[-FAST_SP - 1] && s > q[FAST_SP - 1];
            if (hr) ok = ok && s > q[1] && s > q[-FAST_SP + 1] && s > q[FAST_SP + 1];
            f_min = ok;
            f_ini = ok && s >= g.ini_th;
          }
        }
        const uint32_t bm = __ballot_sync(0xffffffffu, f_min), bi = __ballot_sync(0xffffffffu, f_ini);
        if (lane == 0) { bal[(y * 2 + c) * 2] = bm; bal[(y * 2 + c) * 2 + 1] = bi; }
        n_min += __popc(bm);
        n_ini += __popc(bi);
      }
    }
    __syncwarp();
    const int sel = n_ini > 0 ? 1 : 0;
    const int total = n_ini > 0 ? n_ini : n_min;
    uint32_t* kp_out = kp_row + (long long)cj * ORB_CELL_CAP;
    int base = 0;
    if (total > 0) {
      for (int y = 0; y < dh; ++y) {
        for (int c = 0; c < nch; ++c) {
          const uint32_t m = bal[(y * 2 + c) * 2 + sel];
          if ((m >> lane) & 1u) {
            const int r = base + __popc(m & lt);
            if (r < ORB_CELL_CAP) {
              const int x = cx0 + (c << 5) + lane;
              const uint32_t s = sc[(y + 1) * FAST_SP + x];
              // cell-local FAST coordinates shifted by (j*wCell, i*hCell) (:1540-1541) == level pixel - 16
              kp_out[r] = (uint32_t)(x + xa - ORB_MIN_BORDER) | ((uint32_t)(y + 3 + ci * L.hCell) << 12) | (s << 24);
            }
          }
          base += __popc(m);
        }
      }
    }
    if (lane == 0) {
      cnt_row[cj] = min(total, ORB_CELL_CAP);
      if (total > ORB_CELL_CAP) atomicOr(err_flag, 1);
    }
    __syncwarp();
  }
}

void orb_launch_fast(const uint8_t* pyr, int* cell_cnt, uint32_t* cell_kp, int* err_flag, const OrbGeom& g, int batch,
                     cudaStream_t st, long long* launches) {
  int max_h = 0;
  for (int l = 0; l < g.nlevels; ++l) max_h = max(max_h, g.lv[l].hCell);
  const int win_rows = max_h + 6, sc_rows = max_h + 2;
  const size_t smem = (size_t)win_rows * FAST_WP + (size_t)sc_rows * FAST_SP + FAST_WARPS * FAST_LIST * 2 +
                      (size_t)FAST_WARPS * 4 * max_h * 4;
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaFuncSetAttribute(orb_fast_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  orb_fast_band_kernel<<<dim3(g.fast_tiles_total, batch), FAST_THREADS, smem, st>>>(pyr, cell_cnt, cell_kp, err_flag, win_rows,
                                                                                   sc_rows, g);
  ++*launches;
}
