// Per-cell FAST-9/16 with the iniThFAST / minThFAST retry: the cell loop of
// ORBextractor::ComputeKeyPointsOctTree (src/ORBextractor.cc:1465-1548) and cv::FAST(cell, th, true).
//
// One warp = one 30-px cell. cornerScore is threshold-free (S = max(A,B) - 1), "corner at T" <=> S >= T,
// so ONE score pass serves both thresholds: NMS survivors are computed once on S (neighbours outside the
// cell's detection window count as 0, because the reference runs cv::FAST on the cell sub-image), the
// cell emits the survivors with S >= ini, or, if there are none, those with S >= min. Emission is
// row-major inside the cell (ballot + popc prefix), which is cv::FAST's order.
#include "orb_common.cuh"

#define FAST_WARPS 8

__device__ __forceinline__ int fast_score(const uint8_t* __restrict__ p, int wp, int min_th) {
  // ring offsets (dx,dy), k = 0..15: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
  const int v = p[0];
  int d[16];
  d[0] = v - p[3 * wp];      d[8] = v - p[-3 * wp];
  d[4] = v - p[3];           d[12] = v - p[-3];
  // opposite-pair quick reject on the compass pairs: a 9-arc holds one pixel of every opposite pair
  {
    const bool dark = (max(d[0], d[8]) > min_th) && (max(d[4], d[12]) > min_th);
    const bool bright = (min(d[0], d[8]) < -min_th) && (min(d[4], d[12]) < -min_th);
    if (!dark && !bright) return 0;
  }
  d[1] = v - p[3 * wp + 1];  d[2] = v - p[2 * wp + 2];   d[3] = v - p[wp + 3];
  d[5] = v - p[-wp + 3];     d[6] = v - p[-2 * wp + 2];  d[7] = v - p[-3 * wp + 1];
  d[9] = v - p[-3 * wp - 1]; d[10] = v - p[-2 * wp - 2]; d[11] = v - p[-wp - 3];
  d[13] = v - p[wp - 3];     d[14] = v - p[2 * wp - 2];  d[15] = v - p[3 * wp - 1];
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

// smem per warp: window (win_h x win_p bytes) + score plane ((dom_h + 2) x sc_p bytes, 1-px zero ring)
// + two ballot arrays. Sizes come from the host (max over levels).
__global__ void __launch_bounds__(FAST_WARPS * 32)
orb_fast_cells_kernel(const uint8_t* __restrict__ pyr, int* __restrict__ cell_cnt, uint32_t* __restrict__ cell_kp,
                      int* __restrict__ err_flag, int win_p, int win_rows, int sc_p, int sc_rows, int max_chunks,
                      long long total_cells, const __grid_constant__ OrbGeom g) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_warp = win_p * win_rows + sc_p * sc_rows + max_chunks * 8;
  uint8_t* win = smem + warp * per_warp;
  uint8_t* sc = win + win_p * win_rows;
  uint32_t* bal_ini = reinterpret_cast<uint32_t*>(sc + sc_p * sc_rows);
  uint32_t* bal_min = bal_ini + max_chunks;

  const long long gcell = (long long)blockIdx.x * FAST_WARPS + warp;
  if (gcell >= total_cells) return;
  const int frame = (int)(gcell / g.ncells_total);
  const int cell = (int)(gcell - (long long)frame * g.ncells_total);
  int level = 0;
#pragma unroll 1
  for (int l = 1; l < g.nlevels; ++l)
    if (cell >= g.lv[l].cell_base) level = l;
  const OrbLevel& L = g.lv[level];
  const int ci = (cell - L.cell_base) / L.nCols, cj = (cell - L.cell_base) - ci * L.nCols;
  int* cnt_out = cell_cnt + (long long)frame * g.ncells_total + cell;
  uint32_t* kp_out = cell_kp + ((long long)frame * g.ncells_total + cell) * ORB_CELL_CAP;

  const int iniX = ORB_MIN_BORDER + cj * L.wCell, iniY = ORB_MIN_BORDER + ci * L.hCell;
  if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6) {   // :1479-1503 skip rules
    if (lane == 0) *cnt_out = 0;
    return;
  }
  const int ww = min(iniX + L.wCell + 6, L.maxBX) - iniX, wh = min(iniY + L.hCell + 6, L.maxBY) - iniY;
  const int dw = ww - 6, dh = wh - 6;   // detection domain [3, dim-3)
  if (dw <= 0 || dh <= 0) {
    if (lane == 0) *cnt_out = 0;
    return;
  }
  const uint8_t* src = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off + (long long)iniY * L.pitch + iniX;
  for (int i = lane; i < ww * wh; i += 32) {
    const int y = i / ww, x = i - y * ww;
    win[y * win_p + x] = src[(long long)y * L.pitch + x];
  }
  for (int i = lane; i < sc_p * (dh + 2); i += 32) sc[i] = 0;
  __syncwarp();
  const int npx = dw * dh;
  for (int i = lane; i < npx; i += 32) {
    const int y = i / dw, x = i - y * dw;
    const int s = fast_score(win + (y + 3) * win_p + (x + 3), win_p, g.min_th);
    if (s) sc[(y + 1) * sc_p + (x + 1)] = (uint8_t)s;
  }
  __syncwarp();
  const int nchunks = (npx + 31) >> 5;
  int n_ini = 0, n_min = 0;
  for (int c = 0; c < nchunks; ++c) {
    const int i = (c << 5) + lane;
    bool f_min = false, f_ini = false;
    if (i < npx) {
      const int y = i / dw, x = i - y * dw;
      const uint8_t* q = sc + (y + 1) * sc_p + (x + 1);
      const int s = q[0];
      if (s >= g.min_th && s > q[-1] && s > q[1] && s > q[-sc_p - 1] && s > q[-sc_p] && s > q[-sc_p + 1] &&
          s > q[sc_p - 1] && s > q[sc_p] && s > q[sc_p + 1]) {
        f_min = true;
        f_ini = s >= g.ini_th;
      }
    }
    const uint32_t bm = __ballot_sync(0xffffffffu, f_min), bi = __ballot_sync(0xffffffffu, f_ini);
    if (lane == 0) { bal_min[c] = bm; bal_ini[c] = bi; }
    n_min += __popc(bm);
    n_ini += __popc(bi);
  }
  __syncwarp();
  const uint32_t* bal = n_ini > 0 ? bal_ini : bal_min;
  const int total = n_ini > 0 ? n_ini : n_min;
  int base = 0;
  const uint32_t lt = (1u << lane) - 1u;
  for (int c = 0; c < nchunks; ++c) {
    const uint32_t m = bal[c];
    if ((m >> lane) & 1u) {
      const int r = base + __popc(m & lt);
      if (r < ORB_CELL_CAP) {
        const int i = (c << 5) + lane;
        const int y = i / dw, x = i - y * dw;
        const uint32_t s = sc[(y + 1) * sc_p + (x + 1)];
        // cell-local (x+3, y+3) shifted by (j*wCell, i*hCell)  (:1540-1541)
        kp_out[r] = (uint32_t)(x + 3 + cj * L.wCell) | ((uint32_t)(y + 3 + ci * L.hCell) << 12) | (s << 24);
      }
    }
    base += __popc(m);
  }
  if (lane == 0) {
    *cnt_out = min(total, ORB_CELL_CAP);
    if (total > ORB_CELL_CAP) atomicOr(err_flag, 1);
  }
}

void orb_launch_fast(const uint8_t* pyr, int* cell_cnt, uint32_t* cell_kp, int* err_flag, const OrbGeom& g, int batch,
                     cudaStream_t st, long long* launches) {
  int max_w = 0, max_h = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    max_w = max(max_w, g.lv[l].wCell + 6);
    max_h = max(max_h, g.lv[l].hCell + 6);
  }
  const int win_p = (max_w + 3) & ~3, win_rows = max_h;
  const int sc_p = (max_w - 6 + 2 + 3) & ~3, sc_rows = max_h - 6 + 2;
  const int max_chunks = ((max_w - 6) * (max_h - 6) + 31) / 32;
  const int per_warp = win_p * win_rows + sc_p * sc_rows + max_chunks * 8;
  const size_t smem = (size_t)per_warp * FAST_WARPS;
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaFuncSetAttribute(orb_fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  const long long total = (long long)g.ncells_total * batch;
  const unsigned grid = (unsigned)((total + FAST_WARPS - 1) / FAST_WARPS);
  orb_fast_cells_kernel<<<grid, FAST_WARPS * 32, smem, st>>>(pyr, cell_cnt, cell_kp, err_flag, win_p, win_rows, sc_p,
                                                            sc_rows, max_chunks, total, g);
  ++*launches;
}
