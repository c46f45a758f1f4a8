// Shared device/host definitions for the B200 (sm_100a) ORB front-end kernels.
// Data layout in HBM (per extractor handle, batch capacity B frames):
//   pyramid  : [B] frames, each a block of nlevels padded u8 planes (rows_l x pitch_l, pitch 64-byte aligned)
//   blurred  : same layout
//   cell_cnt : [B][ncells_total] int32, cell_kp : [B][ncells_total][CELL_CAP] packed u32 (x:12 | y:12 | score:8)
//   quadtree scratch + kept keypoints per (frame, level)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define ORB_MAX_LEVELS 16
#define ORB_EDGE 19          // EDGE_THRESHOLD, src/ORBextractor.cc:82
#define ORB_HALF_PATCH 15    // HALF_PATCH_SIZE, :81
#define ORB_MIN_BORDER 16    // EDGE_THRESHOLD - 3, :1443
#define ORB_APRON 4          // reflect-101 apron columns written around every padded plane (3 used by the blur)
#define ORB_APRON_ROWS 4     // allocated apron rows above/below a plane (3 written; the 4th is only ever multiplied by 0)
#define ORB_FAST_TILE_W 219  // max detection-domain width of one FAST tile: 16 + 15 + 219 + 6 = the 256-byte window pitch in fast.cu
#define ORB_RS_TW 128        // tiled resize: output tile, TMA box of the source level
#define ORB_RS_TH 32
#define ORB_RS_BW 192
#define ORB_RS_BH 52         // this fork resizes the PADDED level l-1 to the interior of level l: row ratio 1.32 .. 1.56 -> <= 51 source rows per 32
// NMS survivors per FAST cell: no two 8-neighbours both survive the strict 3x3 maximum test, so a cell of w x h detection pixels holds at most
// ceil(w / 2) * ceil(h / 2) of them. OrbGeom::cell_cap is that bound for the largest cell of the shape (KITTI: 16 x 19 = 304): a cell can never
// overflow, like the reference's growing std::vector (src/ORBextractor.cc:1451).

struct OrbLevel {
  int cols, rows, pitch;     // padded level (mvImagePyramid[l] of this fork)
  int iw, ih;                // interior size cvRound(W * invScale)
  unsigned long long plane_off;  // byte offset of the plane inside one frame's pyramid block
  int maxBX, maxBY;          // cols-16, rows-16
  int nCols, nRows, wCell, hCell;
  int cell_base;             // first cell of this level in the per-frame cell arrays
  int nIni;                  // quadtree roots
  float hX;
  int quota;                 // mnFeaturesPerLevel[l]
  int raw_cap, raw_off;      // raw keypoint capacity / offset in per-frame raw arrays
  int node_cap, node_off;    // node arrays
  int kept_off;              // offset into per-frame kept arrays (capacity node_cap)
  float scale;               // mvScaleFactor[l]
  float inv_scale;           // mvInvScaleFactor[l]
  float patch_size;          // (int)(31*scale)
  int tab_off;               // level >= 1: resize tables in padded destination coordinates: x table [xtab_len], then y table [rows]
  int xtab_len;              // cols rounded up to a multiple of 4
  int pyr_fast;              // 1: 4-pixel window kernel applies (horizontal ratio <= 2)
  int pyr_tiled;             // 1: the shared-memory tiled resize applies (every 128 x 32 tile's source window fits one 192 x 52-byte TMA box)
  int rs_ntx, rs_nty;        // tiles per row / column of the (cols + 8) x (rows + 6) output incl. apron
  int rs_off;                // first int of this level's tile table: xlo[rs_ntx], ylo[rs_nty] (box origin in the source level, xlo % 16 == 0), ynr[rs_nty] (source rows needed)
  int fast_G, fast_nseg;     // FAST tile = one cell row x fast_G cells; segments per cell row
  int fast_tile_base;        // first FAST tile of this level
  unsigned fast_inv_nseg;    // 2^20 / fast_nseg + 1: t / fast_nseg == (t * inv) >> 20
  unsigned inv_wcell;        // 2^32 / wCell + 1: p / wCell == __umulhi(p, inv) for p < 2^16 (wCell >= 30)
  int blur_tile_base, blur_tiles_x;   // 128 x 32 blur tiles: first tile of this level, tiles per row
  unsigned blur_inv_tx;      // t / blur_tiles_x == (t * blur_inv_tx) >> 20
};

struct OrbGeom {
  int nlevels, width, height;
  int pad;                   // border stored around every level: ORB_EDGE (this fork: mvImagePyramid[l] is the padded image, F1) or 0 (ORB_FLAG_UPSTREAM_PYRAMID:
                             // raulmur's pyramid, mvImagePyramid[l] is the w_l x h_l image; its reflect-101 border lives in the apron only)
  int ncells_total;
  int cell_cap;              // entries per cell in cell_kp (see above), multiple of 4
  int raw_total, node_total;
  int fast_tiles_total;      // FAST tiles per frame (all levels)
  unsigned long long frame_stride;   // bytes per frame in pyramid / blurred
  int ini_th, min_th;
  int blur_taps[7];
  unsigned blur_k[10];       // packed tap words of the blur kernel: tapA, tapB (DP4A), e0..e3, o0..o3 (DP2A)
  int blur_tiles_total;
  int blur_th;               // rows per blur tile: 64 (default), 32 or 128 (ORB_B200_BLUR_TH)
  // first tile of every level (INT_MAX beyond nlevels): a 4-step binary search finds a CTA's level (orb_level_of)
  int blur_base[ORB_MAX_LEVELS];
  int fast_base[ORB_MAX_LEVELS];
  OrbLevel lv[ORB_MAX_LEVELS];
};

#ifdef __CUDACC__
// level l with base[l] <= tile < base[l + 1]; base is ascending, base[0] == 0, padded with INT_MAX to 16 entries
__device__ __forceinline__ int orb_level_of(const int* base, int tile) {
  int l = tile >= base[8] ? 8 : 0;
  if (tile >= base[l + 4]) l += 4;
  if (tile >= base[l + 2]) l += 2;
  if (tile >= base[l + 1]) l += 1;
  return l;
}
#endif

struct ResizeTab {   // per PADDED destination index of cv::resize INTER_LINEAR: source offset + 11-bit coefficients
  int ofs;           // first source pixel (second = ofs + 1; its coefficient is 0 where that would leave the image)
  unsigned coef;     // c0 | c1 << 16   (c0 + c1 == 2048)
};

struct OrbKeypoint {  // == cv::KeyPoint / orb_keypoint
  float x, y, size, angle, response;
  int octave, class_id;
};

struct QtScratch {          // per-frame strides are the *_total fields of OrbGeom
  uint32_t* key;            // [B][raw_total]   packed x | y<<12 | score<<24
  int* knode;               // [B][raw_total]
  uint8_t* kchild;          // [B][raw_total]
  int4* bbox;               // [B][node_total]  ulx, uly, brx, bry
  int* ncount;              // [B][node_total]
  int* pidx;                // [B][node_total]  index in P this round, or -1
  unsigned* best;           // [B][node_total]
  int* ordA;                // [B][node_total]
  int* ordB;                // [B][node_total]
  int* par;                 // [B][node_total]
  unsigned long long* cand; // [B][node_total*2] (size<<40 | seq<<20 | node), second half = sort padding
  int* ccnt;                // [B][node_total*4] child key counts per parent index
  int* cslot;               // [B][node_total*4] child slots per parent index
  unsigned long long* scan; // [B][node_total]   packed scan: low = nonempty children, high = multi-key children
  int* raw_n;               // [B][nlevels]      raw keypoint count (test hook / diagnostics)
};

__host__ __device__ __forceinline__ int orb_reflect101(int p, int n) {
  // BORDER_REFLECT_101 for |overshoot| < n (border 19 / 3 << level sizes, checked on the host)
  p = p < 0 ? -p : p;
  return p >= n ? 2 * (n - 1) - p : p;
}
