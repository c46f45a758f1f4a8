// C ABI (include/orb_b200.h) over the sm_100a kernels. Host logic only: parameter tables exactly as the
// reference constructor computes them (src/ORBextractor.cc:776-925), per-shape geometry, workspaces, launches.
// No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
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
void orb_launch_pyramid(const uint8_t*, int, long long, uint8_t*, const ResizeTab*, const OrbGeom&, int, cudaStream_t, long long*, cudaEvent_t);
void orb_launch_fast(const uint8_t*, int*, uint32_t*, int*, const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_quadtree(const int*, const uint32_t*, const QtScratch&, uint32_t*, int*, int*, const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_blur(const uint8_t*, uint8_t*, const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_describe(const uint8_t*, const uint8_t*, const uint32_t*, const int*, OrbKeypoint*, uint8_t*, int, int*, int*,
                         const OrbGeom&, int, cudaStream_t, long long*);
void orb_launch_stereo(const uint8_t*, const OrbKeypoint*, const uint8_t*, const int*, int, float, float, float*, float*, int*, int*, int*,
                       unsigned short*, const OrbGeom&, int, cudaStream_t, long long*);
int orb_knn_num_chunks(int Q, long long D, int num_sms);
void orb_launch_knn(const uint8_t*, int, const uint8_t*, long long, long long, int, unsigned long long*, int, OrbKnn2*, cudaStream_t);
void orb_launch_knn_merge_shards(const OrbKnn2*, int, int, int, OrbKnn2*, cudaStream_t);
void orb_launch_distance_batch(const uint8_t*, const uint8_t*, long long, int*, cudaStream_t);
long long orb_knn_launches();
void orb_knn_count_launches(long long n);
void orb_launch_search_by_bow(const uint8_t*, const OrbKeypoint*, const uint8_t*, const int*, const int*, const int*, const int*, int, const uint8_t*,
                              const OrbKeypoint*, const int*, const int*, const int*, const int*, int, int, int, float, int, void*, int*, uint2*, int*,
                              int*, int*, int*, cudaStream_t, long long*);
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
double orb_popc_peak(int num_sms);

static_assert(sizeof(orb_keypoint) == 28 && sizeof(OrbKeypoint) == 28, "cv::KeyPoint layout");
static_assert(sizeof(orb_knn2) == 24, "orb_knn2 layout");

// ---- errors ------------------------------------------------------------------------------------------
static thread_local char t_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) return fail(ORB_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

static inline int cv_round_f(float v) { return (int)lrintf(v); }

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

#define ORB_MAX_WS 4
struct orb_extractor {
  orb_params prm;
  int device = 0, num_sms = 148;
  cudaStream_t stream = nullptr;
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> quota;
  OrbGeom geom;            // valid when geom.width > 0
  int chunk = 0;           // frames per internal pass (workspace capacity)
  int cap_ws = 0;
  long long launches = 0;
  DevBuf tabs, err;
  // Two pass workspaces on two compute streams: consecutive internal passes alternate between them, so the latency-bound
  // tail of pass p (quadtree, descriptors) overlaps the bandwidth/issue-bound head of pass p+1 (pyramid, FAST, blur).
  // ws[0].st is the handle's public stream.
  struct PassWs {
    cudaStream_t st = nullptr;
    cudaEvent_t ev_join = nullptr;
    int frames = 0;          // frames of the last pass run here
    DevBuf pyr, blr, cell_cnt, cell_kp, kept, kept_cnt, out_cnt;
    DevBuf st_sad;   // stereo: SAD of every kept match [chunk/2][cap]
    DevBuf st_rows, st_sorted;   // stereo: right keypoints sorted by image row: starts [chunk/2][rows0 + 1], indices [chunk/2][cap]
    DevBuf q_key, q_knode, q_kchild, q_bbox, q_ncount, q_pidx, q_best, q_ordA, q_ordB, q_par, q_cand, q_ccnt, q_cslot, q_scan, q_rawn;
    DevBuf* all[25] = {&st_rows, &st_sorted, &st_sad, &pyr, &blr, &cell_cnt, &cell_kp, &kept, &kept_cnt, &out_cnt, &q_key, &q_knode, &q_kchild, &q_bbox, &q_ncount,
                       &q_pidx, &q_best, &q_ordA, &q_ordB, &q_par, &q_cand, &q_ccnt, &q_cslot, &q_scan, &q_rawn};
  } ws[ORB_MAX_WS];
  int nws = 2;             // compute streams / workspaces in use (ORB_B200_STREAMS, 1..ORB_MAX_WS)
  int last_ws = 0;         // workspace of the most recent pass (mvImagePyramid / test hooks read it)
  cudaEvent_t ev_fork = nullptr;
  // host-facing pipeline (orb_extract_batch): H2D of pass p+1 and D2H of pass p-1 overlap the kernels of pass p
  cudaStream_t s_in = nullptr, s_out = nullptr;
  // Output staging has one buffer more than there are workspaces: pass p + nws runs on the workspace of pass p and must not
  // wait for pass p's D2H copy.
  DevBuf stage_in[2], out_kps[ORB_MAX_WS + 1], out_desc[ORB_MAX_WS + 1], out_ur[ORB_MAX_WS + 1], out_dp[ORB_MAX_WS + 1];
  DevBuf cnt_all;   // per call: counts [batch], level counts [batch][nlevels], stereo matches [batch/2]; one D2H at the end
  cudaEvent_t ev_in[2] = {}, ev_done[ORB_MAX_WS] = {}, ev_out[ORB_MAX_WS + 1] = {}, ev_staged[2] = {};   // ev_staged: level 0 has consumed stage_in[b]
  size_t h_cnt_cap = 0;    // ints in h_cnt
  int* h_err = nullptr;    // pinned
  int* h_cnt = nullptr;    // pinned [batch * (1 + nlevels)]: counts, then per-level counts
  bool profile = false;    // per-stage CUDA-event timing (orb_extractor_set_profiling)
  std::vector<cudaEvent_t> ev;   // (ORB_NUM_STAGES + 1) events per profiled pass, resolved lazily (no sync in the pass)
  size_t ev_used = 0;
  double stage_ms[ORB_NUM_STAGES] = {};
  long long stage_passes = 0;
  std::mutex mu;
};

// src/ORBextractor.cc:776-845 with the reference's types: scaleFactor member is double, tables are float.
static void build_tables(orb_extractor* e) {
  const int nl = e->prm.nlevels;
  const double sf = (double)e->prm.scale_factor;
  e->scale.assign(nl, 1.f); e->sigma2.assign(nl, 1.f); e->invScale.assign(nl, 1.f); e->invSigma2.assign(nl, 1.f);
  for (int i = 1; i < nl; ++i) {
    e->scale[i] = (float)(e->scale[i - 1] * sf);
    e->sigma2[i] = e->scale[i] * e->scale[i];
  }
  for (int i = 0; i < nl; ++i) { e->invScale[i] = 1.0f / e->scale[i]; e->invSigma2[i] = 1.0f / e->sigma2[i]; }
  e->quota.assign(nl, 0);
  const float factor = (float)(1.0f / sf);
  float per = e->prm.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
  int sum = 0;
  for (int l = 0; l < nl - 1; ++l) {
    e->quota[l] = cv_round_f(per);
    sum += e->quota[l];
    per *= factor;
  }
  e->quota[nl - 1] = std::max(e->prm.nfeatures - sum, 0);
}

// cv::resize INTER_LINEAR coefficient table for one axis (ssz -> dsz interior pixels), 11-bit fixed point, laid out in
// APRON coordinates of the padded destination: entry t describes padded index p = t - lead, i.e. interior pixel
// reflect101(reflect101(p, dsz + 38) - 19, dsz). The in-place copyMakeBorder of src/ORBextractor.cc:2138 is an index remap
// of the interior, and the apron (p < 0, p >= dsz + 38) is the BORDER_REFLECT_101 extension of the padded image that the
// 7x7 blur reads (:2042). `len` entries are written; entries past the apron repeat the last valid one.
static void resize_table(int ssz, int dsz, int lead, int len, ResizeTab* out) {
  const double scale = 1.0 / ((double)dsz / ssz);
  const int padded = dsz + 2 * ORB_EDGE;
  for (int t = 0; t < len; ++t) {
    const int p = std::min(t - lead, padded + ORB_APRON - 1);
    const int d = orb_reflect101(orb_reflect101(p, padded) - ORB_EDGE, dsz);
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= ssz - 1) { s = ssz - 1; f = 0.f; }
    out[t].ofs = s;
    out[t].coef = (unsigned)cv_round_f((1.f - f) * 2048.f) | ((unsigned)cv_round_f(f * 2048.f) << 16);
  }
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Per-shape geometry: level sizes (:2099-2102), FAST cell grid (:1443-1462), quadtree roots (:1059-1063).
static int build_geometry(orb_extractor* e, int W, int H, std::vector<ResizeTab>& tabs) {
  OrbGeom g;
  memset(&g, 0, sizeof(g));
  const int nl = e->prm.nlevels;
  g.nlevels = nl; g.width = W; g.height = H;
  g.ini_th = e->prm.ini_th_fast; g.min_th = e->prm.min_th_fast;
  static const int taps_new[7] = {18, 34, 48, 56, 48, 34, 18}, taps_old[7] = {18, 34, 49, 55, 49, 34, 18};
  memcpy(g.blur_taps, (e->prm.flags & ORB_FLAG_LEGACY_BLUR) ? taps_old : taps_new, sizeof(g.blur_taps));
  {
    const unsigned k0 = g.blur_taps[0], k1 = g.blur_taps[1], k2 = g.blur_taps[2], k3 = g.blur_taps[3];
    const unsigned kk[10] = {k0 | (k1 << 8) | (k2 << 16) | (k3 << 24), k2 | (k1 << 8) | (k0 << 16),        // H pass, DP4A
                             k0 << 8, k1 | (k2 << 8), k3 | (k2 << 8), k1 | (k0 << 8),                     // V pass, even rows
                             k0 | (k1 << 8), k2 | (k3 << 8), k2 | (k1 << 8), k0};                         // V pass, odd rows
    memcpy(g.blur_k, kk, sizeof(kk));
  }
  int btiles = 0;
  size_t plane = 0;
  int cells = 0, raw = 0, nodes = 0, tab = 0, ftiles = 0;
  tabs.clear();
  for (int l = 0; l < nl; ++l) {
    OrbLevel& L = g.lv[l];
    L.iw = cv_round_f((float)W * e->invScale[l]);
    L.ih = cv_round_f((float)H * e->invScale[l]);
    if (L.iw < 2 * ORB_EDGE + 2 || L.ih < 2 * ORB_EDGE + 2)
      return fail(ORB_E_SHAPE, "level %d is %dx%d: too small for the 19-px reflect border", l, L.iw, L.ih);
    L.cols = L.iw + 2 * ORB_EDGE; L.rows = L.ih + 2 * ORB_EDGE;
    // plane = [4 apron rows][rows][4 apron rows], row = [16 B lead][cols][>= 8 B]: the pyramid kernels also write the
    // reflect-101 extension of the padded image (4 columns / 3 rows each side) so the blur never remaps indices
    L.pitch = (int)align_up(16 + L.cols + 8, 64);
    L.plane_off = plane + (size_t)ORB_APRON_ROWS * L.pitch + 16;
    plane += align_up((size_t)L.pitch * (L.rows + 2 * ORB_APRON_ROWS), 256);
    L.maxBX = L.cols - ORB_EDGE + 3; L.maxBY = L.rows - ORB_EDGE + 3;
    const float width = (float)(L.maxBX - ORB_MIN_BORDER), height = (float)(L.maxBY - ORB_MIN_BORDER);
    L.nCols = (int)(width / 30.f); L.nRows = (int)(height / 30.f);
    if (L.nCols < 1 || L.nRows < 1) return fail(ORB_E_SHAPE, "level %d has no 30-px cell", l);
    L.wCell = (int)ceilf(width / L.nCols); L.hCell = (int)ceilf(height / L.nRows);
    if (L.maxBX - ORB_MIN_BORDER > 4095 || L.maxBY - ORB_MIN_BORDER > 4095)
      return fail(ORB_E_SHAPE, "level %d exceeds the 4095-px packed coordinate range", l);
    L.cell_base = cells;
    cells += L.nCols * L.nRows;
    L.fast_G = std::max(1, ORB_FAST_TILE_W / L.wCell);
    L.fast_nseg = (L.nCols + L.fast_G - 1) / L.fast_G;
    L.fast_tile_base = ftiles;
    ftiles += L.fast_nseg * L.nRows;
    L.blur_tiles_x = (L.cols + 127) / 128;
    L.blur_tile_base = btiles;
    L.blur_inv_tx = (1u << 20) / (unsigned)L.blur_tiles_x + 1u;
    btiles += L.blur_tiles_x * ((L.rows + 31) / 32);
    L.nIni = (int)roundf((float)(L.maxBX - ORB_MIN_BORDER) / (L.maxBY - ORB_MIN_BORDER));
    if (L.nIni < 1) return fail(ORB_E_SHAPE, "aspect ratio < 0.5: the reference divides by nIni == 0 (src/ORBextractor.cc:1059)");
    L.hX = (float)(L.maxBX - ORB_MIN_BORDER) / L.nIni;
    L.quota = e->quota[l];
    const long long by_cells = (long long)L.nCols * L.nRows * ORB_CELL_CAP;
    // a level can never hold more raw keypoints than its cells can (cell overflow is flagged separately); dense 4K frames
    // exceed any quota-based bound (the reference's vector simply grows, src/ORBextractor.cc:1451)
    L.raw_cap = (int)by_cells;
    L.raw_off = raw; raw += (int)align_up(L.raw_cap, 32);
    L.node_cap = std::max(L.quota + 3, 4 * L.nIni) + L.nIni + 8;
    L.node_off = nodes; nodes += (int)align_up(L.node_cap, 32);
    L.kept_off = L.node_off;
    L.scale = e->scale[l];
    L.inv_scale = e->invScale[l];
    L.patch_size = (float)(int)(31 * e->scale[l]);
    L.tab_off = tab;
    if (l > 0) {
      const OrbLevel& P = g.lv[l - 1];
      L.xtab_len = (L.cols + 2 * ORB_APRON + 3) & ~3;                   // columns -4 .. cols + 4
      const int yrows = L.rows + 2 * (ORB_APRON - 1);                   // rows -3 .. rows + 3
      const int ytab_len = (yrows + 3) & ~3;   // keeps every table 16-byte aligned
      tabs.resize(tab + L.xtab_len + ytab_len);
      resize_table(P.cols, L.iw, ORB_APRON, L.xtab_len, tabs.data() + tab);          // source = PADDED level l-1 (this fork, :2119)
      resize_table(P.rows, L.ih, ORB_APRON - 1, ytab_len, tabs.data() + tab + L.xtab_len);
      // the 4-pixel kernel reads an 8-byte source window per thread: needs ofs(x+3) + 1 - ofs(x) <= 7
      L.pyr_fast = 1;
      for (int x = 0; x + 3 < L.xtab_len && L.pyr_fast; x += 4) {
        int lo = tabs[tab + x].ofs, hi = lo;
        for (int k = 1; k < 4; ++k) { lo = std::min(lo, tabs[tab + x + k].ofs); hi = std::max(hi, tabs[tab + x + k].ofs); }
        if (hi + 1 - lo > 7) L.pyr_fast = 0;
      }
      if (getenv("ORB_B200_PYR_GENERIC")) L.pyr_fast = 0;
      tab += L.xtab_len + ytab_len;
    }
  }
  g.blur_tiles_total = btiles;
  g.ncells_total = cells; g.raw_total = raw; g.node_total = nodes; g.fast_tiles_total = ftiles;
  g.frame_stride = plane;
  e->geom = g;
  return ORB_OK;
}

// ---- workspace -------------------------------------------------------------------------------------------
static int default_chunk(const OrbGeom& g) {
  // Frames per internal pass. Every stage is issue/latency bound rather than DRAM bound (profiles/), so L2 residency of a
  // pass buys nothing and large passes win: they fill the 148 SMs for the small pyramid levels and the one-CTA-per-
  // (frame, level) quadtree. Bounded by ~1.5 GB of workspace per pass (two workspaces exist) and 64 frames, which keeps
  // >= 8 passes in a 512-frame call for the copy/compute pipeline of orb_extract_batch.
  const char* env = getenv("ORB_B200_CHUNK");
  if (env && atoi(env) > 0) return atoi(env);
  const double per_frame = 2.0 * (double)g.frame_stride + (double)g.ncells_total * (ORB_CELL_CAP + 1) * 4 +
                           (double)g.raw_total * 9 + (double)g.node_total * 100;
  int c = (int)(1.5e9 / per_frame);
  return std::max(1, std::min(c, 64));
}

static int ensure_shape(orb_extractor* e, int W, int H, int batch, int cap) {
  int rc;
  if (e->geom.width != W || e->geom.height != H) {
    std::vector<ResizeTab> tabs;
    if ((rc = build_geometry(e, W, H, tabs)) != ORB_OK) return rc;
    if ((rc = e->tabs.ensure(std::max<size_t>(tabs.size(), 1) * sizeof(ResizeTab))) != ORB_OK) return rc;
    if (!tabs.empty()) CU(cudaMemcpyAsync(e->tabs.p, tabs.data(), tabs.size() * sizeof(ResizeTab), cudaMemcpyHostToDevice, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    e->chunk = 0;
  }
  const OrbGeom& g = e->geom;
  const int want = std::min(batch, default_chunk(g));
  const int chunk_now = std::max(want, e->chunk);
  const int ws_needed = std::min(e->nws, (batch + chunk_now - 1) / chunk_now);   // later workspaces only exist once a call needs them
  if (want > e->chunk || cap > e->cap_ws || e->ws[ws_needed - 1].pyr.p == nullptr) {
    for (int w = 0; w < ORB_MAX_WS; ++w) CU(cudaStreamSynchronize(e->ws[w].st));
    const size_t B = (size_t)chunk_now;
    const int capw = std::max(cap, e->cap_ws);
    const int nl = g.nlevels;
    if ((rc = e->err.ensure(sizeof(int))) != ORB_OK) return rc;
    int nws = ws_needed;
    for (int w = 0; w < ORB_MAX_WS; ++w)
      if (e->ws[w].pyr.p) nws = std::max(nws, w + 1);   // existing workspaces grow with the others
    for (int w = 0; w < nws; ++w) {
      orb_extractor::PassWs& P = e->ws[w];
#define ENS(buf, bytes) if ((rc = P.buf.ensure(bytes)) != ORB_OK) return rc
    ENS(pyr, B * g.frame_stride + 256);   // +256: the resize window loads may run 12 bytes past a row end
    ENS(blr, B * g.frame_stride + 256);
    ENS(cell_cnt, B * g.ncells_total * sizeof(int));
    ENS(cell_kp, B * g.ncells_total * ORB_CELL_CAP * sizeof(uint32_t));
    ENS(kept, B * g.node_total * sizeof(uint32_t));
    ENS(kept_cnt, B * nl * sizeof(int));
    ENS(out_cnt, B * sizeof(int));
    ENS(q_key, B * g.raw_total * sizeof(uint32_t));
    ENS(q_knode, B * g.raw_total * sizeof(int));
    ENS(q_kchild, B * g.raw_total);
    ENS(q_bbox, B * g.node_total * sizeof(int4));
    ENS(q_ncount, B * g.node_total * sizeof(int));
    ENS(q_pidx, B * g.node_total * sizeof(int));
    ENS(q_best, B * g.node_total * sizeof(unsigned));
    ENS(q_ordA, B * g.node_total * sizeof(int));
    ENS(q_ordB, B * g.node_total * sizeof(int));
    ENS(q_par, B * g.node_total * sizeof(int));
    ENS(q_cand, B * g.node_total * 2 * sizeof(unsigned long long));
    ENS(q_ccnt, B * g.node_total * 4 * sizeof(int));
    ENS(q_cslot, B * g.node_total * 4 * sizeof(int));
    ENS(q_scan, B * g.node_total * sizeof(unsigned long long));
    ENS(q_rawn, B * nl * sizeof(int));
      P.frames = 0;
    }
#undef ENS
    CU(cudaMemsetAsync(e->err.p, 0, sizeof(int), e->stream));
    e->chunk = (int)B;
    e->cap_ws = capw;
  }
  return ORB_OK;
}

// One internal pass over `n` (<= chunk) device-resident frames; outputs to device pointers.
static void run_pass(orb_extractor* e, int w, const uint8_t* images, int step, long long image_stride, int n, OrbKeypoint* kps,
                     uint8_t* desc, int cap, int* counts_dev, cudaEvent_t ev_input_consumed = nullptr, int* level_counts_dev = nullptr) {
  const OrbGeom& g = e->geom;
  orb_extractor::PassWs& P = e->ws[w];
  cudaStream_t st = P.st;
  int* kept_cnt = level_counts_dev ? level_counts_dev : P.kept_cnt.as<int>();
  QtScratch S;
  S.key = P.q_key.as<uint32_t>(); S.knode = P.q_knode.as<int>(); S.kchild = P.q_kchild.as<uint8_t>();
  S.bbox = P.q_bbox.as<int4>(); S.ncount = P.q_ncount.as<int>(); S.pidx = P.q_pidx.as<int>();
  S.best = P.q_best.as<unsigned>(); S.ordA = P.q_ordA.as<int>(); S.ordB = P.q_ordB.as<int>(); S.par = P.q_par.as<int>();
  S.cand = P.q_cand.as<unsigned long long>(); S.ccnt = P.q_ccnt.as<int>(); S.cslot = P.q_cslot.as<int>();
  S.scan = P.q_scan.as<unsigned long long>(); S.raw_n = P.q_rawn.as<int>();
  bool prof = e->profile;
  cudaEvent_t* ev = nullptr;
  if (prof) {
    if (e->ev_used + ORB_NUM_STAGES + 1 > e->ev.size()) {
      const size_t old = e->ev.size();
      e->ev.resize(old + 64 * (ORB_NUM_STAGES + 1), nullptr);
      for (size_t i = old; i < e->ev.size(); ++i)
        if (cudaEventCreate(&e->ev[i]) != cudaSuccess) { e->ev.resize(old); prof = false; break; }
    }
    if (prof) { ev = e->ev.data() + e->ev_used; e->ev_used += ORB_NUM_STAGES + 1; }
  }
  if (prof) cudaEventRecord(ev[0], st);
  orb_launch_pyramid(images, step, image_stride, P.pyr.as<uint8_t>(), e->tabs.as<ResizeTab>(), g, n, st, &e->launches, ev_input_consumed);
  if (prof) cudaEventRecord(ev[1], st);
  orb_launch_fast(P.pyr.as<uint8_t>(), P.cell_cnt.as<int>(), P.cell_kp.as<uint32_t>(), e->err.as<int>(), g, n, st, &e->launches);
  if (prof) cudaEventRecord(ev[2], st);
  orb_launch_blur(P.pyr.as<uint8_t>(), P.blr.as<uint8_t>(), g, n, st, &e->launches);
  if (prof) cudaEventRecord(ev[3], st);
  orb_launch_quadtree(P.cell_cnt.as<int>(), P.cell_kp.as<uint32_t>(), S, P.kept.as<uint32_t>(), kept_cnt,
                      e->err.as<int>(), g, n, st, &e->launches);
  if (prof) cudaEventRecord(ev[4], st);
  orb_launch_describe(P.pyr.as<uint8_t>(), P.blr.as<uint8_t>(), P.kept.as<uint32_t>(), kept_cnt, kps, desc, cap,
                      counts_dev, e->err.as<int>(), g, n, st, &e->launches);
  if (prof) cudaEventRecord(ev[5], st);
  P.frames = n;
  e->last_ws = w;
}

static void resolve_stage_events(orb_extractor* e) {
  if (e->ev_used == 0) return;
  for (int w = 0; w < ORB_MAX_WS; ++w) cudaStreamSynchronize(e->ws[w].st);
  for (size_t b = 0; b + ORB_NUM_STAGES < e->ev_used; b += ORB_NUM_STAGES + 1) {
    for (int i = 0; i < ORB_NUM_STAGES; ++i) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, e->ev[b + i], e->ev[b + i + 1]) == cudaSuccess) e->stage_ms[i] += ms;
    }
    ++e->stage_passes;
  }
  e->ev_used = 0;
}

static int check_err_flag(orb_extractor* e) {
  int flag = 0;
  for (int w = 1; w < ORB_MAX_WS; ++w) CU(cudaStreamSynchronize(e->ws[w].st));
  CU(cudaMemcpyAsync(&flag, e->err.p, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  CU(cudaGetLastError());
  if (flag) {
    cudaMemsetAsync(e->err.p, 0, sizeof(int), e->stream);
    return fail(ORB_E_CAPACITY, "capacity overflow (flags %d: 1 cell cap, 2 raw keypoints, 4 nodes, 8 output cap)", flag);
  }
  return ORB_OK;
}

extern "C" {

const char* orb_last_error(void) { return t_err; }
int orb_abi_version(void) { return ORB_B200_ABI_VERSION; }
int orb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int orb_extractor_create(const orb_params* p, int device, orb_extractor** out) {
  if (!p || !out) return fail(ORB_E_BADARG, "null argument");
  *out = nullptr;
  if (p->nfeatures < 1 || p->nlevels < 1 || p->nlevels > ORB_MAX_LEVELS || !(p->scale_factor > 1.0f) || p->min_th_fast < 1 ||
      p->ini_th_fast < p->min_th_fast || p->ini_th_fast > 254)
    return fail(ORB_E_BADARG, "unsupported ORB parameters (nfeatures>=1, 1<=nlevels<=%d, scale>1, 1<=minTh<=iniTh<=254)", ORB_MAX_LEVELS);
  if (p->flags & ORB_FLAG_UPSTREAM_PYRAMID) return fail(ORB_E_BADARG, "ORB_FLAG_UPSTREAM_PYRAMID is not implemented on device");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  if (device < 0 || device >= ndev) return fail(ORB_E_BADARG, "device %d out of range (%d devices)", device, ndev);
  CU(cudaSetDevice(device));
  orb_extractor* e = new orb_extractor();
  e->prm = *p;
  e->device = device;
  memset(&e->geom, 0, sizeof(e->geom));
  build_tables(e);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) e->num_sms = prop.multiProcessorCount;
  cudaError_t ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
  for (int w = 1; w < ORB_MAX_WS; ++w) {
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->ws[w].st, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ws[w].ev_join, cudaEventDisableTiming);
  }
  if (const char* s = getenv("ORB_B200_STREAMS")) e->nws = std::min(std::max(atoi(s), 1), ORB_MAX_WS);
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming);
  e->ws[0].st = e->stream;
  if (ce != cudaSuccess) { orb_extractor_destroy(e); return fail(ORB_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(ce)); }
  *out = e;
  return ORB_OK;
}

void orb_extractor_destroy(orb_extractor* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) { cudaStreamSynchronize(e->stream); cudaStreamDestroy(e->stream); }
  for (int w = 1; w < ORB_MAX_WS; ++w) {
    if (e->ws[w].st) { cudaStreamSynchronize(e->ws[w].st); cudaStreamDestroy(e->ws[w].st); }
    if (e->ws[w].ev_join) cudaEventDestroy(e->ws[w].ev_join);
  }
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  for (int w = 0; w < ORB_MAX_WS; ++w)
    for (DevBuf* b : e->ws[w].all) b->release();
  DevBuf* bufs[] = {&e->tabs, &e->err, &e->cnt_all, &e->stage_in[0], &e->stage_in[1]};
  for (DevBuf* b : bufs) b->release();
  for (int b = 0; b <= ORB_MAX_WS; ++b) { e->out_kps[b].release(); e->out_desc[b].release(); e->out_ur[b].release(); e->out_dp[b].release(); }
  if (e->h_cnt) cudaFreeHost(e->h_cnt);
  if (e->s_in) cudaStreamDestroy(e->s_in);
  if (e->s_out) cudaStreamDestroy(e->s_out);
  for (int b = 0; b < 2; ++b) {
    if (e->ev_in[b]) cudaEventDestroy(e->ev_in[b]);
    if (e->ev_staged[b]) cudaEventDestroy(e->ev_staged[b]);
  }
  for (int b = 0; b < ORB_MAX_WS; ++b)
    if (e->ev_done[b]) cudaEventDestroy(e->ev_done[b]);
  for (int b = 0; b <= ORB_MAX_WS; ++b)
    if (e->ev_out[b]) cudaEventDestroy(e->ev_out[b]);
  for (cudaEvent_t ev : e->ev)
    if (ev) cudaEventDestroy(ev);
  delete e;
}

int orb_extractor_get_levels(const orb_extractor* e) { return e ? e->prm.nlevels : 0; }
float orb_extractor_get_scale_factor(const orb_extractor* e) { return e ? (float)(double)e->prm.scale_factor : 0.f; }
int orb_extractor_get_table(const orb_extractor* e, int what, float* out) {
  if (!e || !out || what < 0 || what > 3) return fail(ORB_E_BADARG, "bad argument");
  const std::vector<float>& t = what == 0 ? e->scale : what == 1 ? e->invScale : what == 2 ? e->sigma2 : e->invSigma2;
  std::copy(t.begin(), t.end(), out);
  return ORB_OK;
}
int orb_extractor_get_features_per_level(const orb_extractor* e, int32_t* out) {
  if (!e || !out) return fail(ORB_E_BADARG, "bad argument");
  std::copy(e->quota.begin(), e->quota.end(), out);
  return ORB_OK;
}
int orb_extractor_max_keypoints(const orb_extractor* e) {
  if (!e) return 0;
  // per level the list can end at max(quota + 3, 4 * nIni); nIni <= 4 for aspect <= 4.5 -> generous fixed slack
  return e->prm.nfeatures + 16 * e->prm.nlevels;
}
void* orb_extractor_stream(orb_extractor* e) { return e ? (void*)e->stream : nullptr; }

int orb_extractor_set_profiling(orb_extractor* e, int on) {
  if (!e) return fail(ORB_E_BADARG, "null argument");
  std::lock_guard<std::mutex> lk(e->mu);
  CU(cudaSetDevice(e->device));
  resolve_stage_events(e);
  e->profile = on != 0;
  for (int i = 0; i < ORB_NUM_STAGES; ++i) e->stage_ms[i] = 0.0;
  e->stage_passes = 0;
  return ORB_OK;
}

int orb_extractor_stage_times(orb_extractor* e, double* ms, int64_t* passes) {
  if (!e || !ms) return fail(ORB_E_BADARG, "null argument");
  std::lock_guard<std::mutex> lk(e->mu);
  CU(cudaSetDevice(e->device));
  resolve_stage_events(e);
  for (int i = 0; i < ORB_NUM_STAGES; ++i) ms[i] = e->stage_ms[i];
  if (passes) *passes = e->stage_passes;
  return ORB_NUM_STAGES;
}
int64_t orb_extractor_launch_count(const orb_extractor* e) { return e ? e->launches : 0; }

int orb_extract_batch_device(orb_extractor* e, const uint8_t* images, int step, int width, int height, int batch,
                             orb_keypoint* kps, uint8_t* desc, int cap, int32_t* counts_dev, int sync) {
  if (!e || !images || !kps || !desc || !counts_dev) return fail(ORB_E_BADARG, "null argument");
  if (width <= 0 || height <= 0 || step < width || batch <= 0 || cap <= 0) return fail(ORB_E_BADARG, "bad size");
  std::lock_guard<std::mutex> lk(e->mu);
  CU(cudaSetDevice(e->device));
  int rc = ensure_shape(e, width, height, batch, cap);
  if (rc != ORB_OK) return rc;
  const long long istride = (long long)step * height;
  // passes rotate over the workspaces / streams; stage profiling keeps everything on the public stream
  const int used = e->profile ? 1 : std::min(e->nws, (batch + e->chunk - 1) / e->chunk);
  if (used > 1) {
    CU(cudaEventRecord(e->ev_fork, e->stream));
    for (int w = 1; w < used; ++w) CU(cudaStreamWaitEvent(e->ws[w].st, e->ev_fork, 0));
  }
  int pass = 0;
  for (int f0 = 0; f0 < batch; f0 += e->chunk, ++pass) {
    const int n = std::min(e->chunk, batch - f0);
    run_pass(e, pass % used, images + (long long)f0 * istride, step, istride, n,
             reinterpret_cast<OrbKeypoint*>(kps) + (long long)f0 * cap, desc + (long long)f0 * cap * 32, cap, counts_dev + f0);
  }
  for (int w = 1; w < used; ++w) {
    CU(cudaEventRecord(e->ws[w].ev_join, e->ws[w].st));
    CU(cudaStreamWaitEvent(e->stream, e->ws[w].ev_join, 0));
  }
  CU(cudaGetLastError());
  if (sync) return check_err_flag(e);
  return ORB_OK;
}

}  // extern "C"

struct StereoReq {   // Frame::ComputeStereoMatches for the pairs (2p, 2p+1) of the batch
  float bf, b;
  float* uright;     // [batch / 2][cap], out_mem memory
  float* depth;
  int32_t* matched;  // [batch / 2] host, optional
};

static int extract_batch_impl(orb_extractor* e, const uint8_t* const* images, int in_mem, int step, int width, int height, int batch,
                              orb_keypoint* kps, uint8_t* desc, int out_mem, int cap, int32_t* counts, int32_t* level_counts,
                              const StereoReq* sr) {
  if (!e || !images || !kps || !desc || !counts) return fail(ORB_E_BADARG, "null argument");
  if (sr && ((batch & 1) || !sr->uright || !sr->depth || !(sr->b > 0.f) || !(sr->bf > 0.f) || cap > 65535))
    return fail(ORB_E_BADARG, "stereo: even batch (L,R interleaved), uright/depth buffers, bf > 0, b > 0 and cap <= 65535 required");
  if (width <= 0 || height <= 0 || step < width || batch <= 0 || cap <= 0) return fail(ORB_E_BADARG, "bad size");
  for (int i = 0; i < batch; ++i)
    if (!images[i]) return fail(ORB_E_BADARG, "null image %d", i);   // the reference returns silently on an empty image (:1966)
  std::lock_guard<std::mutex> lk(e->mu);
  CU(cudaSetDevice(e->device));
  int rc = ensure_shape(e, width, height, batch, cap);
  if (rc != ORB_OK) return rc;
  const OrbGeom& g = e->geom;
  const int nl = g.nlevels;
  const bool host_out = out_mem != ORB_MEM_DEVICE;
  if (sr && e->chunk < 2) return fail(ORB_E_BADARG, "stereo needs passes of at least 2 frames (ORB_B200_CHUNK)");
  // A copy-in and a copy-out stream beside the compute streams: the H2D copy of pass p+1 and the D2H copy of pass p-1 overlap the
  // kernels of pass p.
  if (!e->s_in) {
    CU(cudaStreamCreateWithFlags(&e->s_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&e->s_out, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      CU(cudaEventCreateWithFlags(&e->ev_in[b], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&e->ev_staged[b], cudaEventDisableTiming));
    }
    for (int b = 0; b < ORB_MAX_WS; ++b) CU(cudaEventCreateWithFlags(&e->ev_done[b], cudaEventDisableTiming));
    for (int b = 0; b <= ORB_MAX_WS; ++b) CU(cudaEventCreateWithFlags(&e->ev_out[b], cudaEventDisableTiming));
  }
  const size_t fbytes = (size_t)step * height;   // frames are staged with the caller's pitch (level 0 reads any pitch)
  const int nbuf = batch > e->chunk ? 2 : 1;
  const int nused = e->profile ? 1 : std::min(e->nws, (batch + e->chunk - 1) / e->chunk);   // workspaces ensure_shape has sized
  const int nout = batch > e->chunk ? nused + 1 : 1;
  if ((rc = e->cnt_all.ensure((size_t)batch * (2 + nl) * sizeof(int))) != ORB_OK) return rc;
  for (int b = 0; b < std::max(nout, nused); ++b) {
    if (b < nbuf && (rc = e->stage_in[b].ensure((size_t)e->chunk * fbytes)) != ORB_OK) return rc;
    if (host_out && b < nout) {
      if ((rc = e->out_kps[b].ensure((size_t)e->chunk * cap * sizeof(OrbKeypoint))) != ORB_OK) return rc;
      if ((rc = e->out_desc[b].ensure((size_t)e->chunk * cap * 32)) != ORB_OK) return rc;
      if (sr && (rc = e->out_ur[b].ensure((size_t)e->chunk / 2 * cap * sizeof(float) + 16)) != ORB_OK) return rc;
      if (sr && (rc = e->out_dp[b].ensure((size_t)e->chunk / 2 * cap * sizeof(float) + 16)) != ORB_OK) return rc;
    }
    if (sr && b < nused) {
      if ((rc = e->ws[b].st_sad.ensure((size_t)e->chunk / 2 * cap * sizeof(int) + 16)) != ORB_OK) return rc;
      if ((rc = e->ws[b].st_rows.ensure((size_t)e->chunk / 2 * (g.lv[0].rows + 1) * sizeof(int) + 16)) != ORB_OK) return rc;
      if ((rc = e->ws[b].st_sorted.ensure((size_t)e->chunk / 2 * cap * sizeof(unsigned short) + 16)) != ORB_OK) return rc;
    }
  }
  if (e->h_cnt_cap < (size_t)batch * (2 + nl)) {
    if (e->h_cnt) cudaFreeHost(e->h_cnt);
    e->h_cnt = nullptr; e->h_cnt_cap = 0;
    CU(cudaMallocHost(&e->h_cnt, (size_t)batch * (2 + nl) * sizeof(int)));
    e->h_cnt_cap = (size_t)batch * (2 + nl);
  }
  int* h_counts = e->h_cnt;
  int* h_levels = e->h_cnt + batch;
  int* h_matched = e->h_cnt + (size_t)batch * (1 + nl);
  int* d_counts = e->cnt_all.as<int>();
  int* d_levels = d_counts + batch;
  int* d_matched = d_counts + (size_t)batch * (1 + nl);
  const cudaMemcpyKind kin = in_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  // ORB_B200_TRACE=1: per-pass timeline of the three pipelines on stderr (development aid; adds timing events only)
  const bool trace = getenv("ORB_B200_TRACE") != nullptr;
  std::vector<cudaEvent_t> tev;
  std::vector<int> tn;
  std::vector<double> thost;   // host clock when each pass had been issued
  const auto t_entry = std::chrono::steady_clock::now();
  auto host_ms = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_entry).count(); };
  auto tmark = [&](cudaStream_t st) {
    if (!trace) return;
    cudaEvent_t ev;
    cudaEventCreate(&ev);
    cudaEventRecord(ev, st);
    tev.push_back(ev);
  };
  int pass = 0;
  for (int f0 = 0, n = 0; f0 < batch; f0 += n, ++pass) {
    // ramp-up: the first passes are short so that the first kernels start after a short copy (pipeline fill); the rest of
    // the batch is split evenly so that no short pass is left over at the end (its latency would be exposed)
    const int full = e->chunk;
    if (batch > 2 * full && pass < 2) {
      n = std::max(full >> (2 - pass), 1);
    } else {
      const int left = batch - f0, passes = (left + full - 1) / full;
      n = (left + passes - 1) / passes;
    }
    if (sr && (n & 1)) n = n < full ? n + 1 : n - 1;   // a pair never straddles two passes (batch is even, chunk >= 2)
    n = std::min(n, batch - f0);
    const int b = pass & 1, ob = pass % nout;
    // ---- copy-in stream: stage_in[b] is free once level 0 of pass-2 has read it (the rest of that pass works on the pyramid)
    if (pass >= 2) CU(cudaStreamWaitEvent(e->s_in, e->ev_staged[b], 0));
    tmark(e->s_in);
    bool contiguous = true;   // frames back to back with the same pitch -> one linear copy for the whole pass
    for (int i = 1; i < n && contiguous; ++i) contiguous = images[f0 + i] == images[f0] + (size_t)i * fbytes;
    if (contiguous) {
      CU(cudaMemcpyAsync(e->stage_in[b].p, images[f0], (size_t)n * fbytes - (step - width), kin, e->s_in));
    } else {
      for (int i = 0; i < n; ++i)
        CU(cudaMemcpyAsync(e->stage_in[b].as<uint8_t>() + (size_t)i * fbytes, images[f0 + i], fbytes - (step - width), kin, e->s_in));
    }
    CU(cudaEventRecord(e->ev_in[b], e->s_in));
    tmark(e->s_in);
    // ---- compute stream of workspace w (stage profiling serialises everything on the public stream)
    const int w = pass % nused;
    orb_extractor::PassWs& P = e->ws[w];
    CU(cudaStreamWaitEvent(P.st, e->ev_in[b], 0));
    if (pass >= nout) CU(cudaStreamWaitEvent(P.st, e->ev_out[ob], 0));   // output staging [ob] drained (pass - nout)
    tmark(P.st);
    OrbKeypoint* dk = host_out ? e->out_kps[ob].as<OrbKeypoint>() : reinterpret_cast<OrbKeypoint*>(kps) + (long long)f0 * cap;
    uint8_t* dd = host_out ? e->out_desc[ob].as<uint8_t>() : desc + (long long)f0 * cap * 32;
    run_pass(e, w, e->stage_in[b].as<uint8_t>(), step, (long long)fbytes, n, dk, dd, cap, d_counts + f0, e->ev_staged[b], d_levels + (size_t)f0 * nl);
    float* dur = nullptr;
    float* ddp = nullptr;
    if (sr) {
      dur = host_out ? e->out_ur[ob].as<float>() : sr->uright + (long long)(f0 / 2) * cap;
      ddp = host_out ? e->out_dp[ob].as<float>() : sr->depth + (long long)(f0 / 2) * cap;
      orb_launch_stereo(P.pyr.as<uint8_t>(), dk, dd, d_counts + f0, cap, sr->bf, sr->b, dur, ddp, P.st_sad.as<int>(),
                        d_matched + f0 / 2, P.st_rows.as<int>(), P.st_sorted.as<unsigned short>(), g, n / 2, P.st, &e->launches);
    }
    CU(cudaEventRecord(e->ev_done[w], P.st));
    tmark(P.st);
    // ---- copy-out stream; the count arrays of all passes are fetched once, after the last pass
    CU(cudaStreamWaitEvent(e->s_out, e->ev_done[w], 0));
    if (host_out) {
      CU(cudaMemcpyAsync(kps + (long long)f0 * cap, dk, (size_t)n * cap * sizeof(OrbKeypoint), cudaMemcpyDeviceToHost, e->s_out));
      CU(cudaMemcpyAsync(desc + (long long)f0 * cap * 32, dd, (size_t)n * cap * 32, cudaMemcpyDeviceToHost, e->s_out));
    }
    if (sr) {
      if (host_out) {
        CU(cudaMemcpyAsync(sr->uright + (long long)(f0 / 2) * cap, dur, (size_t)(n / 2) * cap * sizeof(float), cudaMemcpyDeviceToHost, e->s_out));
        CU(cudaMemcpyAsync(sr->depth + (long long)(f0 / 2) * cap, ddp, (size_t)(n / 2) * cap * sizeof(float), cudaMemcpyDeviceToHost, e->s_out));
      }
    }
    CU(cudaEventRecord(e->ev_out[ob], e->s_out));
    tmark(e->s_out);
    tn.push_back(n);
    if (trace) thost.push_back(host_ms());
  }
  // s_out has waited for every pass (ev_done of each), so all counts are final
  CU(cudaMemcpyAsync(h_counts, d_counts, (size_t)batch * (2 + nl) * sizeof(int), cudaMemcpyDeviceToHost, e->s_out));
  CU(cudaStreamSynchronize(e->s_out));
  if (trace) {
    fprintf(stderr, "host: all passes issued at %.3f ms, outputs on the host at %.3f ms after entry; pass issue times:", thost.back(), host_ms());
    for (double t : thost) fprintf(stderr, " %.3f", t);
    fprintf(stderr, "\n");
    for (int w = 0; w < ORB_MAX_WS; ++w) cudaStreamSynchronize(e->ws[w].st);
    fprintf(stderr, "pass frames | h2d start..end | compute start..end | d2h end   (ms since the first copy was issued)\n");
    for (size_t p = 0; p * 5 + 4 < tev.size(); ++p) {
      float t[5];
      for (int k = 0; k < 5; ++k) cudaEventElapsedTime(&t[k], tev[0], tev[p * 5 + k]);
      fprintf(stderr, "%4zu %4d | %7.3f %7.3f | %7.3f %7.3f | %7.3f\n", p, tn[p], t[0], t[1], t[2], t[3], t[4]);
    }
    for (cudaEvent_t ev : tev) cudaEventDestroy(ev);
  }
  if ((rc = check_err_flag(e)) != ORB_OK) return rc;   // synchronises the compute stream (counts are on the host after it)
  memcpy(counts, h_counts, (size_t)batch * sizeof(int));
  if (level_counts) memcpy(level_counts, h_levels, (size_t)batch * nl * sizeof(int));
  if (sr && sr->matched) memcpy(sr->matched, h_matched, (size_t)(batch / 2) * sizeof(int));
  return ORB_OK;
}

extern "C" {

int orb_extract_batch(orb_extractor* e, const uint8_t* const* images, int in_mem, int step, int width, int height, int batch,
                      orb_keypoint* kps, uint8_t* desc, int out_mem, int cap, int32_t* counts, int32_t* level_counts) {
  return extract_batch_impl(e, images, in_mem, step, width, height, batch, kps, desc, out_mem, cap, counts, level_counts, nullptr);
}

int orb_extract_stereo_batch(orb_extractor* e, const uint8_t* const* images, int in_mem, int step, int width, int height, int pairs,
                             orb_keypoint* kps, uint8_t* desc, int out_mem, int cap, int32_t* counts, int32_t* level_counts, float bf,
                             float b, float* uright, float* depth, int32_t* matched) {
  const StereoReq sr{bf, b, uright, depth, matched};
  if (pairs <= 0) return fail(ORB_E_BADARG, "bad size");
  return extract_batch_impl(e, images, in_mem, step, width, height, 2 * pairs, kps, desc, out_mem, cap, counts, level_counts, &sr);
}

int orb_extractor_pyramid_dims(const orb_extractor* e, int level, int* cols, int* rows) {
  if (!e || !cols || !rows || e->geom.width == 0 || level < 0 || level >= e->geom.nlevels) return fail(ORB_E_BADARG, "bad argument / no image processed yet");
  *cols = e->geom.lv[level].cols;
  *rows = e->geom.lv[level].rows;
  return ORB_OK;
}

int orb_extractor_pyramid_read(orb_extractor* e, int index, int level, int blurred, uint8_t* dst, int dst_step) {
  if (!e || !dst || e->geom.width == 0 || level < 0 || level >= e->geom.nlevels) return fail(ORB_E_BADARG, "bad argument");
  std::lock_guard<std::mutex> lk(e->mu);
  orb_extractor::PassWs& P = e->ws[e->last_ws];
  if (index == -1) index = P.frames - 1;   // the last image of the last call
  if (index < 0 || index >= P.frames) return fail(ORB_E_BADARG, "image %d is not resident (last pass held %d frames)", index, P.frames);
  const OrbLevel& L = e->geom.lv[level];
  if (dst_step < L.cols) return fail(ORB_E_BADARG, "dst_step too small");
  CU(cudaSetDevice(e->device));
  const uint8_t* src = (blurred ? P.blr.as<uint8_t>() : P.pyr.as<uint8_t>()) + (size_t)index * e->geom.frame_stride + L.plane_off;
  CU(cudaMemcpy2DAsync(dst, dst_step, src, L.pitch, L.cols, L.rows, cudaMemcpyDeviceToHost, P.st));
  CU(cudaStreamSynchronize(P.st));
  return ORB_OK;
}

int orb_extractor_raw_keypoints(orb_extractor* e, int index, int level, orb_keypoint* out, int cap) {
  if (!e || !out || e->geom.width == 0 || level < 0 || level >= e->geom.nlevels) return fail(ORB_E_BADARG, "bad argument");
  std::lock_guard<std::mutex> lk(e->mu);
  orb_extractor::PassWs& P = e->ws[e->last_ws];
  if (index < 0 || index >= P.frames) return fail(ORB_E_BADARG, "image %d is not resident", index);
  CU(cudaSetDevice(e->device));
  const OrbGeom& g = e->geom;
  const OrbLevel& L = g.lv[level];
  int n = 0;
  CU(cudaMemcpyAsync(&n, P.q_rawn.as<int>() + (size_t)index * g.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost, P.st));
  CU(cudaStreamSynchronize(P.st));
  n = std::min(n, L.raw_cap);
  if (n > cap) return fail(ORB_E_CAPACITY, "need room for %d raw keypoints", n);
  std::vector<uint32_t> keys(std::max(n, 1));
  CU(cudaMemcpyAsync(keys.data(), P.q_key.as<uint32_t>() + (size_t)index * g.raw_total + L.raw_off, (size_t)n * 4, cudaMemcpyDeviceToHost, P.st));
  CU(cudaStreamSynchronize(P.st));
  for (int i = 0; i < n; ++i) {
    out[i].x = (float)(keys[i] & 0xfffu); out[i].y = (float)((keys[i] >> 12) & 0xfffu);
    out[i].size = 7.f; out[i].angle = -1.f; out[i].response = (float)(keys[i] >> 24);
    out[i].octave = 0; out[i].class_id = -1;
  }
  return n;
}

// ---- matcher ------------------------------------------------------------------------------------------------
struct KnnWorkspace {   // per-device scratch, grown on demand; guarded by a mutex (calls on one device serialise their setup)
  std::mutex mu;
  DevBuf partial, q, db, out;
  // orb_search_by_projection: sorted keypoints + cell starts + per-query top-2, and the staging copies of host arguments
  DevBuf g_bq, g_bnq, g_stage[16];   // orb_search_by_bow: device-built queries, staging of host arguments
  DevBuf g_items, g_cells, g_lists, g_lcnt, g_kps, g_desc, g_ur, g_claim, g_nk, g_q, g_qd, g_nq, g_assigned, g_qmatch, g_nm;
  int num_sms = 0;
};
static KnnWorkspace g_knn_ws[16];

static int current_device_ws(KnnWorkspace** ws) {
  int dev = 0;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16) return fail(ORB_E_BADARG, "device index %d unsupported", dev);
  *ws = &g_knn_ws[dev];
  if ((*ws)->num_sms == 0) {
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    (*ws)->num_sms = prop.multiProcessorCount;
  }
  return ORB_OK;
}

int orb_hamming_knn(const uint8_t* query, int q_mem, int Q, const uint8_t* db, int db_mem, int64_t D, int64_t index_base,
                    int sentinel, orb_knn2* out, int out_mem, void* stream) {
  if (!query || !out || Q <= 0 || D < 0 || (D > 0 && !db) || sentinel < 1) return fail(ORB_E_BADARG, "bad argument");
  if (index_base < 0 || index_base + D >= (1LL << 40)) return fail(ORB_E_BADARG, "index range exceeds 2^40");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  const uint8_t* dq = query;
  const uint8_t* ddb = db;
  orb_knn2* dout = out;
  const bool any_host = q_mem != ORB_MEM_DEVICE || db_mem != ORB_MEM_DEVICE || out_mem != ORB_MEM_DEVICE;
  if (q_mem != ORB_MEM_DEVICE || ((uintptr_t)query & 15)) {
    if ((rc = ws->q.ensure((size_t)Q * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->q.p, query, (size_t)Q * 32, q_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    dq = ws->q.as<uint8_t>();
  }
  if (D > 0 && (db_mem != ORB_MEM_DEVICE || ((uintptr_t)db & 15))) {
    if ((rc = ws->db.ensure((size_t)D * 32)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->db.p, db, (size_t)D * 32, db_mem == ORB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    ddb = ws->db.as<uint8_t>();
  }
  if (out_mem != ORB_MEM_DEVICE) {
    if ((rc = ws->out.ensure((size_t)Q * sizeof(orb_knn2))) != ORB_OK) return rc;
    dout = ws->out.as<orb_knn2>();
  }
  const int nchunks = orb_knn_num_chunks(Q, D, ws->num_sms);
  if ((rc = ws->partial.ensure((size_t)nchunks * Q * 2 * sizeof(unsigned long long))) != ORB_OK) return rc;
  orb_launch_knn(dq, Q, ddb, D, index_base, sentinel, ws->partial.as<unsigned long long>(), nchunks, (OrbKnn2*)dout, st);
  CU(cudaGetLastError());
  if (out_mem != ORB_MEM_DEVICE) CU(cudaMemcpyAsync(out, dout, (size_t)Q * sizeof(orb_knn2), cudaMemcpyDeviceToHost, st));
  if (any_host) CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

int orb_hamming_knn_merge(const orb_knn2* shards, int nshards, int Q, int sentinel, orb_knn2* out, void* stream) {
  if (!shards || !out || nshards < 1 || Q <= 0) return fail(ORB_E_BADARG, "bad argument");
  orb_launch_knn_merge_shards((const OrbKnn2*)shards, nshards, Q, sentinel, (OrbKnn2*)out, (cudaStream_t)stream);
  CU(cudaGetLastError());
  return ORB_OK;
}

int orb_descriptor_distance_batch(const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out, void* stream) {
  if (!a || !b || !out || n < 0) return fail(ORB_E_BADARG, "bad argument");
  if (((uintptr_t)a & 15) || ((uintptr_t)b & 15)) return fail(ORB_E_BADARG, "descriptor arrays must be 16-byte aligned");
  if (n == 0) return ORB_OK;
  orb_launch_distance_batch(a, b, n, out, (cudaStream_t)stream);
  CU(cudaGetLastError());
  return ORB_OK;
}

int orb_descriptor_distance_host(const uint8_t* a, const uint8_t* b, int64_t n, int32_t* out) {
  if (!a || !b || !out || n < 0) return fail(ORB_E_BADARG, "bad argument");
  if (n == 0) return ORB_OK;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  std::lock_guard<std::mutex> lk(ws->mu);
  if ((rc = ws->q.ensure((size_t)n * 32)) != ORB_OK) return rc;
  if ((rc = ws->db.ensure((size_t)n * 32)) != ORB_OK) return rc;
  if ((rc = ws->out.ensure((size_t)n * sizeof(int))) != ORB_OK) return rc;
  CU(cudaMemcpyAsync(ws->q.p, a, (size_t)n * 32, cudaMemcpyHostToDevice, 0));
  CU(cudaMemcpyAsync(ws->db.p, b, (size_t)n * 32, cudaMemcpyHostToDevice, 0));
  orb_launch_distance_batch(ws->q.as<uint8_t>(), ws->db.as<uint8_t>(), n, ws->out.as<int>(), 0);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out, ws->out.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, 0));
  CU(cudaStreamSynchronize(0));
  return ORB_OK;
}

int64_t orb_matcher_launch_count(void) { return orb_knn_launches(); }

// ---- vocabulary / ComputeBoW ------------------------------------------------------------------------------------------------

struct orb_vocabulary {
  int device = 0, k = 0, L = 0, n_nodes = 0, n_words = 0;
  DevBuf child_start, child_id, child_desc, is_leaf, word_of, word_weight;
  DevBuf s_desc, s_n, s_word, s_node, s_bw, s_bv, s_bc, s_fn, s_fs, s_ff, s_fc;   // staging for host-memory calls
  std::mutex mu;
};

int orb_vocabulary_create(int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc,
                          const double* weight, int device, orb_vocabulary** out) {
  if (!out) return fail(ORB_E_BADARG, "null argument");
  *out = nullptr;
  if (!parent || !is_leaf || !desc || !weight || k < 1 || L < 1 || n_nodes < 2) return fail(ORB_E_BADARG, "bad vocabulary arguments");
  std::vector<int> cnt(n_nodes + 1, 0);
  for (int i = 1; i < n_nodes; ++i) {
    if (parent[i] < 0 || parent[i] >= n_nodes || parent[i] == i) return fail(ORB_E_BADARG, "node %d: parent %d out of range", i, parent[i]);
    if (is_leaf[parent[i]] && parent[i] != 0) return fail(ORB_E_BADARG, "node %d hangs below leaf %d", i, parent[i]);
    ++cnt[parent[i] + 1];
  }
  for (int i = 1; i < n_nodes; ++i)
    if (!is_leaf[i] && cnt[i + 1] == 0) return fail(ORB_E_BADARG, "inner node %d has no children", i);
  if (cnt[1] == 0) return fail(ORB_E_BADARG, "root has no children");
  std::vector<int> start(n_nodes + 1, 0);
  for (int i = 0; i < n_nodes; ++i) start[i + 1] = start[i] + cnt[i + 1];
  const int slots = start[n_nodes];   // == n_nodes - 1
  std::vector<int> cursor(start.begin(), start.end() - 1), child_id(slots), word_of(n_nodes, -1);
  std::vector<uint8_t> child_desc((size_t)slots * 32), leaf(n_nodes, 0);
  std::vector<double> word_weight;
  for (int i = 1; i < n_nodes; ++i) {   // file order: children in push_back order, words numbered as the leaves appear
    const int s = cursor[parent[i]]++;
    child_id[s] = i;
    memcpy(&child_desc[(size_t)s * 32], desc + (size_t)i * 32, 32);
    leaf[i] = is_leaf[i] ? 1 : 0;
    if (leaf[i]) { word_of[i] = (int)word_weight.size(); word_weight.push_back(weight[i]); }
  }
  // a walk deeper than n_nodes levels would mean a cycle; parents precede nothing in particular, so check by depth
  for (int i = 1; i < n_nodes; ++i) {
    int d = 0, p = i;
    while (p != 0 && d <= n_nodes) { p = parent[p]; ++d; }
    if (p != 0) return fail(ORB_E_BADARG, "node %d is not connected to the root", i);
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  CU(cudaSetDevice(device));
  orb_vocabulary* v = new orb_vocabulary;
  v->device = device; v->k = k; v->L = L; v->n_nodes = n_nodes; v->n_words = (int)word_weight.size();
  struct Up { DevBuf* b; const void* src; size_t bytes; } ups[] = {
      {&v->child_start, start.data(), start.size() * sizeof(int)}, {&v->child_id, child_id.data(), child_id.size() * sizeof(int)},
      {&v->child_desc, child_desc.data(), child_desc.size()},      {&v->is_leaf, leaf.data(), leaf.size()},
      {&v->word_of, word_of.data(), word_of.size() * sizeof(int)}, {&v->word_weight, word_weight.data(), word_weight.size() * sizeof(double)}};
  for (const Up& u : ups) {
    int rc = u.b->ensure(std::max<size_t>(u.bytes, 16));
    if (rc == ORB_OK && cudaMemcpy(u.b->p, u.src, u.bytes, cudaMemcpyHostToDevice) != cudaSuccess) rc = fail(ORB_E_CUDA, "vocabulary upload failed");
    if (rc != ORB_OK) { orb_vocabulary_destroy(v); return rc; }
  }
  *out = v;
  return ORB_OK;
}

void orb_vocabulary_destroy(orb_vocabulary* v) {
  if (!v) return;
  cudaSetDevice(v->device);
  DevBuf* bufs[] = {&v->child_start, &v->child_id, &v->child_desc, &v->is_leaf, &v->word_of, &v->word_weight, &v->s_desc, &v->s_n, &v->s_word,
                    &v->s_node, &v->s_bw, &v->s_bv, &v->s_bc, &v->s_fn, &v->s_fs, &v->s_ff, &v->s_fc};
  for (DevBuf* b : bufs) b->release();
  delete v;
}

int orb_vocabulary_words(const orb_vocabulary* v) { return v ? v->n_words : 0; }

int orb_bow_transform(orb_vocabulary* v, int mem, const uint8_t* desc, const int32_t* n, int cap, int frames, int levelsup, int32_t* word,
                      int32_t* node, int32_t* bow_word, double* bow_value, int32_t* bow_count, int32_t* fv_node, int32_t* fv_start,
                      int32_t* fv_feature, int32_t* fv_count, void* stream) {
  if (!v || !desc || !n || !bow_word || !bow_value || !bow_count || !fv_node || !fv_start || !fv_feature || !fv_count)
    return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap > 16384 || frames <= 0 || levelsup < 0) return fail(ORB_E_BADARG, "bad size (0 < cap <= 16384, frames > 0, levelsup >= 0)");
  std::lock_guard<std::mutex> lk(v->mu);
  CU(cudaSetDevice(v->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t F = (size_t)frames;
  const bool host = mem != ORB_MEM_DEVICE;
  int rc;
  const uint8_t* d_desc = desc; const int* d_n = n;
  int *d_word = word, *d_node = node, *d_bw = bow_word, *d_bc = bow_count, *d_fn = fv_node, *d_fs = fv_start, *d_ff = fv_feature, *d_fc = fv_count;
  double* d_bv = bow_value;
  if (host) {
    struct S { DevBuf* b; size_t bytes; } st_bufs[] = {{&v->s_desc, F * cap * 32}, {&v->s_n, F * 4}, {&v->s_bw, F * cap * 4}, {&v->s_bv, F * cap * 8},
                                                      {&v->s_bc, F * 4}, {&v->s_fn, F * cap * 4}, {&v->s_fs, F * (cap + 1) * 4},
                                                      {&v->s_ff, F * cap * 4}, {&v->s_fc, F * 4}};
    for (const S& s : st_bufs) if ((rc = s.b->ensure(s.bytes)) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(v->s_desc.p, desc, F * cap * 32, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(v->s_n.p, n, F * 4, cudaMemcpyHostToDevice, st));
    d_desc = v->s_desc.as<uint8_t>(); d_n = v->s_n.as<int>();
    d_bw = v->s_bw.as<int>(); d_bv = v->s_bv.as<double>(); d_bc = v->s_bc.as<int>();
    d_fn = v->s_fn.as<int>(); d_fs = v->s_fs.as<int>(); d_ff = v->s_ff.as<int>(); d_fc = v->s_fc.as<int>();
  } else if ((uintptr_t)desc & 15) {
    return fail(ORB_E_BADARG, "descriptor array must be 16-byte aligned device memory");
  }
  if (host || !word) { if ((rc = v->s_word.ensure(F * cap * 4)) != ORB_OK) return rc; d_word = v->s_word.as<int>(); }
  if (host || !node) { if ((rc = v->s_node.ensure(F * cap * 4)) != ORB_OK) return rc; d_node = v->s_node.as<int>(); }
  BowVoc V{v->child_start.as<int>(), v->child_id.as<int>(), v->child_desc.as<uint4>(), v->is_leaf.as<uint8_t>(), v->word_of.as<int>(),
           v->word_weight.as<double>(), v->L};
  long long launched = 0;
  orb_launch_bow(V, levelsup, d_desc, d_n, cap, frames, d_word, d_node, d_bw, d_bv, d_bc, d_fn, d_fs, d_ff, d_fc, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    if (word) CU(cudaMemcpyAsync(word, d_word, F * cap * 4, cudaMemcpyDeviceToHost, st));
    if (node) CU(cudaMemcpyAsync(node, d_node, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(bow_word, d_bw, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(bow_value, d_bv, F * cap * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(bow_count, d_bc, F * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_node, d_fn, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_start, d_fs, F * (cap + 1) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_feature, d_ff, F * cap * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fv_count, d_fc, F * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

int orb_search_by_bow(int mem, const uint8_t* kf_desc, const orb_keypoint* kf_kps_un, const uint8_t* kf_valid, const int32_t* kf_fv_node,
                      const int32_t* kf_fv_start, const int32_t* kf_fv_feature, const int32_t* kf_fv_count, int qcap, const uint8_t* f_desc,
                      const orb_keypoint* f_kps, const int32_t* f_fv_node, const int32_t* f_fv_start, const int32_t* f_fv_feature,
                      const int32_t* f_fv_count, int cap, int frames, int th_dist, float nn_ratio, int check_orientation, int32_t* assigned,
                      int32_t* nmatches, void* stream) {
  const void* in[14] = {kf_desc, kf_kps_un, kf_valid, kf_fv_node, kf_fv_start, kf_fv_feature, kf_fv_count,
                        f_desc,  f_kps,     f_fv_node, f_fv_start, f_fv_feature, f_fv_count, nullptr};
  for (int i = 0; i < 13; ++i)
    if (!in[i]) return fail(ORB_E_BADARG, "null argument");
  if (!assigned || !nmatches) return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap >= (1 << 18) || qcap <= 0 || qcap >= (1 << 18) || frames <= 0) return fail(ORB_E_BADARG, "bad size (0 < cap, qcap < 2^18, frames > 0)");
  if (th_dist < 0 || th_dist > 255 || !(nn_ratio > 0.f)) return fail(ORB_E_BADARG, "th_dist in [0, 255] and nn_ratio > 0 required");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  const size_t F = (size_t)frames, Q = (size_t)qcap, N = (size_t)cap;
  const bool host = mem != ORB_MEM_DEVICE;
  int* d_assigned = assigned;
  int* d_nm = nmatches;
  if (host) {
    const size_t bytes[13] = {F * Q * 32, F * Q * sizeof(orb_keypoint), F * Q, F * Q * 4, F * (Q + 1) * 4, F * Q * 4, F * 4,
                              F * N * 32, F * N * sizeof(orb_keypoint), F * N * 4, F * (N + 1) * 4, F * N * 4, F * 4};
    for (int i = 0; i < 13; ++i) {
      if ((rc = ws->g_stage[i].ensure(bytes[i])) != ORB_OK) return rc;
      CU(cudaMemcpyAsync(ws->g_stage[i].p, in[i], bytes[i], cudaMemcpyHostToDevice, st));
      in[i] = ws->g_stage[i].p;
    }
    if ((rc = ws->g_assigned.ensure(F * N * sizeof(int))) != ORB_OK) return rc;
    if ((rc = ws->g_nm.ensure(F * sizeof(int))) != ORB_OK) return rc;
    d_assigned = ws->g_assigned.as<int>();
    d_nm = ws->g_nm.as<int>();
  }
  if ((rc = ws->g_bq.ensure(F * Q * 32)) != ORB_OK) return rc;
  if ((rc = ws->g_bnq.ensure(F * sizeof(int))) != ORB_OK) return rc;
  if ((rc = ws->g_qmatch.ensure(F * Q * sizeof(int))) != ORB_OK) return rc;
  if ((rc = ws->g_lists.ensure(F * Q * 32 * sizeof(uint2))) != ORB_OK) return rc;
  if ((rc = ws->g_lcnt.ensure(F * Q * sizeof(int))) != ORB_OK) return rc;
  long long launched = 0;
  orb_launch_search_by_bow((const uint8_t*)in[0], (const OrbKeypoint*)in[1], (const uint8_t*)in[2], (const int*)in[3], (const int*)in[4],
                           (const int*)in[5], (const int*)in[6], qcap, (const uint8_t*)in[7], (const OrbKeypoint*)in[8], (const int*)in[9],
                           (const int*)in[10], (const int*)in[11], (const int*)in[12], cap, frames, th_dist, nn_ratio, check_orientation,
                           ws->g_bq.p, ws->g_bnq.as<int>(), ws->g_lists.as<uint2>(), ws->g_lcnt.as<int>(), d_assigned, ws->g_qmatch.as<int>(),
                           d_nm, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    CU(cudaMemcpyAsync(assigned, d_assigned, F * N * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(nmatches, d_nm, F * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

int orb_undistort_keypoints(int mem, const orb_camera* cam, const orb_keypoint* kps, const int32_t* n, int cap, int frames,
                            orb_keypoint* kps_un, void* stream) {
  if (!cam || !kps || !n || !kps_un || cap <= 0 || frames <= 0) return fail(ORB_E_BADARG, "bad argument");
  if (cam->n_dist != 4 && cam->n_dist != 5 && cam->n_dist != 8) return fail(ORB_E_BADARG, "n_dist must be 4, 5 or 8");
  if (!(cam->fx != 0.f) || !(cam->fy != 0.f)) return fail(ORB_E_BADARG, "fx / fy must be non-zero");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  const size_t bytes = (size_t)frames * cap * sizeof(orb_keypoint);
  if (mem == ORB_MEM_DEVICE) {
    orb_launch_undistort(cam, (const OrbKeypoint*)kps, n, cap, frames, (OrbKeypoint*)kps_un, st);
    orb_knn_count_launches(1);
    CU(cudaGetLastError());
    return ORB_OK;
  }
  if ((rc = ws->g_kps.ensure(bytes)) != ORB_OK) return rc;
  if ((rc = ws->g_nk.ensure((size_t)frames * sizeof(int))) != ORB_OK) return rc;
  CU(cudaMemcpyAsync(ws->g_kps.p, kps, bytes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ws->g_nk.p, n, (size_t)frames * sizeof(int), cudaMemcpyHostToDevice, st));
  orb_launch_undistort(cam, ws->g_kps.as<OrbKeypoint>(), ws->g_nk.as<int>(), cap, frames, ws->g_kps.as<OrbKeypoint>(), st);
  orb_knn_count_launches(1);
  CU(cudaGetLastError());
  if (kps_un != kps) memcpy(kps_un, kps, bytes);   // rows >= n[f] keep the caller's content (the kernel only rewrites rows < n[f])
  for (int f = 0; f < frames; ++f) {
    const int m = std::min(std::max(n[f], 0), cap);
    if (m > 0)
      CU(cudaMemcpyAsync(kps_un + (size_t)f * cap, ws->g_kps.as<orb_keypoint>() + (size_t)f * cap, (size_t)m * sizeof(orb_keypoint),
                         cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

int orb_image_bounds(const orb_camera* cam, int cols, int rows, float* bounds) {
  if (!cam || !bounds || cols <= 0 || rows <= 0) return fail(ORB_E_BADARG, "bad argument");
  if (cam->dist[0] == 0.0f) {   // src/Frame.cc:861-867
    bounds[0] = 0.f; bounds[1] = 0.f; bounds[2] = (float)cols; bounds[3] = (float)rows;
    return ORB_OK;
  }
  orb_keypoint c[4];
  memset(c, 0, sizeof(c));
  c[1].x = (float)cols; c[2].y = (float)rows; c[3].x = (float)cols; c[3].y = (float)rows;   // :838-845
  const int32_t four = 4;
  const int rc = orb_undistort_keypoints(ORB_MEM_HOST, cam, c, &four, 4, 1, c, nullptr);
  if (rc != ORB_OK) return rc;
  bounds[0] = std::min(c[0].x, c[2].x);   // mnMinX (:855)
  bounds[2] = std::max(c[1].x, c[3].x);   // mnMaxX
  bounds[1] = std::min(c[0].y, c[1].y);   // mnMinY
  bounds[3] = std::max(c[2].y, c[3].y);   // mnMaxY
  return ORB_OK;
}

int orb_distinctive_descriptors(int mem, const uint8_t* desc, const int64_t* offsets, int points, int32_t* best_idx,
                                int32_t* best_median, void* stream) {
  if (!offsets || !best_idx || points < 0) return fail(ORB_E_BADARG, "bad argument");
  if (points == 0) return ORB_OK;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  const bool host = mem != ORB_MEM_DEVICE;
  if (!host) {
    if (!desc || ((uintptr_t)desc & 15)) return fail(ORB_E_BADARG, "descriptor array must be 16-byte aligned device memory");
    orb_launch_distinctive(desc, (const long long*)offsets, points, best_idx, best_median, st);
    CU(cudaGetLastError());
    return ORB_OK;
  }
  const int64_t total = offsets[points];
  if (offsets[0] != 0 || total < 0 || (total > 0 && !desc)) return fail(ORB_E_BADARG, "offsets must start at 0 and be ascending");
  for (int p = 0; p < points; ++p)
    if (offsets[p + 1] < offsets[p] || offsets[p + 1] - offsets[p] >= (1 << 22)) return fail(ORB_E_BADARG, "offsets not ascending / more than 2^22 observations");
  if ((rc = ws->db.ensure((size_t)std::max<int64_t>(total, 1) * 32)) != ORB_OK) return rc;
  if ((rc = ws->q.ensure((size_t)(points + 1) * sizeof(int64_t))) != ORB_OK) return rc;
  if ((rc = ws->out.ensure((size_t)points * 2 * sizeof(int))) != ORB_OK) return rc;
  if (total > 0) CU(cudaMemcpyAsync(ws->db.p, desc, (size_t)total * 32, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(ws->q.p, offsets, (size_t)(points + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  int* d_idx = ws->out.as<int>();
  int* d_med = d_idx + points;
  orb_launch_distinctive(ws->db.as<uint8_t>(), ws->q.as<long long>(), points, d_idx, d_med, st);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(best_idx, d_idx, (size_t)points * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (best_median) CU(cudaMemcpyAsync(best_median, d_med, (size_t)points * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return ORB_OK;
}

int orb_search_by_projection(int mem, const orb_keypoint* kps_un, const uint8_t* desc, const float* uright, const uint8_t* claimed,
                             const int32_t* n_kps, int cap, const orb_proj_query* queries, const uint8_t* qdesc,
                             const int32_t* n_queries, int qcap, int frames, const orb_search_params* prm, int32_t* assigned,
                             int32_t* query_match, int32_t* nmatches, void* stream) {
  static_assert(sizeof(orb_proj_query) == 32 && sizeof(orb_search_params) == 32, "ABI struct layout");
  if (!kps_un || !desc || !n_kps || !queries || !qdesc || !n_queries || !prm || !assigned || !nmatches)
    return fail(ORB_E_BADARG, "null argument");
  if (cap <= 0 || cap >= (1 << 18) || qcap <= 0 || frames <= 0) return fail(ORB_E_BADARG, "bad size (0 < cap < 2^18, qcap > 0, frames > 0)");
  if (!(prm->max_x > prm->min_x) || !(prm->max_y > prm->min_y)) return fail(ORB_E_BADARG, "empty image bounds");
  if (prm->th_dist < 0 || prm->th_dist > 255) return fail(ORB_E_BADARG, "th_dist must be in [0, 255]");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  KnnWorkspace* ws;
  int rc = current_device_ws(&ws);
  if (rc != ORB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(ws->mu);
  const size_t F = (size_t)frames;
  const bool host = mem != ORB_MEM_DEVICE;
  // device views of the arguments (staged when the caller's arrays are host memory)
  const void* d_in[7] = {kps_un, desc, uright, claimed, n_kps, queries, qdesc};
  const void* d_nq = n_queries;
  int* d_assigned = assigned;
  int* d_qmatch = query_match;
  int* d_nm = nmatches;
  if (host) {
    DevBuf* bufs[7] = {&ws->g_kps, &ws->g_desc, &ws->g_ur, &ws->g_claim, &ws->g_nk, &ws->g_q, &ws->g_qd};
    const size_t bytes[7] = {F * cap * sizeof(orb_keypoint), F * cap * 32, F * cap * sizeof(float), F * cap, F * sizeof(int),
                             F * qcap * sizeof(orb_proj_query), F * qcap * 32};
    for (int i = 0; i < 7; ++i) {
      if (!d_in[i]) continue;
      if ((rc = bufs[i]->ensure(bytes[i])) != ORB_OK) return rc;
      CU(cudaMemcpyAsync(bufs[i]->p, d_in[i], bytes[i], cudaMemcpyHostToDevice, st));
      d_in[i] = bufs[i]->p;
    }
    if ((rc = ws->g_nq.ensure(F * sizeof(int))) != ORB_OK) return rc;
    CU(cudaMemcpyAsync(ws->g_nq.p, n_queries, F * sizeof(int), cudaMemcpyHostToDevice, st));
    d_nq = ws->g_nq.p;
    if ((rc = ws->g_assigned.ensure(F * cap * sizeof(int))) != ORB_OK) return rc;
    if ((rc = ws->g_nm.ensure(F * sizeof(int))) != ORB_OK) return rc;
    d_assigned = ws->g_assigned.as<int>();
    d_nm = ws->g_nm.as<int>();
  }
  if (host || !query_match) {   // the resolve kernel always keeps the per-query record (rotation filter)
    if ((rc = ws->g_qmatch.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
    d_qmatch = ws->g_qmatch.as<int>();
  }
  if ((rc = ws->g_items.ensure(F * cap * sizeof(uint4))) != ORB_OK) return rc;
  if ((rc = ws->g_cells.ensure(F * (64 * 48 + 1) * sizeof(int))) != ORB_OK) return rc;
  if ((rc = ws->g_lists.ensure(F * qcap * 32 * sizeof(uint2))) != ORB_OK) return rc;
  if ((rc = ws->g_lcnt.ensure(F * qcap * sizeof(int))) != ORB_OK) return rc;
  long long launched = 0;
  orb_launch_guided((const OrbKeypoint*)d_in[0], (const uint8_t*)d_in[1], (const float*)d_in[2], (const uint8_t*)d_in[3], (const int*)d_in[4],
                    cap, d_in[5], (const uint8_t*)d_in[6], (const int*)d_nq, qcap, frames, prm->min_x, prm->min_y, prm->max_x, prm->max_y,
                    prm->th_dist, prm->nn_ratio, prm->check_orientation, ws->g_items.as<uint4>(), ws->g_cells.as<int>(),
                    ws->g_lists.as<uint2>(), ws->g_lcnt.as<int>(), d_assigned, d_qmatch, d_nm, st, &launched);
  orb_knn_count_launches(launched);
  CU(cudaGetLastError());
  if (host) {
    CU(cudaMemcpyAsync(assigned, d_assigned, F * cap * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (query_match) CU(cudaMemcpyAsync(query_match, d_qmatch, F * qcap * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(nmatches, d_nm, F * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return ORB_OK;
}

double orb_measure_popc_peak(int device) {
  if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return 0.0; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 0.0;
  return orb_popc_peak(prop.multiProcessorCount);
}

}  // extern "C"
