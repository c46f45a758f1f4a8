// C ABI (include/orb_b200.h) over the sm_100a kernels, extractor part. Host logic only: parameter tables exactly as the
// reference constructor computes them (src/ORBextractor.cc:776-925), per-shape geometry, workspaces, launches.
// No CPU fallback: every compute entry point needs a CUDA device.
#include "orb_host.h"

// ---- errors ------------------------------------------------------------------------------------------
static thread_local char t_err[512] = "";
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
  return code;
}

static inline int cv_round_f(float v) { return (int)lrintf(v); }

#define ORB_MAX_WS 4
struct orb_extractor {
  orb_params prm;
  int device = 0, num_sms = 148;
  cudaStream_t stream = nullptr;
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> quota;
  OrbGeom geom;            // valid when geom.width > 0
  int chunk = 0;           // frames per internal pass (workspace capacity)
  int pass_frames = 0;     // frames per pass of the current call (<= chunk; ensure_shape)
  int cap_ws = 0;
  long long launches = 0;
  DevBuf tabs, err, fast_tiles, rs_tiles;
  // Two pass workspaces on two compute streams: consecutive internal passes alternate between them, so the latency-bound
  // tail of pass p (quadtree, descriptors) overlaps the bandwidth/issue-bound head of pass p+1 (pyramid, FAST, blur).
  // ws[0].st is the handle's public stream.
  struct PassWs {
    cudaStream_t st = nullptr;
    cudaEvent_t ev_join = nullptr;
    cudaStream_t st_side = nullptr;                       // high-priority side stream of the pass: quadtree (latency-bound) beside blur
    cudaEvent_t ev_fast = nullptr, ev_side = nullptr, ev_pyr = nullptr;   // ev_pyr: the pyramid of the pass is complete
    cudaEvent_t ev_pfork = nullptr, ev_pjoin = nullptr;                   // pyramid chain on the side stream (default; ORB_B200_PYR_MAIN=1 turns it off)
    int frames = 0;          // frames of the last pass run here
    // per-level TMA descriptors of `pyr` / `blr`: FAST band windows (256 B x hCell + 6), IC_Angle windows (48 B x 31), descriptor patches (64 B x 37)
    alignas(64) unsigned char fast_maps[ORB_LEVEL_MAPS_BYTES], ic_maps[ORB_LEVEL_MAPS_BYTES], pt_maps[ORB_LEVEL_MAPS_BYTES],
        blur_maps[ORB_LEVEL_MAPS_BYTES],   // blur tiles: 160 B x 40 rows of pyr
        rs_maps[ORB_LEVEL_MAPS_BYTES];     // tiled resize: 192 B x 44 rows of the source level
    bool maps_ok = false;
    DevBuf pyr, blr, cell_cnt, cell_kp, kept, kept_cnt, out_cnt;
    DevBuf st_sad;   // stereo: SAD of every kept match [chunk/2][cap]
    DevBuf st_rows, st_sorted;   // stereo: right keypoints sorted by image row: starts [chunk/2][rows0 + 1], indices [chunk/2][cap]
    DevBuf q_key, q_knode, q_kchild, q_bbox, q_ncount, q_pidx, q_best, q_ordA, q_ordB, q_par, q_cand, q_ccnt, q_cslot, q_scan, q_rawn;
    DevBuf* all[25] = {&st_rows, &st_sorted, &st_sad, &pyr, &blr, &cell_cnt, &cell_kp, &kept, &kept_cnt, &out_cnt, &q_key, &q_knode, &q_kchild, &q_bbox, &q_ncount,
                       &q_pidx, &q_best, &q_ordA, &q_ordB, &q_par, &q_cand, &q_ccnt, &q_cslot, &q_scan, &q_rawn};
  } ws[ORB_MAX_WS];
  int nws = 4;             // compute streams / workspaces in use (ORB_B200_STREAMS, 1..ORB_MAX_WS)
  int side = 2;            // ORB_B200_SIDE: 0 = off, 1 = quadtree, 2 (default) = quadtree + descriptors on the pass's high-priority side stream
  int last_ws = 0;         // workspace of the most recent pass (mvImagePyramid / test hooks read it)
  cudaEvent_t ev_fork = nullptr;
  // orb_extractor_set_pyramid_download: host destinations of the padded levels of the LAST image of every host-facing call; the copy is issued on
  // the copy-out stream as soon as the pyramid kernels of the last pass are done, so it overlaps FAST / quadtree / descriptors
  uint8_t* pyr_dst[ORB_MAX_LEVELS] = {};
  int pyr_dst_step[ORB_MAX_LEVELS] = {};
  bool pyr_dst_on = false;
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
static void resize_table(int ssz, int dsz, int pad, int lead, int len, ResizeTab* out) {
  const double scale = 1.0 / ((double)dsz / ssz);
  const int padded = dsz + 2 * pad;
  for (int t = 0; t < len; ++t) {
    const int p = std::min(t - lead, padded + ORB_APRON - 1);
    const int d = orb_reflect101(orb_reflect101(p, padded) - pad, dsz);
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
static int build_geometry(orb_extractor* e, int W, int H, std::vector<ResizeTab>& tabs, std::vector<int>& rs_tiles) {
  OrbGeom g;
  memset(&g, 0, sizeof(g));
  const int nl = e->prm.nlevels;
  g.nlevels = nl; g.width = W; g.height = H;
  g.ini_th = e->prm.ini_th_fast; g.min_th = e->prm.min_th_fast;
  g.pad = (e->prm.flags & ORB_FLAG_UPSTREAM_PYRAMID) ? 0 : ORB_EDGE;
  {
    const char* th = getenv("ORB_B200_BLUR_TH");
    g.blur_th = (th && (atoi(th) == 32 || atoi(th) == 128)) ? atoi(th) : 64;   // 64: +3.4 % on the step against 32 (longer-lived CTAs, two rounds of loads in flight), 128: the same
  }
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
  rs_tiles.clear();
  for (int l = 0; l < nl; ++l) {
    OrbLevel& L = g.lv[l];
    L.iw = cv_round_f((float)W * e->invScale[l]);
    L.ih = cv_round_f((float)H * e->invScale[l]);
    if (L.iw < 2 * ORB_EDGE + 2 || L.ih < 2 * ORB_EDGE + 2)
      return fail(ORB_E_SHAPE, "level %d is %dx%d: too small for the 19-px reflect border", l, L.iw, L.ih);
    L.cols = L.iw + 2 * g.pad; L.rows = L.ih + 2 * g.pad;
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
    L.fast_G = (L.nCols + L.fast_nseg - 1) / L.fast_nseg;          // same number of tiles, balanced widths (41 cells: 7,7,7,7,7,6 instead of 8 x 5 + 1)
    if (L.fast_G * L.wCell > ORB_FAST_TILE_W) return fail(ORB_E_SHAPE, "level %d: FAST cell wider than a tile", l);   // wCell < 60 by construction
    L.fast_tile_base = ftiles;
    L.fast_inv_nseg = (1u << 20) / (unsigned)L.fast_nseg + 1u;   // t / nseg == (t * inv) >> 20 while t * nseg < 2^20 (nseg == 1 included)
    if ((long long)L.fast_nseg * L.nRows * L.fast_nseg >= (1 << 20)) return fail(ORB_E_SHAPE, "level %d: FAST tile index out of the reciprocal's range", l);
    L.inv_wcell = (unsigned)((1ull << 32) / (unsigned)L.wCell + 1ull);
    ftiles += L.fast_nseg * L.nRows;
    L.blur_tiles_x = (L.cols + 127) / 128;
    L.blur_tile_base = btiles;
    L.blur_inv_tx = (1u << 20) / (unsigned)L.blur_tiles_x + 1u;
    btiles += L.blur_tiles_x * ((L.rows + g.blur_th - 1) / g.blur_th);
    L.nIni = (int)roundf((float)(L.maxBX - ORB_MIN_BORDER) / (L.maxBY - ORB_MIN_BORDER));
    if (L.nIni < 1) return fail(ORB_E_SHAPE, "aspect ratio < 0.5: the reference divides by nIni == 0 (src/ORBextractor.cc:1059)");
    L.hX = (float)(L.maxBX - ORB_MIN_BORDER) / L.nIni;
    L.quota = e->quota[l];
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
      resize_table(P.cols, L.iw, g.pad, ORB_APRON, L.xtab_len, tabs.data() + tab);   // source = stored level l-1: PADDED in this fork (:2119), the plain image upstream
      resize_table(P.rows, L.ih, g.pad, ORB_APRON - 1, ytab_len, tabs.data() + tab + L.xtab_len);
      // the 4-pixel kernel reads an 8-byte source window per thread: needs ofs(x+3) + 1 - ofs(x) <= 7
      L.pyr_fast = 1;
      for (int x = 0; x + 3 < L.xtab_len && L.pyr_fast; x += 4) {
        int lo = tabs[tab + x].ofs, hi = lo;
        for (int k = 1; k < 4; ++k) { lo = std::min(lo, tabs[tab + x + k].ofs); hi = std::max(hi, tabs[tab + x + k].ofs); }
        if (hi + 1 - lo > 7) L.pyr_fast = 0;
      }
      if (getenv("ORB_B200_PYR_GENERIC")) L.pyr_fast = 0;
      // tiled variant: per 128-column / 32-row tile the origin of its source window; usable if every window fits the TMA box
      {
        const ResizeTab* xt = tabs.data() + tab;
        const ResizeTab* yt = xt + L.xtab_len;
        L.rs_ntx = (L.xtab_len + ORB_RS_TW - 1) / ORB_RS_TW;
        L.rs_nty = (yrows + ORB_RS_TH - 1) / ORB_RS_TH;
        L.rs_off = (int)rs_tiles.size();
        std::vector<int> ynr;   // source rows a tile row needs (the y table folds back inside the reflected border: not derivable from its ends)
        L.pyr_tiled = L.pyr_fast;
        for (int tx = 0; tx < L.rs_ntx; ++tx) {
          int lo = INT_MAX, hi = 0;
          for (int x = tx * ORB_RS_TW; x < std::min((tx + 1) * ORB_RS_TW, L.xtab_len); ++x) { lo = std::min(lo, xt[x].ofs); hi = std::max(hi, xt[x].ofs + 1); }
          lo &= ~15;
          if (hi - lo + 8 > ORB_RS_BW) L.pyr_tiled = 0;   // + 8: a thread reads an aligned 12-byte window around its 8 source bytes
          rs_tiles.push_back(lo);
        }
        for (int ty = 0; ty < L.rs_nty; ++ty) {
          int lo = INT_MAX, hi = 0;
          for (int y = ty * ORB_RS_TH; y < std::min((ty + 1) * ORB_RS_TH, yrows); ++y) { lo = std::min(lo, yt[y].ofs); hi = std::max(hi, std::min(yt[y].ofs + 1, P.rows - 1)); }
          if (hi - lo + 1 > ORB_RS_BH) L.pyr_tiled = 0;
          rs_tiles.push_back(lo);
          ynr.push_back(hi - lo + 1);
        }
        rs_tiles.insert(rs_tiles.end(), ynr.begin(), ynr.end());
      }
      tab += L.xtab_len + ytab_len;
    }
  }
  g.blur_tiles_total = btiles;
  for (int l = 0; l < ORB_MAX_LEVELS; ++l) {
    g.blur_base[l] = l < g.nlevels ? g.lv[l].blur_tile_base : INT_MAX;
    g.fast_base[l] = l < g.nlevels ? g.lv[l].fast_tile_base : INT_MAX;
  }
  // Capacities that cannot overflow: NMS survivors are pairwise non-adjacent (8-neighbourhood), so a cell of wCell x hCell detection pixels holds at
  // most ceil(wCell / 2) * ceil(hCell / 2) of them, and a level at most that times its cells (the reference's vectors simply grow, :1451).
  g.cell_cap = 4;
  for (int l = 0; l < nl; ++l) g.cell_cap = std::max(g.cell_cap, ((g.lv[l].wCell + 1) / 2) * ((g.lv[l].hCell + 1) / 2));
  g.cell_cap = (int)align_up(g.cell_cap, 4);
  for (int l = 0; l < nl; ++l) {
    OrbLevel& L = g.lv[l];
    const long long by_cells = (long long)L.nCols * L.nRows * (((L.wCell + 1) / 2) * ((L.hCell + 1) / 2));
    if (by_cells > (1LL << 30)) return fail(ORB_E_SHAPE, "level %d: raw keypoint capacity out of range", l);
    L.raw_cap = (int)by_cells;
    L.raw_off = raw; raw += (int)align_up(L.raw_cap, 32);
  }
  g.ncells_total = cells; g.raw_total = raw; g.node_total = nodes; g.fast_tiles_total = ftiles;
  g.frame_stride = plane;
  e->geom = g;
  return ORB_OK;
}

// ---- workspace -------------------------------------------------------------------------------------------
static int default_chunk(const OrbGeom& g) {
  // Frames per internal pass. Every stage is issue/latency bound rather than DRAM bound (profiles/), so L2 residency of a pass buys nothing;
  // what counts is that a pass fills the 148 SMs for the small pyramid levels and the one-CTA-per-(frame, level) quadtree, and that several
  // passes are in flight so that the chains of small launches hide behind the other passes' large kernels. With the pyramid on the
  // high-priority stream, four workspaces of 32-frame passes are the best point on every shape (profiles/r2_stream_sweep.md: KITTI 95.4 k
  // against 92.1 k images/s host to host with two 64-frame passes, TUM 152.0 k against 146.7 k, EuRoC 123.9 k against 118.8 k), bounded by
  // ~0.9 GB of workspace per pass (4K: 6 frames).
  const char* env = getenv("ORB_B200_CHUNK");
  if (env && atoi(env) > 0) return atoi(env);
  const double per_frame = 2.0 * (double)g.frame_stride + (double)g.ncells_total * (g.cell_cap + 1) * 4 +
                           (double)g.raw_total * 9 + (double)g.node_total * 100;
  int c = (int)(0.9e9 / per_frame);
  return std::max(1, std::min(c, 32));
}

static int ensure_shape(orb_extractor* e, int W, int H, int batch, int cap, bool copies = false) {
  int rc;
  if (e->geom.width != W || e->geom.height != H) {
    std::vector<ResizeTab> tabs;
    std::vector<int> rs_tiles;
    if ((rc = build_geometry(e, W, H, tabs, rs_tiles)) != ORB_OK) return rc;
    if ((rc = e->rs_tiles.ensure(std::max<size_t>(rs_tiles.size(), 1) * sizeof(int))) != ORB_OK) return rc;
    if (!rs_tiles.empty()) CU(cudaMemcpy(e->rs_tiles.p, rs_tiles.data(), rs_tiles.size() * sizeof(int), cudaMemcpyHostToDevice));
    if ((rc = e->tabs.ensure(std::max<size_t>(tabs.size(), 1) * sizeof(ResizeTab))) != ORB_OK) return rc;
    if (!tabs.empty()) CU(cudaMemcpyAsync(e->tabs.p, tabs.data(), tabs.size() * sizeof(ResizeTab), cudaMemcpyHostToDevice, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    std::vector<int4> tiles;
    orb_fast_build_tiles(e->geom, tiles);
    if ((rc = e->fast_tiles.ensure(std::max<size_t>(tiles.size(), 1) * sizeof(int4))) != ORB_OK) return rc;
    CU(cudaMemcpy(e->fast_tiles.p, tiles.data(), tiles.size() * sizeof(int4), cudaMemcpyHostToDevice));
    e->chunk = 0;
    e->pyr_dst_on = false;   // level sizes changed: the caller registers new destinations
  }
  const OrbGeom& g = e->geom;
  const int want = std::min(batch, default_chunk(g));
  const int chunk_now = std::max(want, e->chunk);
  // Frames per pass of THIS call. A call that copies its frames in or its results out overlaps those copies with the kernels pass by pass, so a
  // mid-size call is cut into about eight passes instead of one or two full ones (same box, host to host: KITTI 64 frames 52.3 k -> 56.7 k
  // images/s, 4K 16 frames 4.3 k -> 5.4 k), as long as a pass keeps ~3.5 M pixels to fill the SMs with; large calls keep the full pass size.
  int psz = chunk_now;
  if (copies && !e->profile) {
    const int fill = (int)std::min<long long>((3500000LL + (long long)W * H - 1) / ((long long)W * H), chunk_now);
    psz = std::min(chunk_now, std::max({fill, (batch + 7) / 8, std::min(2, chunk_now)}));
  }
  e->pass_frames = psz;
  const int ws_needed = std::min(e->nws, (batch + psz - 1) / psz);   // later workspaces only exist once a call needs them
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
    ENS(cell_kp, B * g.ncells_total * g.cell_cap * sizeof(uint32_t));
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
      // TMA descriptors follow the (re)allocated pyramid buffer; without the driver entry point the kernel stages with cp.async
      P.maps_ok = orb_build_level_maps(P.pyr.as<uint8_t>(), (int)B, g, 256, 0, P.fast_maps) == 0 &&
                  orb_build_level_maps(P.pyr.as<uint8_t>(), (int)B, g, 48, 31, P.ic_maps) == 0 &&
                  orb_build_level_maps(P.pyr.as<uint8_t>(), (int)B, g, 160, 40, P.blur_maps) == 0 &&
                  orb_build_level_maps(P.pyr.as<uint8_t>(), (int)B, g, ORB_RS_BW, ORB_RS_BH, P.rs_maps) == 0 &&
                  orb_build_level_maps(P.blr.as<uint8_t>(), (int)B, g, 64, 37, P.pt_maps) == 0;
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
                     uint8_t* desc, int cap, int* counts_dev, cudaEvent_t ev_input_consumed = nullptr, int* level_counts_dev = nullptr,
                     bool mark_pyramid = false) {
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
  // The pyramid's chain of small dependent launches runs on the pass's high-priority side stream: its CTAs are dispatched ahead of the queued
  // FAST / blur CTAs of the other pass instead of waiting for that grid to drain, the chain's latency hides behind them, and the input staging
  // buffer is released earlier for the next copy (118.3 k vs 116.7 k images/s resident, 91.9 k vs 88.5 k host to host; ORB_B200_PYR_MAIN=1 = before)
  static const bool pyr_side = getenv("ORB_B200_PYR_MAIN") == nullptr;
  const bool ps = pyr_side && !prof && e->side >= 1;
  if (ps) { cudaEventRecord(P.ev_pfork, st); cudaStreamWaitEvent(P.st_side, P.ev_pfork, 0); }
  orb_launch_pyramid(images, step, image_stride, P.pyr.as<uint8_t>(), e->tabs.as<ResizeTab>(), e->rs_tiles.as<int>(), P.maps_ok ? P.rs_maps : nullptr,
                     e->num_sms, g, n, ps ? P.st_side : st, &e->launches, ev_input_consumed);
  if (ps) { cudaEventRecord(P.ev_pjoin, P.st_side); cudaStreamWaitEvent(st, P.ev_pjoin, 0); }
  if (mark_pyramid) cudaEventRecord(P.ev_pyr, st);
  if (prof) cudaEventRecord(ev[1], st);
  static const bool fast_v1 = getenv("ORB_B200_FAST_V1") != nullptr;   // round-1 kernel, same-box A/B only
  if (fast_v1) orb_launch_fast_v1(P.pyr.as<uint8_t>(), P.cell_cnt.as<int>(), P.cell_kp.as<uint32_t>(), e->err.as<int>(), g, n, st, &e->launches);
  else orb_launch_fast(P.pyr.as<uint8_t>(), e->fast_tiles.p, P.maps_ok ? P.fast_maps : nullptr, P.cell_cnt.as<int>(), P.cell_kp.as<uint32_t>(),
                       e->err.as<int>(), g, n, st, &e->launches);
  if (prof) cudaEventRecord(ev[2], st);
  // The quadtree is latency-bound (one CTA per frame x level, ~30 % issue-active) and only needs the FAST cells; the blur only needs
  // the pyramid. Outside stage profiling the quadtree (side >= 1) and the descriptors (side >= 2) go to the pass's high-priority
  // side stream: their CTAs are dispatched ahead of the queued blur / FAST CTAs of both pass streams and run beside them.
  const int side = prof ? 0 : e->side;
  cudaStream_t sq = side >= 1 ? P.st_side : st;
  if (side >= 1) { cudaEventRecord(P.ev_fast, st); cudaStreamWaitEvent(sq, P.ev_fast, 0); }
  if (side == 0) {
    orb_launch_blur(P.pyr.as<uint8_t>(), P.blr.as<uint8_t>(), P.maps_ok ? P.blur_maps : nullptr, g, n, st, &e->launches);
    if (prof) cudaEventRecord(ev[3], st);
  }
  orb_launch_quadtree(P.cell_cnt.as<int>(), P.cell_kp.as<uint32_t>(), S, P.kept.as<uint32_t>(), kept_cnt,
                      e->err.as<int>(), g, n, sq, &e->launches);
  if (prof) cudaEventRecord(ev[4], st);
  if (side >= 1) orb_launch_blur(P.pyr.as<uint8_t>(), P.blr.as<uint8_t>(), P.maps_ok ? P.blur_maps : nullptr, g, n, st, &e->launches);
  cudaStream_t sd = side >= 2 ? P.st_side : st;
  if (side == 1) { cudaEventRecord(P.ev_side, sq); cudaStreamWaitEvent(st, P.ev_side, 0); }
  if (side >= 2) { cudaEventRecord(P.ev_fast, st); cudaStreamWaitEvent(sd, P.ev_fast, 0); }      // blur done
  orb_launch_describe(P.pyr.as<uint8_t>(), P.blr.as<uint8_t>(), P.maps_ok ? P.ic_maps : nullptr, P.maps_ok ? P.pt_maps : nullptr, P.kept.as<uint32_t>(), kept_cnt, kps, desc, cap,
                      counts_dev, e->err.as<int>(), g, n, sd, &e->launches);
  if (side >= 2) { cudaEventRecord(P.ev_side, sd); cudaStreamWaitEvent(st, P.ev_side, 0); }
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
    return fail(ORB_E_CAPACITY, "capacity overflow (flags %d: 1 cell cap, 2 raw keypoints, 4 nodes, 8 output cap, 16 TMA window load failed)", flag);
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
  if (p->flags & ~(ORB_FLAG_UPSTREAM_PYRAMID | ORB_FLAG_LEGACY_BLUR)) return fail(ORB_E_BADARG, "unknown bits in orb_params.flags (%d)", p->flags);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(ORB_E_CUDA, "no CUDA device (there is no CPU fallback)"); }
  if (device < 0 || device >= ndev) return fail(ORB_E_BADARG, "device %d out of range (%d devices)", device, ndev);
  ORB_ON_DEVICE(device);
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
  if (const char* s = getenv("ORB_B200_SIDE")) e->side = atoi(s);
  {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);   // hi = numerically lowest = greatest priority
    for (int w = 0; w < ORB_MAX_WS; ++w) {
      if (ce == cudaSuccess) ce = cudaStreamCreateWithPriority(&e->ws[w].st_side, cudaStreamNonBlocking, hi);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ws[w].ev_fast, cudaEventDisableTiming);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ws[w].ev_side, cudaEventDisableTiming);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ws[w].ev_pyr, cudaEventDisableTiming);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ws[w].ev_pfork, cudaEventDisableTiming);
      if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ws[w].ev_pjoin, cudaEventDisableTiming);
    }
  }
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming);
  e->ws[0].st = e->stream;
  if (ce != cudaSuccess) { orb_extractor_destroy(e); return fail(ORB_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(ce)); }
  *out = e;
  return ORB_OK;
}

void orb_extractor_destroy(orb_extractor* e) {
  if (!e) return;
  OrbDeviceGuard guard(e->device);
  if (e->stream) { cudaStreamSynchronize(e->stream); cudaStreamDestroy(e->stream); }
  for (int w = 0; w < ORB_MAX_WS; ++w) {
    if (e->ws[w].st_side) { cudaStreamSynchronize(e->ws[w].st_side); cudaStreamDestroy(e->ws[w].st_side); }
    if (e->ws[w].ev_fast) cudaEventDestroy(e->ws[w].ev_fast);
    if (e->ws[w].ev_side) cudaEventDestroy(e->ws[w].ev_side);
    if (e->ws[w].ev_pyr) cudaEventDestroy(e->ws[w].ev_pyr);
    if (e->ws[w].ev_pfork) cudaEventDestroy(e->ws[w].ev_pfork);
    if (e->ws[w].ev_pjoin) cudaEventDestroy(e->ws[w].ev_pjoin);
  }
  for (int w = 1; w < ORB_MAX_WS; ++w) {
    if (e->ws[w].st) { cudaStreamSynchronize(e->ws[w].st); cudaStreamDestroy(e->ws[w].st); }
    if (e->ws[w].ev_join) cudaEventDestroy(e->ws[w].ev_join);
  }
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  for (int w = 0; w < ORB_MAX_WS; ++w)
    for (DevBuf* b : e->ws[w].all) b->release();
  DevBuf* bufs[] = {&e->tabs, &e->err, &e->fast_tiles, &e->rs_tiles, &e->cnt_all, &e->stage_in[0], &e->stage_in[1]};
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
  ORB_ON_DEVICE(e->device);
  resolve_stage_events(e);
  e->profile = on != 0;
  for (int i = 0; i < ORB_NUM_STAGES; ++i) e->stage_ms[i] = 0.0;
  e->stage_passes = 0;
  return ORB_OK;
}

int orb_extractor_stage_times(orb_extractor* e, double* ms, int64_t* passes) {
  if (!e || !ms) return fail(ORB_E_BADARG, "null argument");
  std::lock_guard<std::mutex> lk(e->mu);
  ORB_ON_DEVICE(e->device);
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
  ORB_ON_DEVICE(e->device);
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
  ORB_ON_DEVICE(e->device);
  const bool host_out = out_mem != ORB_MEM_DEVICE;
  int rc = ensure_shape(e, width, height, batch, cap, host_out || in_mem != ORB_MEM_DEVICE);
  if (rc != ORB_OK) return rc;
  const OrbGeom& g = e->geom;
  const int nl = g.nlevels;
  const int psz = e->pass_frames;
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
  const int nbuf = batch > psz ? 2 : 1;
  const int nused = e->profile ? 1 : std::min(e->nws, (batch + psz - 1) / psz);   // workspaces ensure_shape has sized
  const int nout = batch > psz ? nused + 1 : 1;
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
    const int full = psz;
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
    const bool last_pass = f0 + n >= batch;
    const bool pyr_out = last_pass && e->pyr_dst_on;
    run_pass(e, w, e->stage_in[b].as<uint8_t>(), step, (long long)fbytes, n, dk, dd, cap, d_counts + f0, e->ev_staged[b], d_levels + (size_t)f0 * nl,
             pyr_out);
    if (pyr_out) {   // the padded levels of the call's last image go home while the rest of the pass still runs
      CU(cudaStreamWaitEvent(e->s_out, P.ev_pyr, 0));
      for (int l = 0; l < nl; ++l) {
        const OrbLevel& L = g.lv[l];
        CU(cudaMemcpy2DAsync(e->pyr_dst[l], (size_t)e->pyr_dst_step[l], P.pyr.as<uint8_t>() + (size_t)(n - 1) * g.frame_stride + L.plane_off, L.pitch,
                             L.cols, L.rows, cudaMemcpyDeviceToHost, e->s_out));
      }
    }
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

int orb_extractor_check(orb_extractor* e) {
  if (!e) return fail(ORB_E_BADARG, "null argument");
  std::lock_guard<std::mutex> lk(e->mu);
  ORB_ON_DEVICE(e->device);
  if (!e->err.p) return ORB_OK;   // nothing has run yet
  return check_err_flag(e);
}

int orb_host_alloc(size_t bytes, void** out) {
  if (!out || bytes == 0) return fail(ORB_E_BADARG, "bad argument");
  *out = nullptr;
  CU(cudaMallocHost(out, bytes));
  return ORB_OK;
}
void orb_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

int orb_extractor_set_pyramid_download(orb_extractor* e, uint8_t* const* level_dst, const int32_t* level_steps) {
  if (!e) return fail(ORB_E_BADARG, "null argument");
  std::lock_guard<std::mutex> lk(e->mu);
  if (!level_dst) { e->pyr_dst_on = false; return ORB_OK; }
  if (!level_steps || e->geom.width == 0) return fail(ORB_E_BADARG, "bad argument / no image processed yet (level sizes unknown)");
  for (int l = 0; l < e->geom.nlevels; ++l) {
    if (!level_dst[l] || level_steps[l] < e->geom.lv[l].cols) return fail(ORB_E_BADARG, "level %d: null destination or step < %d", l, e->geom.lv[l].cols);
    e->pyr_dst[l] = level_dst[l];
    e->pyr_dst_step[l] = level_steps[l];
  }
  e->pyr_dst_on = true;
  return ORB_OK;
}

int orb_extractor_pyramid_read_all(orb_extractor* e, int index, int blurred, uint8_t* const* level_dst, const int32_t* level_steps) {
  if (!e || !level_dst || !level_steps || e->geom.width == 0) return fail(ORB_E_BADARG, "bad argument / no image processed yet");
  std::lock_guard<std::mutex> lk(e->mu);
  orb_extractor::PassWs& P = e->ws[e->last_ws];
  if (index == -1) index = P.frames - 1;
  if (index < 0 || index >= P.frames) return fail(ORB_E_BADARG, "image %d is not resident (last pass held %d frames)", index, P.frames);
  ORB_ON_DEVICE(e->device);
  for (int l = 0; l < e->geom.nlevels; ++l) {
    const OrbLevel& L = e->geom.lv[l];
    if (!level_dst[l] || level_steps[l] < L.cols) return fail(ORB_E_BADARG, "level %d: null destination or step < %d", l, L.cols);
    const uint8_t* src = (blurred ? P.blr.as<uint8_t>() : P.pyr.as<uint8_t>()) + (size_t)index * e->geom.frame_stride + L.plane_off;
    CU(cudaMemcpy2DAsync(level_dst[l], (size_t)level_steps[l], src, L.pitch, L.cols, L.rows, cudaMemcpyDeviceToHost, P.st));
  }
  CU(cudaStreamSynchronize(P.st));   // one synchronisation for all levels
  return ORB_OK;
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
  ORB_ON_DEVICE(e->device);
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
  ORB_ON_DEVICE(e->device);
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

}  // extern "C"
