// Pyramid kernels: ORBextractor::ComputePyramid (src/ORBextractor.cc:2093-2168) of THIS fork:
// level 0 = copyMakeBorder(image, 19, REFLECT_101); level l = resize(PADDED level l-1 -> w_l x h_l,
// INTER_LINEAR) + in-place copyMakeBorder(REFLECT_101|ISOLATED); mvImagePyramid[l] = padded image.
// Integer-only restatement of OpenCV's 8-bit fixed-point resize -> bit-exact by construction.
// Border pixels are pure index remaps of interior pixels, so every padded output pixel is computed
// directly from level l-1 (no second pass over the plane).
#include <algorithm>
#include <cstdlib>

#include "orb_common.cuh"
#include "orb_tma.cuh"

// Programmatic dependent launch (small passes only, see orb_launch_pyramid): the next kernel of the level chain may be scheduled while
// this one drains; it blocks in orb_pdl_wait() until this grid has completed and its writes are visible. Both are no-ops for a
// kernel launched the ordinary way. Every kernel of the chain waits before it reads OR exits, so completion stays transitive.
__device__ __forceinline__ void orb_pdl_begin() {
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Every plane is written together with its apron: columns [-4, cols + 4) and rows [-3, rows + 3) hold the
// BORDER_REFLECT_101 extension of the PADDED image (what cv::GaussianBlur reads beyond it), so the blur kernel is branch-free.
// Kernel x index t <-> padded column t - ORB_APRON, y index t <-> padded row t - (ORB_APRON - 1).

// Level 0 = copyMakeBorder(image, 19, REFLECT_101) + apron, in 16-byte chunks of the destination row (chunk c = columns
// [-16 + 16c, 16c)). Interior chunks (every source byte and the aligned words around it inside the source row) are one thread
// each: five aligned 32-bit loads from the arbitrarily aligned source row, four funnel shifts, one 16-byte store. The few
// border chunks per row go through a separate small launch of the byte-remapping kernel, so no warp mixes the two paths.
__global__ void __launch_bounds__(256) orb_pyr_level0_interior_kernel(const uint8_t* __restrict__ images, int step, long long image_stride,
                                                                      uint8_t* __restrict__ pyr, int c_lo, int c_hi,
                                                                      const __grid_constant__ OrbGeom g) {
  orb_pdl_begin();
  const OrbLevel& L = g.lv[0];
  const int frame = blockIdx.z;
  const int c = c_lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (c > c_hi) return;
  // four rows per thread (y, y + 8, y + 16, y + 24), all twenty loads issued before the first store: the CTA lives long enough to amortise its launch
  const int x0 = 16 * c - 16;
  const int ylast = L.rows + ORB_APRON - 2;
  unsigned w[4][5], sh[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = min((int)(blockIdx.y * 32 + threadIdx.y + 8 * r) - (ORB_APRON - 1), ylast);
    const int sy = orb_reflect101(orb_reflect101(y, L.rows) - g.pad, g.height);
    const uint8_t* p = images + (long long)frame * image_stride + (long long)sy * step + (x0 - g.pad);
    sh[r] = 8u * (unsigned)((uintptr_t)p & 3);
    const unsigned* ap = reinterpret_cast<const unsigned*>((uintptr_t)p & ~(uintptr_t)3);
#pragma unroll
    for (int k = 0; k < 5; ++k) w[r][k] = __ldg(ap + k);
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = (int)(blockIdx.y * 32 + threadIdx.y + 8 * r) - (ORB_APRON - 1);
    if (y > ylast) break;
    uint8_t* dst = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off + (long long)y * L.pitch + x0;
    *reinterpret_cast<uint4*>(dst) = make_uint4(__funnelshift_r(w[r][0], w[r][1], sh[r]), __funnelshift_r(w[r][1], w[r][2], sh[r]),
                                                __funnelshift_r(w[r][2], w[r][3], sh[r]), __funnelshift_r(w[r][3], w[r][4], sh[r]));
  }
}

// Border chunks: thread = (row, border chunk e, word k of the chunk); chunk index = e < n_left ? e : c_hi + 1 + (e - n_left).
__global__ void __launch_bounds__(256) orb_pyr_level0_border_kernel(const uint8_t* __restrict__ images, int step, long long image_stride,
                                                                    uint8_t* __restrict__ pyr, int n_left, int c_hi, int n_edge,
                                                                    const __grid_constant__ OrbGeom g) {
  orb_pdl_begin();
  const OrbLevel& L = g.lv[0];
  const int frame = blockIdx.z;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int nrows = L.rows + 2 * (ORB_APRON - 1);
  if (i >= nrows * n_edge * 4) return;
  const int k = i & 3, e = (i >> 2) % n_edge, y = (i >> 2) / n_edge - (ORB_APRON - 1);
  const int c = e < n_left ? e : c_hi + 1 + (e - n_left);
  const int x4 = 16 * c - 16 + 4 * k;
  const int sy = orb_reflect101(orb_reflect101(y, L.rows) - g.pad, g.height);
  const uint8_t* srow = images + (long long)frame * image_stride + (long long)sy * step;
  uint32_t out = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    const int x = max(min(x4 + b, L.cols + ORB_APRON - 1), -ORB_APRON);   // columns outside the apron are never read
    out |= (uint32_t)__ldg(srow + orb_reflect101(orb_reflect101(x, L.cols) - g.pad, g.width)) << (8 * b);
  }
  uint8_t* dst = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off;
  *reinterpret_cast<uint32_t*>(dst + (long long)y * L.pitch + x4) = out;
}

// Generic path (any scale factor): one thread = 4 adjacent padded output pixels, byte gathers.
__global__ void __launch_bounds__(256) orb_pyr_resize_kernel(uint8_t* __restrict__ pyr, const ResizeTab* __restrict__ tabs,
                                                             int level, const __grid_constant__ OrbGeom g) {
  orb_pdl_begin();
  const OrbLevel& D = g.lv[level];
  const OrbLevel& S = g.lv[level - 1];
  const int frame = blockIdx.z;
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;   // table index; padded column x4 - ORB_APRON
  const int y = blockIdx.y * blockDim.y + threadIdx.y;           // table index; padded row y - (ORB_APRON - 1)
  if (y >= D.rows + 2 * (ORB_APRON - 1) || x4 >= D.xtab_len) return;
  uint8_t* base = pyr + (unsigned long long)frame * g.frame_stride;
  const uint8_t* src = base + S.plane_off;
  const ResizeTab* xt = tabs + D.tab_off;
  const ResizeTab ty = xt[D.xtab_len + y];
  const int cy0 = ty.coef & 0xffff, cy1 = ty.coef >> 16;
  const uint8_t* r0 = src + (long long)ty.ofs * S.pitch;
  const uint8_t* r1 = src + (long long)min(ty.ofs + 1, S.rows - 1) * S.pitch;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const ResizeTab tx = xt[x4 + k];
    const int c0 = tx.coef & 0xffff, c1 = tx.coef >> 16;
    const int sx1 = min(tx.ofs + 1, S.cols - 1);
    const int t0 = r0[tx.ofs] * c0 + r0[sx1] * c1;   // HResizeLinear: int32, 11-bit coefficients
    const int t1 = r1[tx.ofs] * c0 + r1[sx1] * c1;
    const uint32_t v = (uint32_t)((((cy0 * (t0 >> 4)) >> 16) + ((cy1 * (t1 >> 4)) >> 16) + 2) >> 2);  // VResizeLinear
    out |= (v & 0xffu) << (8 * k);
  }
  *reinterpret_cast<uint32_t*>(base + D.plane_off + (long long)(y - (ORB_APRON - 1)) * D.pitch + x4 - ORB_APRON) = out;
}

// Fast path (horizontal ratio <= 2, checked on the host): one thread = 4 adjacent padded output pixels x 4 rows.
// The 4 pixels read source bytes inside one 8-byte window [lo, lo+8) per source row: three aligned 32-bit loads,
// two funnel shifts to byte-align the window, one PRMT per pixel to pick its (s, s+1) byte pair and one DP2A for
// the horizontal lerp (c0*s0 + c1*s1 with the 11-bit coefficient pair packed as 16-bit lanes).
__global__ void __launch_bounds__(256) orb_pyr_resize4_kernel(uint8_t* __restrict__ pyr, const ResizeTab* __restrict__ tabs,
                                                              int level, const __grid_constant__ OrbGeom g) {
  orb_pdl_begin();
  const OrbLevel& D = g.lv[level];
  const OrbLevel& S = g.lv[level - 1];
  const int frame = blockIdx.z;
  const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
  if (x4 >= D.xtab_len) return;
  uint8_t* base = pyr + (unsigned long long)frame * g.frame_stride;
  const uint8_t* src = base + S.plane_off;
  uint8_t* dst = base + D.plane_off;
  const ResizeTab* xt = tabs + D.tab_off;
  const ResizeTab* yt = xt + D.xtab_len;
  const uint4 e01 = *reinterpret_cast<const uint4*>(xt + x4), e23 = *reinterpret_cast<const uint4*>(xt + x4 + 2);
  const int o0 = (int)e01.x, o1 = (int)e01.z, o2 = (int)e23.x, o3 = (int)e23.z;
  const int lo = min(min(o0, o1), min(o2, o3));
  const unsigned sh = 8u * (unsigned)(lo & 3);
  const int wbase = lo & ~3;
  const unsigned s0 = 17u * (unsigned)(o0 - lo) + 16u, s1 = 17u * (unsigned)(o1 - lo) + 16u, s2 = 17u * (unsigned)(o2 - lo) + 16u,
                 s3 = 17u * (unsigned)(o3 - lo) + 16u;   // PRMT selectors: bytes (r, r + 1) of the window
  // all source words of the thread's four rows are requested before the first one is used (rows past the end re-read the last row)
  const int nrows = D.rows + 2 * (ORB_APRON - 1);
  unsigned a0[4], a1[4], a2[4], b0[4], b1[4], b2[4], cy0[4], cy1[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = min((int)(blockIdx.y * 32 + threadIdx.y + 8 * r), nrows - 1);
    const ResizeTab ty = yt[y];
    cy0[r] = (ty.coef & 0xffffu) << 16; cy1[r] = ty.coef & 0xffff0000u;   // pre-shifted: (cy * t) >> 16 == umulhi(cy << 16, t)
    const unsigned* r0 = reinterpret_cast<const unsigned*>(src + (long long)ty.ofs * S.pitch + wbase);
    const unsigned* r1 = reinterpret_cast<const unsigned*>(src + (long long)min(ty.ofs + 1, S.rows - 1) * S.pitch + wbase);
    a0[r] = r0[0]; a1[r] = r0[1]; a2[r] = r0[2]; b0[r] = r1[0]; b1[r] = r1[1]; b2[r] = r1[2];
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = blockIdx.y * 32 + threadIdx.y + 8 * r;
    if (y >= nrows) break;
    const unsigned al = __funnelshift_r(a0[r], a1[r], sh), ah = __funnelshift_r(a1[r], a2[r], sh);
    const unsigned bl = __funnelshift_r(b0[r], b1[r], sh), bh = __funnelshift_r(b1[r], b2[r], sh);
    unsigned v[4];
#define ORB_PX(k, sel, coef)                                                                     \
    {                                                                                            \
      const unsigned t0 = __dp2a_lo(coef, __byte_perm(al, ah, sel), 0u);                         \
      const unsigned t1 = __dp2a_lo(coef, __byte_perm(bl, bh, sel), 0u);                         \
      v[k] = (__umulhi(cy0[r], t0 >> 4) + __umulhi(cy1[r], t1 >> 4) + 2u) >> 2;                   \
    }
    ORB_PX(0, s0, e01.y) ORB_PX(1, s1, e01.w) ORB_PX(2, s2, e23.y) ORB_PX(3, s3, e23.w)
#undef ORB_PX
    *reinterpret_cast<unsigned*>(dst + (long long)(y - (ORB_APRON - 1)) * D.pitch + x4 - ORB_APRON) =
        __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
  }
}

// Tiled, separable form of the same resize for large passes (TMA + shared memory), ORB_B200_PYR_TILED=1: bit-exact, but measured 15 % slower than
// the register-window kernel above (39 against 27 thread instructions per pixel: the shared-memory round trip and per-row addressing cost more
// than the 1.5x reuse of horizontal lerps saves), so it is the alternative, not the default. A 128 x 32 output tile (table coordinates: apron
// included) needs <= 192 x 52 bytes of the source level. Persistent CTAs loop over (frame, tile) items with a two-stage pipeline: ONE thread issues the
// cp.async.bulk.tensor box of the NEXT item while all threads work on the current one, so the box latency is hidden (a single-stage box per
// tile is not: profiles/r2_tma_ab.md).
//   H pass: every source row of the box is lerped ONCE to the tile's 128 output columns (the 4-pixel kernel above does it for both source rows of
//           every output row, i.e. 2 / 1.2 = 1.7 times per source row): thread = pixel quad x every 8th row; 8-byte window + PRMT + DP2A as
//           above; (t >> 4) is kept as u16 (<= 32640) in shared memory;
//   V pass: thread = pixel quad x every 8th output row: two LDS.64 fetch the quad's values of its two source rows, (cy * t) >> 16 by IMAD.HI,
//           + 2 >> 2, one 32-bit store. Identical integer arithmetic, hence identical bytes.
#define RS_THREADS 256
__global__ void __launch_bounds__(RS_THREADS) orb_pyr_resize_tiled_kernel(uint8_t* __restrict__ pyr, const ResizeTab* __restrict__ tabs,
                                                                         const int* __restrict__ rs_tiles, int level, int batch,
                                                                         const __grid_constant__ OrbGeom g,
                                                                         const __grid_constant__ OrbLevelMaps maps) {
  __shared__ __align__(128) uint8_t s_raw[2][ORB_RS_BW * ORB_RS_BH];
  __shared__ __align__(16) unsigned short s_h[ORB_RS_BH][ORB_RS_TW];
  __shared__ __align__(8) unsigned long long s_bar[2];
  const OrbLevel& D = g.lv[level];
  const OrbLevel& S = g.lv[level - 1];
  const int tid = threadIdx.x, q = tid & 31, grp = tid >> 5;
  const ResizeTab* xt = tabs + D.tab_off;
  const ResizeTab* yt = xt + D.xtab_len;
  const int* xlo_t = rs_tiles + D.rs_off;
  const int* ylo_t = xlo_t + D.rs_ntx;
  const int per_frame = D.rs_ntx * D.rs_nty;
  const int total = per_frame * batch;
  const int nrows = D.rows + 2 * (ORB_APRON - 1);
  const unsigned bar0 = orb_smem_u32(&s_bar[0]);
  if (tid == 0) { orb_mbar_init(bar0, 1); orb_mbar_init(bar0 + 8, 1); }
  __syncthreads();
  auto issue = [&](int item, int buf) {   // one thread: the source window of `item` into s_raw[buf]
    const int frame = item / per_frame, t = item - frame * per_frame;
    const int ty = t / D.rs_ntx, tx = t - ty * D.rs_ntx;
    orb_mbar_expect_tx(bar0 + 8 * buf, ORB_RS_BW * ORB_RS_BH);
    // tensor coordinates = source-level coordinates + (16-byte row lead, apron rows); bytes outside the allocation arrive as zeros (never selected)
    orb_tma_load_3d(orb_smem_u32(s_raw[buf]), &maps.m[level - 1], xlo_t[tx] + 16, ylo_t[ty] + ORB_APRON_ROWS, frame, bar0 + 8 * buf);
  };
  if (tid == 0 && (int)blockIdx.x < total) issue(blockIdx.x, 0);
  int it = 0;
  for (int item = blockIdx.x; item < total; item += gridDim.x, ++it) {
    const int buf = it & 1;
    // s_raw[buf ^ 1] was last read in the H pass of the previous iteration, which every thread left before that iteration's first barrier
    if (tid == 0 && item + (int)gridDim.x < total) issue(item + gridDim.x, buf ^ 1);
    const int frame = item / per_frame, t = item - frame * per_frame;
    const int ty = t / D.rs_ntx, tx = t - ty * D.rs_ntx;
    const int x4 = tx * ORB_RS_TW + 4 * q;                // first table column of this thread's quad
    const int xlo = xlo_t[tx], ylo = ylo_t[ty];
    const bool xact = x4 < D.xtab_len;
    // ---- H pass
    uint4 e01 = make_uint4(0, 0, 0, 0), e23 = make_uint4(0, 0, 0, 0);
    if (xact) { e01 = *reinterpret_cast<const uint4*>(xt + x4); e23 = *reinterpret_cast<const uint4*>(xt + x4 + 2); }
    const int o0 = (int)e01.x, o1 = (int)e01.z, o2 = (int)e23.x, o3 = (int)e23.z;
    const int lo = min(min(o0, o1), min(o2, o3));
    const unsigned sh = 8u * (unsigned)(lo & 3);
    const int wbase = (lo & ~3) - xlo;                     // box origin is 16-byte aligned: same word alignment as in global memory
    const unsigned s0 = 17u * (unsigned)(o0 - lo) + 16u, s1 = 17u * (unsigned)(o1 - lo) + 16u, s2 = 17u * (unsigned)(o2 - lo) + 16u,
                   s3 = 17u * (unsigned)(o3 - lo) + 16u;   // PRMT selectors: bytes (r, r + 1) of the window
    const int nsrc = ylo_t[D.rs_nty + ty];                // source rows this tile row needs (host table)
    orb_mbar_wait(bar0 + 8 * buf, (unsigned)(it >> 1) & 1u);
    if (xact) {
      const uint8_t* raw = s_raw[buf] + wbase;
      for (int r = grp; r < nsrc; r += RS_THREADS / 32) {
        const unsigned* w = reinterpret_cast<const unsigned*>(raw + r * ORB_RS_BW);
        const unsigned a0 = w[0], a1 = w[1], a2 = w[2];
        const unsigned al = __funnelshift_r(a0, a1, sh), ah = __funnelshift_r(a1, a2, sh);
        const unsigned t0 = __dp2a_lo(e01.y, __byte_perm(al, ah, s0), 0u) >> 4, t1 = __dp2a_lo(e01.w, __byte_perm(al, ah, s1), 0u) >> 4,
                       t2 = __dp2a_lo(e23.y, __byte_perm(al, ah, s2), 0u) >> 4, t3 = __dp2a_lo(e23.w, __byte_perm(al, ah, s3), 0u) >> 4;
        *reinterpret_cast<uint2*>(&s_h[r][4 * q]) = make_uint2(t0 | (t1 << 16), t2 | (t3 << 16));
      }
    }
    __syncthreads();
    // ---- V pass
    if (xact) {
      uint8_t* dst = pyr + (unsigned long long)frame * g.frame_stride + D.plane_off;
#pragma unroll
      for (int j = 0; j < ORB_RS_TH / (RS_THREADS / 32); ++j) {
        const int y = ty * ORB_RS_TH + grp + (RS_THREADS / 32) * j;
        if (y < nrows) {
          const ResizeTab tyy = yt[y];
          const unsigned cy0 = (tyy.coef & 0xffffu) << 16, cy1 = tyy.coef & 0xffff0000u;   // pre-shifted: (cy * t) >> 16 == umulhi(cy << 16, t)
          const int r0 = tyy.ofs - ylo, r1 = min(tyy.ofs + 1, S.rows - 1) - ylo;
          const uint2 a = *reinterpret_cast<const uint2*>(&s_h[r0][4 * q]), b = *reinterpret_cast<const uint2*>(&s_h[r1][4 * q]);
          const unsigned v0 = (__umulhi(cy0, a.x & 0xffffu) + __umulhi(cy1, b.x & 0xffffu) + 2u) >> 2;
          const unsigned v1 = (__umulhi(cy0, a.x >> 16) + __umulhi(cy1, b.x >> 16) + 2u) >> 2;
          const unsigned v2 = (__umulhi(cy0, a.y & 0xffffu) + __umulhi(cy1, b.y & 0xffffu) + 2u) >> 2;
          const unsigned v3 = (__umulhi(cy0, a.y >> 16) + __umulhi(cy1, b.y >> 16) + 2u) >> 2;
          *reinterpret_cast<unsigned*>(dst + (long long)(y - (ORB_APRON - 1)) * D.pitch + x4 - ORB_APRON) =
              __byte_perm(__byte_perm(v0, v1, 0x0040), __byte_perm(v2, v3, 0x0040), 0x5410);
        }
      }
    }
    __syncthreads();   // s_h and s_raw[buf] may be overwritten
  }
}

template <typename... KArgs, typename... Args>
static void launch_chain(void (*kern)(KArgs...), dim3 grd, dim3 blk, cudaStream_t st, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grd; cfg.blockDim = blk; cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

void orb_launch_pyramid(const uint8_t* images, int step, long long image_stride, uint8_t* pyr, const ResizeTab* tabs, const int* rs_tiles,
                        const void* rs_maps, int num_sms, const OrbGeom& g, int batch, cudaStream_t st, long long* launches,
                        cudaEvent_t ev_input_consumed) {
  static const bool tiled = getenv("ORB_B200_PYR_TILED") != nullptr;   // measured alternative (15 % slower, profiles/r2_tma_ab.md): off by default
  dim3 blk(32, 8);
  // small passes (a Frame's one or two images) are a chain of launch latencies: overlap each launch with the tail of the previous
  // kernel. Large passes keep ordinary launches: waiting CTAs would hold SM slots the other pass stream can use.
  static const bool pdl_on = !getenv("ORB_B200_NO_PDL");
  const bool pdl = pdl_on && batch <= 8;
  bool first = true;
  {
    const OrbLevel& L = g.lv[0];
    const int nrows = L.rows + 2 * (ORB_APRON - 1);
    const int c_last = (L.cols + ORB_APRON - 1 + 16) / 16;              // chunk holding the last apron column
    // interior chunk c: source columns 16c - 16 - pad .. 16c - 1 - pad and the aligned words around them (3 bytes before, 5 words from there) lie inside
    // the source row: 16c - 19 - pad >= 0 and 16c + 3 - pad <= W - 1 (pad 19: c = 3 .. (W + 15) / 16)
    const int c_lo = (19 + g.pad + 15) / 16;
    const int c_hi = std::min(c_last, (g.width + g.pad - 4) / 16);
    const int n_int = std::max(c_hi - c_lo + 1, 0);
    const int hi_eff = n_int > 0 ? c_hi : c_lo - 1;
    const int n_left = std::min(c_lo, c_last + 1), n_edge = n_left + (c_last - hi_eff);
    if (n_int > 0) {
      dim3 grd((n_int + 31) / 32, (nrows + 31) / 32, batch);
      launch_chain(orb_pyr_level0_interior_kernel, grd, blk, st, false, images, step, image_stride, pyr, c_lo, c_hi, g);
      first = false;
      ++*launches;
    }
    dim3 grd((nrows * n_edge * 4 + 255) / 256, 1, batch);
    launch_chain(orb_pyr_level0_border_kernel, grd, dim3(256), st, pdl && !first, images, step, image_stride, pyr, n_left, hi_eff, n_edge, g);
    ++*launches;
    // only level 0 reads the input frames (with dependent launches the record would break the kernel-to-kernel edge: it moves to the end)
    if (ev_input_consumed && !pdl) cudaEventRecord(ev_input_consumed, st);
  }
  for (int l = 1; l < g.nlevels; ++l) {
    const OrbLevel& L = g.lv[l];
    if (L.pyr_tiled && rs_maps && !pdl && tiled) {   // large passes: persistent CTAs, two-stage TMA pipeline, separable lerp through shared memory
      const int items = L.rs_ntx * L.rs_nty * batch;
      const int ctas = std::min(items, num_sms * 5);     // 28 KB of shared memory each
      orb_pyr_resize_tiled_kernel<<<ctas, RS_THREADS, 0, st>>>(pyr, tabs, rs_tiles, l, batch, g, *reinterpret_cast<const OrbLevelMaps*>(rs_maps));
    } else if (L.pyr_fast) {
      dim3 grd((L.xtab_len / 4 + 31) / 32, (L.rows + 2 * (ORB_APRON - 1) + 31) / 32, batch);
      launch_chain(orb_pyr_resize4_kernel, grd, blk, st, pdl, pyr, tabs, l, g);
    } else {
      dim3 grd((L.xtab_len / 4 + 31) / 32, (L.rows + 2 * (ORB_APRON - 1) + 7) / 8, batch);
      launch_chain(orb_pyr_resize_kernel, grd, blk, st, pdl, pyr, tabs, l, g);
    }
    ++*launches;
  }
  if (ev_input_consumed && pdl) cudaEventRecord(ev_input_consumed, st);
}
