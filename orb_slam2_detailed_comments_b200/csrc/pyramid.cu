// Pyramid kernels: ORBextractor::ComputePyramid (src/ORBextractor.cc:2093-2168) of THIS fork:
// level 0 = copyMakeBorder(image, 19, REFLECT_101); level l = resize(PADDED level l-1 -> w_l x h_l,
// INTER_LINEAR) + in-place copyMakeBorder(REFLECT_101|ISOLATED); mvImagePyramid[l] = padded image.
// Integer-only restatement of OpenCV's 8-bit fixed-point resize -> bit-exact by construction.
// Border pixels are pure index remaps of interior pixels, so every padded output pixel is computed
// directly from level l-1 (no second pass over the plane).
#include "orb_common.cuh"

// Every plane is written together with its apron: columns [-4, cols + 4) and rows [-3, rows + 3) hold the
// BORDER_REFLECT_101 extension of the PADDED image (what cv::GaussianBlur reads beyond it), so the blur kernel is branch-free.
// Kernel x index t <-> padded column t - ORB_APRON, y index t <-> padded row t - (ORB_APRON - 1).

// Level 0 = copyMakeBorder(image, 19, REFLECT_101). One thread = 4 adjacent output pixels (one 32-bit store); interior quads
// read the (arbitrarily aligned) source row as two aligned words + a funnel shift, border quads as 4 remapped bytes.
__global__ void __launch_bounds__(256) orb_pyr_level0_kernel(const uint8_t* __restrict__ images, int step,
                                                             long long image_stride, uint8_t* __restrict__ pyr,
                                                             const __grid_constant__ OrbGeom g) {
  const OrbLevel& L = g.lv[0];
  const int frame = blockIdx.z;
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4 - ORB_APRON;
  const int y = blockIdx.y * blockDim.y + threadIdx.y - (ORB_APRON - 1);
  if (y >= L.rows + ORB_APRON - 1 || x4 >= L.cols + ORB_APRON) return;
  const uint8_t* src = images + (long long)frame * image_stride;
  const int sy = orb_reflect101(orb_reflect101(y, L.rows) - ORB_EDGE, g.height);
  const uint8_t* srow = src + (long long)sy * step;
  uint32_t out = 0;
  const int sx = x4 - ORB_EDGE;
  if (sx >= 4 && sx + 8 <= g.width) {   // both aligned words lie inside this source row
    const uint8_t* p = srow + sx;
    const unsigned sh = 8u * (unsigned)((uintptr_t)p & 3);
    const unsigned* ap = reinterpret_cast<const unsigned*>((uintptr_t)p & ~(uintptr_t)3);
    const unsigned lo = __ldg(ap);
    const unsigned hi = __ldg(ap + 1);
    out = __funnelshift_r(lo, hi, sh);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int x = min(x4 + k, L.cols + ORB_APRON - 1);
      out |= (uint32_t)__ldg(srow + orb_reflect101(orb_reflect101(x, L.cols) - ORB_EDGE, g.width)) << (8 * k);
    }
  }
  uint8_t* dst = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off;
  *reinterpret_cast<uint32_t*>(dst + (long long)y * L.pitch + x4) = out;
}

// Generic path (any scale factor): one thread = 4 adjacent padded output pixels, byte gathers.
__global__ void __launch_bounds__(256) orb_pyr_resize_kernel(uint8_t* __restrict__ pyr, const ResizeTab* __restrict__ tabs,
                                                             int level, const __grid_constant__ OrbGeom g) {
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
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = blockIdx.y * 32 + threadIdx.y + 8 * r;
    if (y >= D.rows + 2 * (ORB_APRON - 1)) break;
    const ResizeTab ty = yt[y];
    const unsigned cy0 = ty.coef & 0xffffu, cy1 = ty.coef >> 16;
    const unsigned* r0 = reinterpret_cast<const unsigned*>(src + (long long)ty.ofs * S.pitch + wbase);
    const unsigned* r1 = reinterpret_cast<const unsigned*>(src + (long long)min(ty.ofs + 1, S.rows - 1) * S.pitch + wbase);
    const unsigned a0 = r0[0], a1 = r0[1], a2 = r0[2], b0 = r1[0], b1 = r1[1], b2 = r1[2];
    const unsigned al = __funnelshift_r(a0, a1, sh), ah = __funnelshift_r(a1, a2, sh);
    const unsigned bl = __funnelshift_r(b0, b1, sh), bh = __funnelshift_r(b1, b2, sh);
    unsigned v[4];
#define ORB_PX(k, sel, coef)                                                                     \
    {                                                                                            \
      const unsigned t0 = __dp2a_lo(coef, __byte_perm(al, ah, sel), 0u);                         \
      const unsigned t1 = __dp2a_lo(coef, __byte_perm(bl, bh, sel), 0u);                         \
      v[k] = (((cy0 * (t0 >> 4)) >> 16) + ((cy1 * (t1 >> 4)) >> 16) + 2u) >> 2;                  \
    }
    ORB_PX(0, s0, e01.y) ORB_PX(1, s1, e01.w) ORB_PX(2, s2, e23.y) ORB_PX(3, s3, e23.w)
#undef ORB_PX
    *reinterpret_cast<unsigned*>(dst + (long long)(y - (ORB_APRON - 1)) * D.pitch + x4 - ORB_APRON) =
        __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
  }
}

void orb_launch_pyramid(const uint8_t* images, int step, long long image_stride, uint8_t* pyr, const ResizeTab* tabs,
                        const OrbGeom& g, int batch, cudaStream_t st, long long* launches) {
  dim3 blk(32, 8);
  {
    const OrbLevel& L = g.lv[0];
    dim3 grd(((L.cols + 2 * ORB_APRON + 3) / 4 + 31) / 32, (L.rows + 2 * (ORB_APRON - 1) + 7) / 8, batch);
    orb_pyr_level0_kernel<<<grd, blk, 0, st>>>(images, step, image_stride, pyr, g);
    ++*launches;
  }
  for (int l = 1; l < g.nlevels; ++l) {
    const OrbLevel& L = g.lv[l];
    if (L.pyr_fast) {
      dim3 grd((L.xtab_len / 4 + 31) / 32, (L.rows + 2 * (ORB_APRON - 1) + 31) / 32, batch);
      orb_pyr_resize4_kernel<<<grd, blk, 0, st>>>(pyr, tabs, l, g);
    } else {
      dim3 grd((L.xtab_len / 4 + 31) / 32, (L.rows + 2 * (ORB_APRON - 1) + 7) / 8, batch);
      orb_pyr_resize_kernel<<<grd, blk, 0, st>>>(pyr, tabs, l, g);
    }
    ++*launches;
  }
}
