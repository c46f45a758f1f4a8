// Pyramid kernels: ORBextractor::ComputePyramid (src/ORBextractor.cc:2093-2168) of THIS fork:
// level 0 = copyMakeBorder(image, 19, REFLECT_101); level l = resize(PADDED level l-1 -> w_l x h_l,
// INTER_LINEAR) + in-place copyMakeBorder(REFLECT_101|ISOLATED); mvImagePyramid[l] = padded image.
// Integer-only restatement of OpenCV's 8-bit fixed-point resize -> bit-exact by construction.
// Border pixels are pure index remaps of interior pixels, so every padded output pixel is computed
// directly from level l-1 (no second pass over the plane).
#include "orb_common.cuh"

// One thread = 4 horizontally adjacent padded output pixels (one 32-bit store).
__global__ void __launch_bounds__(256) orb_pyr_level0_kernel(const uint8_t* __restrict__ images, int step,
                                                             long long image_stride, uint8_t* __restrict__ pyr,
                                                             const __grid_constant__ OrbGeom g) {
  const OrbLevel& L = g.lv[0];
  const int frame = blockIdx.z;
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (y >= L.rows || x4 >= L.pitch) return;
  const uint8_t* src = images + (long long)frame * image_stride;
  const int sy = orb_reflect101(y - ORB_EDGE, g.height);
  const uint8_t* srow = src + (long long)sy * step;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int x = x4 + k;
    uint32_t v = 0;
    if (x < L.cols) v = __ldg(srow + orb_reflect101(x - ORB_EDGE, g.width));
    out |= v << (8 * k);
  }
  uint8_t* dst = pyr + (unsigned long long)frame * g.frame_stride + L.plane_off;
  *reinterpret_cast<uint32_t*>(dst + (long long)y * L.pitch + x4) = out;
}

__global__ void __launch_bounds__(256) orb_pyr_resize_kernel(uint8_t* __restrict__ pyr, const ResizeTab* __restrict__ tabs,
                                                             int level, const __grid_constant__ OrbGeom g) {
  const OrbLevel& D = g.lv[level];
  const OrbLevel& S = g.lv[level - 1];
  const int frame = blockIdx.z;
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (y >= D.rows || x4 >= D.pitch) return;
  uint8_t* base = pyr + (unsigned long long)frame * g.frame_stride;
  const uint8_t* src = base + S.plane_off;
  const ResizeTab* xt = tabs + D.tab_off;
  const ResizeTab* yt = xt + D.iw;
  const ResizeTab ty = yt[orb_reflect101(y - ORB_EDGE, D.ih)];
  const int sy1 = min(ty.ofs + 1, S.rows - 1);
  const uint8_t* r0 = src + (long long)ty.ofs * S.pitch;
  const uint8_t* r1 = src + (long long)sy1 * S.pitch;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int x = x4 + k;
    uint32_t v = 0;
    if (x < D.cols) {
      const ResizeTab tx = xt[orb_reflect101(x - ORB_EDGE, D.iw)];
      const int sx1 = min(tx.ofs + 1, S.cols - 1);
      const int t0 = r0[tx.ofs] * tx.c0 + r0[sx1] * tx.c1;   // HResizeLinear: int32, 11-bit coefficients
      const int t1 = r1[tx.ofs] * tx.c0 + r1[sx1] * tx.c1;
      v = (uint32_t)((((ty.c0 * (t0 >> 4)) >> 16) + ((ty.c1 * (t1 >> 4)) >> 16) + 2) >> 2);  // VResizeLinear
    }
    out |= (v & 0xffu) << (8 * k);
  }
  *reinterpret_cast<uint32_t*>(base + D.plane_off + (long long)y * D.pitch + x4) = out;
}

void orb_launch_pyramid(const uint8_t* images, int step, long long image_stride, uint8_t* pyr, const ResizeTab* tabs,
                        const OrbGeom& g, int batch, cudaStream_t st, long long* launches) {
  dim3 blk(32, 8);
  {
    const OrbLevel& L = g.lv[0];
    dim3 grd((L.pitch / 4 + 31) / 32, (L.rows + 7) / 8, batch);
    orb_pyr_level0_kernel<<<grd, blk, 0, st>>>(images, step, image_stride, pyr, g);
    ++*launches;
  }
  for (int l = 1; l < g.nlevels; ++l) {
    const OrbLevel& L = g.lv[l];
    dim3 grd((L.pitch / 4 + 31) / 32, (L.rows + 7) / 8, batch);
    orb_pyr_resize_kernel<<<grd, blk, 0, st>>>(pyr, tabs, l, g);
    ++*launches;
  }
}
