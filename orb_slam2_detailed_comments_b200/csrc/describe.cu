// Gaussian blur (src/ORBextractor.cc:2039-2047), IC_Angle (:104-161), computeOrbDescriptor (:177-254) and the
// output assembly of ORBextractor::operator() (:2020-2081).
#include "orb_common.cuh"

__constant__ int8_t c_pattern[256 * 4] = {
#include "../../include/orb_pattern_31.inc"
};
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};  // :866-910

// ---- 7x7 sigma-2 Gaussian, OpenCV 8-bit fixed-point path (taps 8.8): out = (sum_ij k_i k_j p_ij + 32768) >> 16,
// BORDER_REFLECT_101 at the border of the PADDED level (this fork blurs the padded image). All intermediate sums are exact
// integers (row sums <= 257*255 fit 16 bits), so the order of the two passes is free. CTA tile = 128 x 32 output pixels:
//   H pass: one thread = 4 adjacent pixels of one of the 38 input rows, read straight from global (3 words), bytes widened
//           to 16-bit lane pairs [p(a), p(a+2)] so that one 32-bit IMAD does two taps -> u16 row sums in shared memory;
//   V pass: one thread = 4 pixels x 4 output rows from 10 staged rows, 32-bit accumulators, one 32-bit store per row.
#define BL_TW 128
#define BL_TH 32
template <bool CLAMP>   // CLAMP: tap sum > 256 (legacy taps sum to 257): saturate like OpenCV's ufixedpoint -> u8 cast
__global__ void __launch_bounds__(256) orb_blur_kernel(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blr,
                                                       const __grid_constant__ OrbGeom g) {
  __shared__ uint2 s_h[BL_TH + 6][BL_TW / 4];   // per pixel quad: .x = [h0, h2], .y = [h1, h3] (16-bit lanes)
  int level = 0, t = blockIdx.x;
  for (; level < g.nlevels; ++level) {
    const int tx = (g.lv[level].cols + BL_TW - 1) / BL_TW, ty = (g.lv[level].rows + BL_TH - 1) / BL_TH;
    if (t < tx * ty) break;
    t -= tx * ty;
  }
  if (level >= g.nlevels) return;
  const OrbLevel& L = g.lv[level];
  const int tiles_x = (L.cols + BL_TW - 1) / BL_TW;
  const int ty_ = t / tiles_x;
  const int x0 = (t - ty_ * tiles_x) * BL_TW, y0 = ty_ * BL_TH;
  const unsigned long long off = (unsigned long long)blockIdx.y * g.frame_stride + L.plane_off;
  const uint8_t* src = pyr + off;
  const unsigned k0 = g.blur_taps[0], k1 = g.blur_taps[1], k2 = g.blur_taps[2], k3 = g.blur_taps[3];
  for (int i = threadIdx.x; i < (BL_TH + 6) * (BL_TW / 4); i += 256) {
    const int ly = i >> 5, q = i & 31;
    const int x = x0 + 4 * q;
    const int sy = orb_reflect101(min(y0 + ly - 3, L.rows + 2), L.rows);
    const uint8_t* row = src + (long long)sy * L.pitch;
    unsigned c0, c1, c2;
    if (x >= 4 && x + 8 <= L.cols) {
      c0 = *reinterpret_cast<const unsigned*>(row + x - 4);
      c1 = *reinterpret_cast<const unsigned*>(row + x);
      c2 = *reinterpret_cast<const unsigned*>(row + x + 4);
    } else {
      c0 = c1 = c2 = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        c0 |= (unsigned)row[orb_reflect101(min(x - 4 + b, L.cols + 2), L.cols)] << (8 * b);
        c1 |= (unsigned)row[orb_reflect101(min(x + b, L.cols + 2), L.cols)] << (8 * b);
        c2 |= (unsigned)row[orb_reflect101(min(x + 4 + b, L.cols + 2), L.cols)] << (8 * b);
      }
    }
    // W(a) = bytes B[a..a+3] of the 12 loaded bytes B[0..11] = pixels x-4 .. x+7
    const unsigned w1 = __byte_perm(c0, c1, 0x4321), w3 = __byte_perm(c0, c1, 0x6543), w5 = __byte_perm(c1, c2, 0x4321),
                   w7 = __byte_perm(c1, c2, 0x6543);
    // P(a) = [B[a], B[a+2]] as 16-bit lanes
    const unsigned p1 = __byte_perm(w1, 0, 0x4240), p2 = __byte_perm(w1, 0, 0x4341), p3 = __byte_perm(w3, 0, 0x4240),
                   p4 = __byte_perm(w3, 0, 0x4341), p5 = __byte_perm(w5, 0, 0x4240), p6 = __byte_perm(w5, 0, 0x4341),
                   p7 = __byte_perm(w7, 0, 0x4240), p8 = __byte_perm(w7, 0, 0x4341);
    uint2 h;
    h.x = k0 * (p1 + p7) + k1 * (p2 + p6) + k2 * (p3 + p5) + k3 * p4;   // [h(x), h(x+2)]
    h.y = k0 * (p2 + p8) + k1 * (p3 + p7) + k2 * (p4 + p6) + k3 * p5;   // [h(x+1), h(x+3)]
    s_h[ly][q] = h;
  }
  __syncthreads();
  const int q = threadIdx.x & 31, seg = threadIdx.x >> 5;
  const int x = x0 + 4 * q;
  if (x >= L.cols) return;
  unsigned v[10][4];
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint2 h = s_h[4 * seg + r][q];
    v[r][0] = h.x & 0xffffu; v[r][2] = h.x >> 16; v[r][1] = h.y & 0xffffu; v[r][3] = h.y >> 16;
  }
  uint8_t* dst = blr + off;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = y0 + 4 * seg + r;
    if (y < L.rows) {
      unsigned c[4];
#pragma unroll
      for (int p = 0; p < 4; ++p)
        c[p] = k0 * (v[r][p] + v[r + 6][p]) + k1 * (v[r + 1][p] + v[r + 5][p]) + k2 * (v[r + 2][p] + v[r + 4][p]) + k3 * v[r + 3][p] + 32768u;
      if (CLAMP) {
#pragma unroll
        for (int p = 0; p < 4; ++p) c[p] = min(c[p], 0xffffffu);
      }
      // byte 2 of each accumulator is the output pixel
      const unsigned lo = __byte_perm(c[0], c[1], 0x0062), hi = __byte_perm(c[2], c[3], 0x0062);
      *reinterpret_cast<unsigned*>(dst + (long long)y * L.pitch + x) = __byte_perm(lo, hi, 0x5410);
    }
  }
}

// cv::fastAtan2 bit replica (float32, no FMA contraction).
__device__ __forceinline__ float orb_fast_atan2(float y, float x) {
  const float s = (float)(180.0 / 3.1415926535897932384626433832795);
  const float p1 = __fmul_rn(0.9997878412794807f, s), p3 = __fmul_rn(-0.3258083974640975f, s),
              p5 = __fmul_rn(0.1555786518463281f, s), p7 = __fmul_rn(-0.04432655554792128f, s);
  const float eps = (float)2.2204460492503131e-16;
  const float ax = fabsf(x), ay = fabsf(y);
  float a;
  if (ax >= ay) {
    const float c = __fdiv_rn(ay, __fadd_rn(ax, eps)), c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    const float c = __fdiv_rn(ax, __fadd_rn(ay, eps)), c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0.f) a = __fsub_rn(180.f, a);
  if (y < 0.f) a = __fsub_rn(360.f, a);
  return a;
}

// Canonical sin/cos of a float angle: fdlibm-style reduction + kernels in double with explicit
// round-to-nearest mul/add (no FMA), one rounding to float. Same operation sequence as the oracle.
__device__ __forceinline__ void orb_sincos(float angle, float* s, float* c) {
  const double x = (double)angle;
  const double invpio2 = 6.36619772367581382433e-01, pio2_1 = 1.57079632673412561417e+00, pio2_1t = 6.07710050650619224932e-11;
  const double kf = floor(__dadd_rn(__dmul_rn(x, invpio2), 0.5));
  const int k = (int)kf;
  const double r = __dsub_rn(__dsub_rn(x, __dmul_rn(kf, pio2_1)), __dmul_rn(kf, pio2_1t));
  const double z = __dmul_rn(r, r);
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03, S3 = -1.98412698298579493134e-04,
               S4 = 2.75573137070700676789e-06, S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03, C3 = 2.48015872894767294178e-05,
               C4 = -2.75573143513906633035e-07, C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
#define DM(a, b) __dmul_rn(a, b)
#define DA(a, b) __dadd_rn(a, b)
  const double ps = DA(S2, DM(z, DA(S3, DM(z, DA(S4, DM(z, DA(S5, DM(z, S6))))))));
  const double sn = DA(r, DM(DM(z, r), DA(S1, DM(z, ps))));
  const double pc = DM(z, DA(C1, DM(z, DA(C2, DM(z, DA(C3, DM(z, DA(C4, DM(z, DA(C5, DM(z, C6)))))))))));
  const double cs = __dsub_rn(1.0, __dsub_rn(DM(0.5, z), DM(z, pc)));
#undef DM
#undef DA
  switch (k & 3) {
    case 0: *s = (float)sn; *c = (float)cs; break;
    case 1: *s = (float)cs; *c = (float)(-sn); break;
    case 2: *s = (float)(-sn); *c = (float)(-cs); break;
    default: *s = (float)(-cs); *c = (float)sn; break;
  }
}

// One warp per kept keypoint: orientation on the un-blurred level, descriptor on the blurred level,
// then the final cv::KeyPoint (pt *= scale for level != 0) at the frame's output offset.
#define DESC_WARPS 8
__global__ void __launch_bounds__(DESC_WARPS * 32)
orb_describe_kernel(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blr, const uint32_t* __restrict__ kept,
                    const int* __restrict__ kept_cnt, OrbKeypoint* __restrict__ kps, uint8_t* __restrict__ desc, int cap,
                    int* __restrict__ counts, int* __restrict__ err_flag, const __grid_constant__ OrbGeom g) {
  __shared__ int8_t s_pat[1024];
  for (int i = threadIdx.x; i < 1024; i += DESC_WARPS * 32) s_pat[i] = c_pattern[i];
  __syncthreads();
  const int frame = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5);   // keypoint slot of this frame in output order
  const int* kc = kept_cnt + frame * g.nlevels;
  int level = -1, base = 0, total = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    const int c = kc[l];
    if (level < 0 && w < total + c) { level = l; base = total; }
    total += c;
  }
  if (w == 0 && lane == 0) {
    counts[frame] = min(total, cap);
    if (total > cap) atomicOr(err_flag, 8);
  }
  if (level < 0 || w >= cap) return;
  const OrbLevel& L = g.lv[level];
  const uint32_t q = kept[(long long)frame * g.node_total + L.node_off + (w - base)];
  const int x = (int)(q & 0xfffu) + ORB_MIN_BORDER, y = (int)((q >> 12) & 0xfffu) + ORB_MIN_BORDER;   // :1573-1589
  const unsigned long long off = (unsigned long long)frame * g.frame_stride + L.plane_off;
  // IC_Angle: lane u = lane-15 sums column u of the disc; m10 = sum u*I, m01 = sum v*I (exact int32)
  const uint8_t* center = pyr + off + (long long)y * L.pitch + x;
  int m10 = 0, m01 = 0;
  {
    const int u = lane - ORB_HALF_PATCH;
    if (lane < 31) {
#pragma unroll
      for (int v = -ORB_HALF_PATCH; v <= ORB_HALF_PATCH; ++v) {
        if (abs(u) <= c_umax[v < 0 ? -v : v]) {
          const int val = center[v * L.pitch + u];
          m10 += u * val;
          m01 += v * val;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
  }
  const float angle = orb_fast_atan2((float)m01, (float)m10);
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  float a, b;
  orb_sincos(__fmul_rn(angle, factorPI), &b, &a);
  const uint8_t* bc = blr + off + (long long)y * L.pitch + x;
  int val = 0;
#pragma unroll
  for (int bit = 0; bit < 8; ++bit) {
    const int8_t* p = &s_pat[(lane * 8 + bit) * 4];
    const float x0 = (float)p[0], y0 = (float)p[1], x1 = (float)p[2], y1 = (float)p[3];
    const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
    const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
    const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
    const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
    const int t0 = bc[r0 * L.pitch + c0], t1 = bc[r1 * L.pitch + c1];
    val |= (t0 < t1) << bit;
  }
  const long long o = (long long)frame * cap + w;
  desc[o * 32 + lane] = (uint8_t)val;
  if (lane == 0) {
    OrbKeypoint k;
    k.x = level ? __fmul_rn((float)x, L.scale) : (float)x;
    k.y = level ? __fmul_rn((float)y, L.scale) : (float)y;
    k.size = L.patch_size;
    k.angle = angle;
    k.response = (float)(q >> 24);
    k.octave = level;
    k.class_id = -1;
    kps[o] = k;
  }
}

void orb_launch_blur(const uint8_t* pyr, uint8_t* blr, const OrbGeom& g, int batch, cudaStream_t st, long long* launches) {
  int tiles = 0;
  for (int l = 0; l < g.nlevels; ++l)
    tiles += ((g.lv[l].cols + BL_TW - 1) / BL_TW) * ((g.lv[l].rows + BL_TH - 1) / BL_TH);
  int sum = g.blur_taps[3];
  for (int i = 0; i < 3; ++i) sum += 2 * g.blur_taps[i];
  if (sum > 256) orb_blur_kernel<true><<<dim3(tiles, batch), 256, 0, st>>>(pyr, blr, g);
  else orb_blur_kernel<false><<<dim3(tiles, batch), 256, 0, st>>>(pyr, blr, g);
  ++*launches;
}

void orb_launch_describe(const uint8_t* pyr, const uint8_t* blr, const uint32_t* kept, const int* kept_cnt, OrbKeypoint* kps,
                         uint8_t* desc, int cap, int* counts, int* err_flag, const OrbGeom& g, int batch, cudaStream_t st,
                         long long* launches) {
  orb_describe_kernel<<<dim3((cap + DESC_WARPS - 1) / DESC_WARPS, batch), DESC_WARPS * 32, 0, st>>>(pyr, blr, kept, kept_cnt, kps,
                                                                                                  desc, cap, counts, err_flag, g);
  ++*launches;
}
