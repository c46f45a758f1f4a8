// Gaussian blur (src/ORBextractor.cc:2039-2047), IC_Angle (:104-161), computeOrbDescriptor (:177-254) and the
// output assembly of ORBextractor::operator() (:2020-2081).
#include "orb_common.cuh"

__device__ const int8_t d_pattern[256 * 4] = {
#include "../../include/orb_pattern_31.inc"
};
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};  // :866-910

// ---- 7x7 sigma-2 Gaussian, OpenCV 8-bit fixed-point path (taps 8.8): out = (sum_ij k_i k_j p_ij + 32768) >> 16,
// BORDER_REFLECT_101 at the border of the PADDED level (this fork blurs the padded image; the pyramid kernels have already
// written that reflection into the plane's apron, so no index is remapped here). All intermediate sums are exact integers
// (row sums <= 257*255 fit 16 bits), so the order of the two passes is free. CTA tile = 128 x 32 output pixels, 320 threads:
//   H pass: one thread = 4 adjacent pixels of two adjacent rows: three 32-bit loads per row, the seven taps of a pixel are
//           two DP4A on byte windows (PRMT); the two rows' sums are stored as vertical 16-bit pairs [h(r), h(r+1)], r even;
//   V pass: one thread = 4 pixels x 4 output rows from 6 staged row pairs; the seven taps are four DP2A on the pairs
//           (tap pairs differ for even and odd output rows), accumulator preloaded with the rounding constant.
#define BL_TW 128
#define BL_TH 32
#define BL_THREADS 320
template <bool CLAMP>   // CLAMP: tap sum > 256 (legacy taps sum to 257): saturate like OpenCV's ufixedpoint -> u8 cast
__global__ void __launch_bounds__(BL_THREADS) orb_blur_kernel(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blr,
                                                              const __grid_constant__ OrbGeom g) {
  __shared__ uint4 s_p[(BL_TH + 8) / 2][BL_TW / 4];   // [row pair][pixel quad]: word j = [h(2k, x+j), h(2k+1, x+j)]
  const int level = orb_level_of(g.blur_base, (int)blockIdx.x);
  const OrbLevel& L = g.lv[level];
  const unsigned t = blockIdx.x - L.blur_tile_base;
  const unsigned ty_ = (t * L.blur_inv_tx) >> 20;            // t / blur_tiles_x (host-checked range)
  const int x0 = (int)(t - ty_ * L.blur_tiles_x) * BL_TW, y0 = (int)ty_ * BL_TH;
  const unsigned long long off = (unsigned long long)blockIdx.y * g.frame_stride + L.plane_off;
  const uint8_t* src = pyr + off;
  const unsigned tapA = g.blur_k[0], tapB = g.blur_k[1];
  const int pitch = L.pitch;
  // local row r <-> plane row y0 - 4 + r (row -4 and rows past rows + 3 only ever meet a zero tap)
  // all 12 loads of this thread's two H-pass items (item = 4 pixels x 2 rows) are issued before any arithmetic
  unsigned cw[2][2][3];
  bool act[2];
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int i = threadIdx.x + it * BL_THREADS;   // 20 row pairs x 32 quads = 640 items
    const int k = i >> 5, q = i & 31;
    const int x = x0 + 4 * q;
    act[it] = x < L.cols;
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const int ry = min(y0 - 4 + 2 * k + rr, L.rows + 3);
      const unsigned* row = reinterpret_cast<const unsigned*>(src + (long long)ry * pitch + x);
      if (act[it]) { cw[it][rr][0] = __ldg(row - 1); cw[it][rr][1] = __ldg(row); cw[it][rr][2] = __ldg(row + 1); }
    }
  }
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int i = threadIdx.x + it * BL_THREADS;
    const int k = i >> 5, q = i & 31;
    if (act[it]) {
      unsigned h[2][4];
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const unsigned c0 = cw[it][rr][0], c1 = cw[it][rr][1], c2 = cw[it][rr][2];
        // W(a) = bytes B[a..a+3] of the 12 loaded bytes B[0..11] = pixels x-4 .. x+7; pixel j: W(j+1).tapA + W(j+5).tapB
        const unsigned w1 = __byte_perm(c0, c1, 0x4321), w2 = __byte_perm(c0, c1, 0x5432), w3 = __byte_perm(c0, c1, 0x6543),
                       w5 = __byte_perm(c1, c2, 0x4321), w6 = __byte_perm(c1, c2, 0x5432), w7 = __byte_perm(c1, c2, 0x6543);
        h[rr][0] = __dp4a(w1, tapA, __dp4a(w5, tapB, 0u));
        h[rr][1] = __dp4a(w2, tapA, __dp4a(w6, tapB, 0u));
        h[rr][2] = __dp4a(w3, tapA, __dp4a(w7, tapB, 0u));
        h[rr][3] = __dp4a(c1, tapA, __dp4a(c2, tapB, 0u));
      }
      s_p[k][q] = make_uint4(__byte_perm(h[0][0], h[1][0], 0x5410), __byte_perm(h[0][1], h[1][1], 0x5410),
                             __byte_perm(h[0][2], h[1][2], 0x5410), __byte_perm(h[0][3], h[1][3], 0x5410));
    }
  }
  __syncthreads();
  if (threadIdx.x >= 256) return;
  const int q = threadIdx.x & 31, seg = threadIdx.x >> 5;   // output rows y0 + 4*seg .. + 3 = local rows 4*seg + 4 .. + 7
  const int x = x0 + 4 * q;
  if (x >= L.cols) return;
  uint4 P[6];                                                // pairs 2*seg .. 2*seg + 5 = local rows 4*seg .. 4*seg + 11
#pragma unroll
  for (int i = 0; i < 6; ++i) P[i] = s_p[2 * seg + i][q];
  // even local row t = 2m: [0,k0].P(m-2) + [k1,k2].P(m-1) + [k3,k2].P(m) + [k1,k0].P(m+1)
  // odd  local row t = 2m+1: [k0,k1].P(m-1) + [k2,k3].P(m) + [k2,k1].P(m+1) + [k0,0].P(m+2)
  const unsigned e0 = g.blur_k[2], e1 = g.blur_k[3], e2 = g.blur_k[4], e3 = g.blur_k[5];
  const unsigned o0 = g.blur_k[6], o1 = g.blur_k[7], o2 = g.blur_k[8], o3 = g.blur_k[9];
  uint8_t* dst = blr + off + (long long)(y0 + 4 * seg) * pitch + x;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int y = y0 + 4 * seg + r;
    if (y < L.rows) {
      const int m = 2 + (r >> 1);                           // pair index (relative to P[0]) holding local row t
      unsigned c[4];
#pragma unroll
      for (int p = 0; p < 4; ++p) {
#define PW(i) (p == 0 ? P[i].x : p == 1 ? P[i].y : p == 2 ? P[i].z : P[i].w)
        if ((r & 1) == 0)
          c[p] = __dp2a_lo(PW(m - 2), e0, __dp2a_lo(PW(m - 1), e1, __dp2a_lo(PW(m), e2, __dp2a_lo(PW(m + 1), e3, 32768u))));
        else
          c[p] = __dp2a_lo(PW(m - 1), o0, __dp2a_lo(PW(m), o1, __dp2a_lo(PW(m + 1), o2, __dp2a_lo(PW(m + 2), o3, 32768u))));
#undef PW
        if (CLAMP) c[p] = min(c[p], 0xffffffu);
      }
      // byte 2 of each accumulator is the output pixel
      const unsigned lo = __byte_perm(c[0], c[1], 0x0062), hi = __byte_perm(c[2], c[3], 0x0062);
      *reinterpret_cast<unsigned*>(dst + r * pitch) = __byte_perm(lo, hi, 0x5410);
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

// Orientation (un-blurred level), descriptor (blurred level) and the final cv::KeyPoint (pt *= scale for level != 0).
// One warp handles DESC_KPW keypoints of one frame back to back, so the per-lane tables stay in registers:
//   * IC_Angle: lane = patch row v (31 rows). The row's 32 bytes u = -16..15 are 9 aligned words (staged through shared
//     memory by coalesced loads) funnel-shifted into 8 words; m10 += dp4a(word, [u masked by the disc]), rowsum += dp4a(word, [disc mask]); m01 = v * rowsum;
//     two REDUX adds finish the moments (exact int32).
//   * rBRIEF: lane = descriptor byte; its 8 test pairs (x0, y0, x1, y1 as floats) come from shared memory, one LDS.128 each;
//     the 512 pixel gathers hit the staged 37 x 40-byte patch in shared memory.
#define DESC_WARPS 4     // 128-thread CTAs, 14.5 KB of shared memory: one fits beside five resident FAST CTAs of the other pass stream
#define DESC_KPW 8
__device__ __forceinline__ int dp4a_u8_s8(unsigned a, unsigned b, int c) {   // a: unsigned bytes, b: signed bytes
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__global__ void __launch_bounds__(DESC_WARPS * 32, 40 / DESC_WARPS)
orb_describe_kernel(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blr, const uint32_t* __restrict__ kept,
                    const int* __restrict__ kept_cnt, OrbKeypoint* __restrict__ kps, uint8_t* __restrict__ desc, int cap,
                    int* __restrict__ counts, int* __restrict__ err_flag, const __grid_constant__ OrbGeom g) {
  const int frame = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // cumulative kept counts: lane l holds the inclusive prefix of level l
  int cinc = lane < g.nlevels ? kept_cnt[frame * g.nlevels + lane] : 0;
  const int cown = cinc;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, cinc, o);
    if (lane >= o) cinc += y;
  }
  const int total = __shfl_sync(0xffffffffu, cinc, g.nlevels - 1);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    counts[frame] = min(total, cap);
    if (total > cap) atomicOr(err_flag, 8);
  }
  const int w_first = (blockIdx.x * DESC_WARPS + warp) * DESC_KPW;
  const int w_end = min(min(w_first + DESC_KPW, total), cap);
  // rBRIEF pattern as floats, [test bit][lane] so that one LDS.128 per test is conflict-free (256 test pairs)
  __shared__ float4 s_pat[8 * 32];
#pragma unroll
  for (int b = warp; b < 8; b += DESC_WARPS) {
    const char4 p = reinterpret_cast<const char4*>(d_pattern)[lane * 8 + b];
    s_pat[b * 32 + lane] = make_float4((float)p.x, (float)p.y, (float)p.z, (float)p.w);
  }
  __syncthreads();
  if (w_first >= w_end) return;

  // per-lane constants
  const int v = lane - ORB_HALF_PATCH;                   // patch row of this lane (lane 31: none)
  unsigned dm[8];                                        // byte mask of the disc in this lane's row, words u = -16 + 4i .. + 3
  {
    const int hw = lane < 31 ? c_umax[v < 0 ? -v : v] : -1;   // half width of the disc in this row (:866-910, symmetric)
    // one bit per byte u = -16..15: bits [16 - hw, 16 + hw]; each nibble is then widened to a byte mask
    const unsigned bits = hw >= 0 ? ((2u << (2 * hw)) - 1u) << (16 - hw) : 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) dm[i] = ((((bits >> (4 * i)) & 15u) * 0x00204081u) & 0x01010101u) * 0xffu;
  }
  // per-warp staging: the 31 x 36-byte IC_Angle window of the un-blurred level and the 37 x 40-byte descriptor patch of the
  // blurred level are fetched with coalesced 32-bit loads (row-major over the lanes) and then gathered from shared memory
  __shared__ unsigned s_ic[DESC_WARPS][31 * 9 + 1];
  __shared__ unsigned s_pt[DESC_WARPS][37 * 10 + 2];
  unsigned* ic = s_ic[warp];
  unsigned* pt = s_pt[warp];
  const uint8_t* ptb = reinterpret_cast<const uint8_t*>(pt);
  const int ic_sub = lane / 9, ic_w = lane - 9 * ic_sub, ic_idx = 9 * ic_sub + ic_w;      // lanes 27..31 idle
  const int pt_sub = lane / 10, pt_w = lane - 10 * pt_sub, pt_idx = 10 * pt_sub + pt_w;   // lanes 30, 31 idle
  const bool ic_act = lane < 27, pt_act = lane < 30;

  for (int w = w_first; w < w_end; ++w) {
    const int level = __popc(__ballot_sync(0xffffffffu, lane < g.nlevels && cinc <= w));
    const int base = __shfl_sync(0xffffffffu, cinc - cown, level);
    const OrbLevel& L = g.lv[level];
    const int pitch = L.pitch;
    const uint32_t q = kept[(long long)frame * g.node_total + L.node_off + (w - base)];
    const int x = (int)(q & 0xfffu) + ORB_MIN_BORDER, y = (int)((q >> 12) & 0xfffu) + ORB_MIN_BORDER;   // :1573-1589
    const unsigned long long off = (unsigned long long)frame * g.frame_stride + L.plane_off;
    // all global loads of this keypoint are issued up front
    const int xi = (x - 16) & ~3, xp = (x - 18) & ~3;          // planes are 16-byte aligned: column alignment == address alignment
    const unsigned* icg = reinterpret_cast<const unsigned*>(pyr + off + (long long)(y - 15) * pitch + xi);
    const unsigned* ptg = reinterpret_cast<const unsigned*>(blr + off + (long long)(y - 18) * pitch + xp);
    const int pw = pitch >> 2;
    // three window rows per load instruction: lane = (row % 3, word); the per-lane offsets need one IMAD per keypoint
    const unsigned* icl = icg + ic_sub * pw + ic_w;
    const unsigned* ptl = ptg + pt_sub * pw + pt_w;
    // global -> shared without a register round trip: one cp.async per window word, the IC_Angle window and the descriptor patch as
    // two groups so that the orientation starts as soon as the first has landed
    __syncwarp();                                              // the previous keypoint's gathers are done
#pragma unroll
    for (int i = 0; i < 11; ++i)
      if (ic_act && 3 * i + ic_sub < 31) cp_async4(&ic[27 * i + ic_idx], icl + 3 * i * pw);
    asm volatile("cp.async.commit_group;");
#pragma unroll
    for (int i = 0; i < 13; ++i)
      if (pt_act && 3 * i + pt_sub < 37) cp_async4(&pt[30 * i + pt_idx], ptl + 3 * i * pw);
    asm volatile("cp.async.commit_group;");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    // IC_Angle (:104-161): lane = row v
    int m10 = 0, rs = 0;
    if (lane < 31) {
      const unsigned sh = 8u * (unsigned)((x - 16) & 3);
      unsigned a[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) a[i] = ic[lane * 9 + i];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const unsigned wv = __funnelshift_r(a[i], a[i + 1], sh) & dm[i];
        const int u0 = -16 + 4 * i;                       // signed byte coefficients u0 .. u0 + 3 (compile-time constant)
        const unsigned cu = ((unsigned)u0 & 0xffu) | (((unsigned)(u0 + 1) & 0xffu) << 8) | (((unsigned)(u0 + 2) & 0xffu) << 16) |
                            (((unsigned)(u0 + 3) & 0xffu) << 24);
        m10 = dp4a_u8_s8(wv, cu, m10);
        rs = (int)__dp4a(wv, 0x01010101u, (unsigned)rs);
      }
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    const int m01 = __reduce_add_sync(0xffffffffu, v * rs);
    const float angle = orb_fast_atan2((float)m01, (float)m10);
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    float a, b;
    orb_sincos(__fmul_rn(angle, factorPI), &b, &a);
    // computeOrbDescriptor (:177-254) on the staged patch: byte (r, c) of the patch is pt[(r + 18) * 40 + c + x - xp]
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    const uint8_t* bc = ptb + 18 * 40 + (x - xp);
    int val = 0;
#pragma unroll
    for (int bit = 0; bit < 8; ++bit) {
      const float4 p = s_pat[bit * 32 + lane];   // x0, y0, x1, y1
      const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)));
      const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)));
      const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)));
      const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)));
      const int t0 = bc[r0 * 40 + c0], t1 = bc[r1 * 40 + c1];
      val |= (t0 < t1) << bit;
    }
    const long long o = (long long)frame * cap + w;
    desc[o * 32 + lane] = (uint8_t)val;
    if (lane < 7) {   // the 28-byte cv::KeyPoint as seven coalesced 32-bit stores
      float f;
      switch (lane) {
        case 0: f = level ? __fmul_rn((float)x, L.scale) : (float)x; break;
        case 1: f = level ? __fmul_rn((float)y, L.scale) : (float)y; break;
        case 2: f = L.patch_size; break;
        case 3: f = angle; break;
        case 4: f = (float)(q >> 24); break;
        case 5: f = __int_as_float(level); break;
        default: f = __int_as_float(-1); break;
      }
      reinterpret_cast<float*>(kps + o)[lane] = f;
    }
  }
}

void orb_launch_blur(const uint8_t* pyr, uint8_t* blr, const OrbGeom& g, int batch, cudaStream_t st, long long* launches) {
  int sum = g.blur_taps[3];
  for (int i = 0; i < 3; ++i) sum += 2 * g.blur_taps[i];
  if (sum > 256) orb_blur_kernel<true><<<dim3(g.blur_tiles_total, batch), BL_THREADS, 0, st>>>(pyr, blr, g);
  else orb_blur_kernel<false><<<dim3(g.blur_tiles_total, batch), BL_THREADS, 0, st>>>(pyr, blr, g);
  ++*launches;
}

void orb_launch_describe(const uint8_t* pyr, const uint8_t* blr, const uint32_t* kept, const int* kept_cnt, OrbKeypoint* kps,
                         uint8_t* desc, int cap, int* counts, int* err_flag, const OrbGeom& g, int batch, cudaStream_t st,
                         long long* launches) {
  const int per_cta = DESC_WARPS * DESC_KPW;
  orb_describe_kernel<<<dim3((cap + per_cta - 1) / per_cta, batch), DESC_WARPS * 32, 0, st>>>(pyr, blr, kept, kept_cnt, kps,
                                                                                                  desc, cap, counts, err_flag, g);
  ++*launches;
}
