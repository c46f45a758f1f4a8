// Device helpers shared by the descriptor kernels.
#pragma once
#include "orb_common.cuh"

static __device__ const int8_t d_pattern[256 * 4] = {
#include "../../include/orb_pattern_31.inc"
};
static __constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};  // :866-910

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


#define DESC_WARPS 4     // 128-thread CTAs
#define DESC_KPW 8
__device__ __forceinline__ int dp4a_u8_s8(unsigned a, unsigned b, int c) {   // a: unsigned bytes, b: signed bytes
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
