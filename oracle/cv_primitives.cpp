// TEST INFRASTRUCTURE ONLY — see cv_primitives.h. Compile with -ffp-contract=off.
#include "cv_primitives.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace orb_oracle {

int cv_round(float v) { return (int)lrintf(v); }
int cv_round(double v) { return (int)lrint(v); }

static inline int reflect101(int p, int n) {
  if (n == 1) return 0;
  while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
  return p;
}

// OpenCV resize, INTER_LINEAR, 8UC1: 11-bit coefficient tables (INTER_RESIZE_COEF_BITS=11),
// HResizeLinear to int32 then VResizeLinear's ((b0*(S0>>4))>>16 + (b1*(S1>>4))>>16 + 2)>>2.
static void linear_tab(int ssz, int dsz, std::vector<int>& ofs, std::vector<short>& coef) {
  const double scale = 1.0 / ((double)dsz / ssz);
  ofs.resize(dsz);
  coef.resize(2 * (size_t)dsz);
  for (int d = 0; d < dsz; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)std::floor(f);
    f -= s;
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= ssz - 1) { s = ssz - 1; f = 0.f; }
    ofs[d] = s;
    coef[2 * d] = (short)cv_round((1.f - f) * 2048.f);
    coef[2 * d + 1] = (short)cv_round(f * 2048.f);
  }
}

void resize_linear_u8(const Img& src, const Img& dst) {
  std::vector<int> xofs, yofs;
  std::vector<short> ca, cb;
  linear_tab(src.cols, dst.cols, xofs, ca);
  linear_tab(src.rows, dst.rows, yofs, cb);
  std::vector<int> r0(dst.cols), r1(dst.cols);
  auto hrow = [&](int sy, std::vector<int>& out) {
    const uint8_t* S = src.ptr(sy);
    for (int dx = 0; dx < dst.cols; ++dx) {
      const int sx = xofs[dx], sx1 = std::min(sx + 1, src.cols - 1);
      out[dx] = S[sx] * ca[2 * dx] + S[sx1] * ca[2 * dx + 1];
    }
  };
  int have0 = -1, have1 = -1;
  for (int dy = 0; dy < dst.rows; ++dy) {
    const int sy0 = yofs[dy], sy1 = std::min(sy0 + 1, src.rows - 1);
    if (have0 != sy0) {
      if (have1 == sy0) { std::swap(r0, r1); std::swap(have0, have1); }
      else { hrow(sy0, r0); have0 = sy0; }
    }
    if (have1 != sy1) { hrow(sy1, r1); have1 = sy1; }
    const int b0 = cb[2 * dy], b1 = cb[2 * dy + 1];
    uint8_t* D = dst.ptr(dy);
    for (int dx = 0; dx < dst.cols; ++dx)
      D[dx] = (uint8_t)((((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2);
  }
}

void copy_make_border_reflect101(const Img& src, const Img& dst, int b) {
  // Interior first (memmove: the reference's level>=1 call is in place on the ROI), then borders
  // as pure index remaps of the interior.
  for (int y = 0; y < src.rows; ++y) {
    uint8_t* d = dst.ptr(y + b) + b;
    if (d != src.ptr(y)) std::memmove(d, src.ptr(y), src.cols);
  }
  for (int y = 0; y < src.rows; ++y) {
    uint8_t* d = dst.ptr(y + b);
    for (int x = 0; x < b; ++x) {
      d[x] = d[b + reflect101(x - b, src.cols)];
      d[b + src.cols + x] = d[b + reflect101(src.cols + x, src.cols)];
    }
  }
  for (int y = 0; y < b; ++y) {
    std::memcpy(dst.ptr(y), dst.ptr(b + reflect101(y - b, src.rows)), dst.cols);
    std::memcpy(dst.ptr(b + src.rows + y), dst.ptr(b + reflect101(src.rows + y, src.rows)), dst.cols);
  }
}

void gaussian_blur_7x7_s2(const Img& src, const Img& dst, bool legacy_taps) {
  static const int k_new[7] = {18, 34, 48, 56, 48, 34, 18};
  static const int k_old[7] = {18, 34, 49, 55, 49, 34, 18};
  const int* k = legacy_taps ? k_old : k_new;
  const int W = src.cols, H = src.rows;
  // Row pass: u8 -> 8.8 fixed point (sum <= 65280 fits u16), reflect-101 via an extended row.
  std::vector<uint16_t> rowpass((size_t)W * H);
  std::vector<uint8_t> ext(W + 6);
  for (int y = 0; y < H; ++y) {
    const uint8_t* S = src.ptr(y);
    for (int x = 0; x < W + 6; ++x) ext[x] = S[reflect101(x - 3, W)];
    uint16_t* R = &rowpass[(size_t)y * W];
    const uint8_t* e = ext.data();
    for (int x = 0; x < W; ++x)
      R[x] = (uint16_t)(k[0] * e[x] + k[1] * e[x + 1] + k[2] * e[x + 2] + k[3] * e[x + 3] + k[4] * e[x + 4] +
                        k[5] * e[x + 5] + k[6] * e[x + 6]);
  }
  // Column pass: 8.8 x 8.8 -> 16.16 in u32, round to nearest: (c + 0x8000) >> 16.
  for (int y = 0; y < H; ++y) {
    const uint16_t* r[7];
    for (int t = 0; t < 7; ++t) r[t] = &rowpass[(size_t)reflect101(y + t - 3, H) * W];
    uint8_t* D = dst.ptr(y);
    for (int x = 0; x < W; ++x) {
      const uint32_t c = (uint32_t)k[0] * r[0][x] + (uint32_t)k[1] * r[1][x] + (uint32_t)k[2] * r[2][x] +
                         (uint32_t)k[3] * r[3][x] + (uint32_t)k[4] * r[4][x] + (uint32_t)k[5] * r[5][x] +
                         (uint32_t)k[6] * r[6][x];
      const uint32_t v = (c + 32768u) >> 16;
      D[x] = (uint8_t)std::min(v, 255u);
    }
  }
}

static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

int fast_corner_score(const uint8_t* p, int step) {
  int d[25];
  const int v = p[0];
  for (int k = 0; k < 16; ++k) d[k] = v - p[kRingDy[k] * step + kRingDx[k]];
  for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
  int A = -255, B = -255;  // A: best dark arc (centre brighter), B: best bright arc
  for (int k = 0; k < 16; ++k) {
    int mn = d[k], mx = d[k];
    for (int j = 1; j < 9; ++j) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
    A = std::max(A, mn);
    B = std::max(B, -mx);
  }
  const int s = std::max(A, B) - 1;
  return s < 0 ? 0 : s;
}

void fast_detect(const Img& img, int threshold, bool nonmax, std::vector<KeyPoint>& out) {
  out.clear();
  const int W = img.cols, H = img.rows;
  if (W < 7 || H < 7) return;
  std::vector<uint8_t> S((size_t)W * H, 0);  // thresholded score, 0 outside the detection domain
  for (int y = 3; y < H - 3; ++y)
    for (int x = 3; x < W - 3; ++x) {
      // Quick reject (same idea as cv::FAST's opposite-pixel test): every 9-arc of the 16-ring
      // holds at least one pixel of each opposite pair (k, k+8), so a corner at `threshold`
      // needs one of each pair beyond the threshold on the same side.
      const uint8_t* p = img.ptr(y) + x;
      const int v = p[0];
      bool dark = true, bright = true;
      for (int k = 0; k < 8 && (dark || bright); ++k) {
        const int d0 = v - p[kRingDy[k] * img.step + kRingDx[k]];
        const int d1 = v - p[kRingDy[k + 8] * img.step + kRingDx[k + 8]];
        dark = dark && (d0 > threshold || d1 > threshold);
        bright = bright && (d0 < -threshold || d1 < -threshold);
      }
      if (!dark && !bright) continue;
      const int s = fast_corner_score(p, img.step);
      if (s >= threshold) S[(size_t)y * W + x] = (uint8_t)s;
    }
  for (int y = 3; y < H - 3; ++y)
    for (int x = 3; x < W - 3; ++x) {
      const int s = S[(size_t)y * W + x];
      if (!s) continue;
      if (nonmax) {
        const uint8_t* c = &S[(size_t)y * W + x];
        if (!(s > c[-1] && s > c[1] && s > c[-W - 1] && s > c[-W] && s > c[-W + 1] && s > c[W - 1] &&
              s > c[W] && s > c[W + 1]))
          continue;
      }
      out.push_back(KeyPoint{(float)x, (float)y, 7.f, -1.f, nonmax ? (float)s : 0.f, 0, -1});
    }
}

float fast_atan2(float y, float x) {
  const float s = (float)(180.0 / 3.1415926535897932384626433832795);
  const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s, p5 = 0.1555786518463281f * s,
              p7 = -0.04432655554792128f * s;
  const float ax = std::fabs(x), ay = std::fabs(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

}  // namespace orb_oracle
