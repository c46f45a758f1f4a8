// TEST INFRASTRUCTURE ONLY — see orb_oracle.h. Compile with -ffp-contract=off (canonical: no FMA).
#include "orb_oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <list>
#include <map>
#include <thread>
#include <utility>
#include <vector>

#include "cv_primitives.h"

namespace orb_oracle {

static const int kPatch = 31, kHalfPatch = 15, kEdge = 19;  // src/ORBextractor.cc:80-82
static const int8_t kPattern[256 * 4] = {
#include "../include/orb_pattern_31.inc"
};

// Canonical sin/cos: fdlibm-style reduction + kernels in double, plain * + - only, one rounding
// to float. Replaces libm cosf/sinf of src/ORBextractor.cc:184-186 (see header).
static void sincos_canonical(float angle, float* s, float* c) {
  const double x = (double)angle;
  const double invpio2 = 6.36619772367581382433e-01, pio2_1 = 1.57079632673412561417e+00,
               pio2_1t = 6.07710050650619224932e-11;
  const double kf = std::floor(x * invpio2 + 0.5);
  const int k = (int)kf;
  const double r = (x - kf * pio2_1) - kf * pio2_1t;
  const double z = r * r;
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03, S3 = -1.98412698298579493134e-04,
               S4 = 2.75573137070700676789e-06, S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03, C3 = 2.48015872894767294178e-05,
               C4 = -2.75573143513906633035e-07, C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  const double ps = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
  const double sn = r + (z * r) * (S1 + z * ps);
  const double pc = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
  const double cs = 1.0 - (0.5 * z - z * pc);
  switch (k & 3) {
    case 0: *s = (float)sn; *c = (float)cs; break;
    case 1: *s = (float)cs; *c = (float)(-sn); break;
    case 2: *s = (float)(-sn); *c = (float)(-cs); break;
    default: *s = (float)(-cs); *c = (float)sn; break;
  }
}

struct Plane {
  std::vector<uint8_t> buf;
  int cols = 0, rows = 0;
  void alloc(int c, int r) { cols = c; rows = r; buf.assign((size_t)c * r, 0); }
  Img view() { return Img{buf.data(), cols, rows, cols}; }
  Img view() const { return Img{const_cast<uint8_t*>(buf.data()), cols, rows, cols}; }
};

class Extractor {
 public:
  Extractor(int nf, float sf, int nl, int ini, int mn, int flags);
  int run(const Img& image, std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc);

  int nfeatures, nlevels, iniTh, minTh, flags;
  double scaleFactor;  // the reference member is double (include/ORBextractor.h:234)
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> quota, umax;
  std::vector<Plane> pyr, blurred;
  std::vector<char> has_blur;
  std::vector<std::vector<KeyPoint>> raw;  // vToDistributeKeys per level
  std::vector<int> levelCounts;

 private:
  void computePyramid(const Img& image);
  void computeKeyPoints(std::vector<std::vector<KeyPoint>>& all);
};

std::vector<KeyPoint> distribute(const std::vector<KeyPoint>& in, int minX, int maxX, int minY, int maxY, int N,
                                 int reserveHint);

// src/ORBextractor.cc:776-925
Extractor::Extractor(int nf, float sf, int nl, int ini, int mn, int fl)
    : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), flags(fl), scaleFactor(sf) {
  scale.resize(nl); invScale.resize(nl); sigma2.resize(nl); invSigma2.resize(nl);
  scale[0] = 1.f; sigma2[0] = 1.f;
  for (int i = 1; i < nl; ++i) {
    scale[i] = (float)(scale[i - 1] * scaleFactor);  // float * double -> double, stored as float
    sigma2[i] = scale[i] * scale[i];
  }
  for (int i = 0; i < nl; ++i) { invScale[i] = 1.f / scale[i]; invSigma2[i] = 1.f / sigma2[i]; }
  quota.resize(nl);
  const float factor = (float)(1.0f / scaleFactor);
  float perScale = (float)(nf * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl)));
  int sum = 0;
  for (int l = 0; l < nl - 1; ++l) {
    quota[l] = cv_round(perScale);
    sum += quota[l];
    perScale *= factor;
  }
  quota[nl - 1] = std::max(nf - sum, 0);
  // umax: quarter-circle row half-widths, made symmetric (:866-910)
  umax.assign(kHalfPatch + 1, 0);
  const int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
  const double hp2 = kHalfPatch * kHalfPatch;
  for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
  pyr.resize(nl); blurred.resize(nl); has_blur.assign(nl, 0); raw.resize(nl); levelCounts.assign(nl, 0);
}

// src/ORBextractor.cc:2093-2168. Default = this fork: level l is resized from the PADDED level l-1
// and mvImagePyramid[l] is the padded image (line 2166). Upstream flag: ROI semantics.
void Extractor::computePyramid(const Img& image) {
  const bool upstream = flags & ORB_ORACLE_UPSTREAM_PYRAMID;
  for (int l = 0; l < nlevels; ++l) {
    const float s = invScale[l];
    const int w = cv_round((float)image.cols * s), h = cv_round((float)image.rows * s);
    pyr[l].alloc(w + 2 * kEdge, h + 2 * kEdge);
    Img whole = pyr[l].view();
    Img inner = whole.roi(kEdge, kEdge, w, h);
    if (l == 0) {
      copy_make_border_reflect101(image, whole, kEdge);
    } else {
      Img prev = pyr[l - 1].view();
      if (upstream) prev = prev.roi(kEdge, kEdge, prev.cols - 2 * kEdge, prev.rows - 2 * kEdge);
      resize_linear_u8(prev, inner);
      copy_make_border_reflect101(inner, whole, kEdge);
    }
  }
}

// src/ORBextractor.cc:1424-1601
void Extractor::computeKeyPoints(std::vector<std::vector<KeyPoint>>& all) {
  all.assign(nlevels, {});
  const bool upstream = flags & ORB_ORACLE_UPSTREAM_PYRAMID;
  const float W = 30;
  std::vector<KeyPoint> cellKps;
  for (int level = 0; level < nlevels; ++level) {
    // Image FAST runs on: the padded level (this fork, F1) or its interior ROI (upstream flag).
    Img base = pyr[level].view();
    int ox = 0, oy = 0, cols = base.cols, rows = base.rows;
    if (upstream) { ox = oy = kEdge; cols -= 2 * kEdge; rows -= 2 * kEdge; }
    const int minBX = kEdge - 3, minBY = minBX;
    const int maxBX = cols - kEdge + 3, maxBY = rows - kEdge + 3;
    std::vector<KeyPoint>& toDist = raw[level];
    toDist.clear();
    toDist.reserve((size_t)nfeatures * 10);
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
    for (int i = 0; i < nRows; ++i) {
      const float iniY = (float)(minBY + i * hCell);
      float maxY = iniY + hCell + 6;
      if (iniY >= maxBY - 3) continue;
      if (maxY > maxBY) maxY = (float)maxBY;
      for (int j = 0; j < nCols; ++j) {
        const float iniX = (float)(minBX + j * wCell);
        float maxX = iniX + wCell + 6;
        if (iniX >= maxBX - 6) continue;
        if (maxX > maxBX) maxX = (float)maxBX;
        const Img cell = base.roi(ox + (int)iniX, oy + (int)iniY, (int)maxX - (int)iniX, (int)maxY - (int)iniY);
        fast_detect(cell, iniTh, true, cellKps);
        if (cellKps.empty()) fast_detect(cell, minTh, true, cellKps);
        for (KeyPoint& k : cellKps) {
          k.x += j * wCell;
          k.y += i * hCell;
          toDist.push_back(k);
        }
      }
    }
    std::vector<KeyPoint>& kps = all[level];
    kps = distribute(toDist, minBX, maxBX, minBY, maxBY, quota[level], nfeatures);
    const int scaledPatch = (int)(kPatch * scale[level]);
    for (KeyPoint& k : kps) {
      k.x += minBX;
      k.y += minBY;
      k.octave = level;
      k.size = (float)scaledPatch;
    }
  }
  // orientation, src/ORBextractor.cc:104-161 + 933-952, on the un-blurred level
  for (int level = 0; level < nlevels; ++level) {
    Img base = pyr[level].view();
    const int off = upstream ? kEdge : 0;
    for (KeyPoint& k : all[level]) {
      const uint8_t* center = base.ptr(cv_round(k.y) + off) + cv_round(k.x) + off;
      const int step = base.step;
      int m01 = 0, m10 = 0;
      for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * center[u];
      for (int v = 1; v <= kHalfPatch; ++v) {
        int vsum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
          const int p = center[u + v * step], m = center[u - v * step];
          vsum += p - m;
          m10 += u * (p + m);
        }
        m01 += v * vsum;
      }
      k.angle = fast_atan2((float)m01, (float)m10);
    }
  }
}

// ---- DistributeOctTree, src/ORBextractor.cc:1050-1417 (+ DivideNode :963-1035) -------------------
// Node list = std::list with push_front / erase like the reference; keys held as indices into `in`
// (in-node order preserved: it decides the "first max response" winner, :1394-1409).
struct Node {
  int ulx, uly, brx, bry;  // UL.x, UL.y, BR.x, BR.y (UR = (brx, uly), BL = (ulx, bry))
  std::vector<int> keys;
  bool noMore = false;
  std::list<Node>::iterator self;
};

static void divide(const Node& n, const std::vector<KeyPoint>& in, Node out[4]) {
  const int halfX = (int)std::ceil((float)(n.brx - n.ulx) / 2);
  const int halfY = (int)std::ceil((float)(n.bry - n.uly) / 2);
  const int mx = n.ulx + halfX, my = n.uly + halfY;
  out[0].ulx = n.ulx; out[0].uly = n.uly; out[0].brx = mx;    out[0].bry = my;
  out[1].ulx = mx;    out[1].uly = n.uly; out[1].brx = n.brx; out[1].bry = my;
  out[2].ulx = n.ulx; out[2].uly = my;    out[2].brx = mx;    out[2].bry = n.bry;
  out[3].ulx = mx;    out[3].uly = my;    out[3].brx = n.brx; out[3].bry = n.bry;
  for (int c = 0; c < 4; ++c) { out[c].keys.clear(); out[c].keys.reserve(n.keys.size()); out[c].noMore = false; }
  for (int idx : n.keys) {
    const KeyPoint& kp = in[idx];
    const int c = (kp.x < (float)mx ? 0 : 1) + (kp.y < (float)my ? 0 : 2);
    out[c].keys.push_back(idx);
  }
  for (int c = 0; c < 4; ++c)
    if (out[c].keys.size() == 1) out[c].noMore = true;
}

std::vector<KeyPoint> distribute(const std::vector<KeyPoint>& in, int minX, int maxX, int minY, int maxY, int N,
                                 int reserveHint) {
  std::vector<KeyPoint> result;
  const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
  if (nIni <= 0) return result;  // the reference divides by zero here (portrait aspect < 0.5); rejected upstream
  const float hX = (float)(maxX - minX) / nIni;
  std::list<Node> nodes;
  std::vector<Node*> ini(nIni);
  for (int i = 0; i < nIni; ++i) {
    Node n;
    n.ulx = (int)(hX * (float)i); n.uly = 0;
    n.brx = (int)(hX * (float)(i + 1)); n.bry = maxY - minY;
    nodes.push_back(n);
    ini[i] = &nodes.back();
  }
  for (size_t i = 0; i < in.size(); ++i) {
    const int slot = (int)(in[i].x / hX);
    if (slot >= 0 && slot < nIni) ini[slot]->keys.push_back((int)i);  // always true for in-domain FAST points
  }
  for (auto it = nodes.begin(); it != nodes.end();) {
    if (it->keys.size() == 1) { it->noMore = true; ++it; }
    else if (it->keys.empty()) it = nodes.erase(it);
    else ++it;
  }
  // (size, seq, node): seq = creation order -> canonical replacement of the pointer tie-break (F2)
  struct Cand { int size; int seq; Node* node; };
  std::vector<Cand> cands;
  int seq = 0;
  Node kids[4];
  auto push_children = [&](bool countExpand, int& nToExpand) {
    for (int c = 0; c < 4; ++c) {
      if (kids[c].keys.empty()) continue;
      nodes.push_front(kids[c]);
      if (kids[c].keys.size() > 1) {
        if (countExpand) ++nToExpand;
        cands.push_back(Cand{(int)kids[c].keys.size(), seq++, &nodes.front()});
        nodes.front().self = nodes.begin();
      }
    }
  };
  bool finish = false;
  while (!finish) {
    int prevSize = (int)nodes.size();
    int nToExpand = 0;
    cands.clear();
    for (auto it = nodes.begin(); it != nodes.end();) {
      if (it->noMore) { ++it; continue; }
      divide(*it, in, kids);
      push_children(true, nToExpand);
      it = nodes.erase(it);
    }
    if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
      finish = true;
    } else if ((int)nodes.size() + nToExpand * 3 > N) {
      while (!finish) {
        prevSize = (int)nodes.size();
        std::vector<Cand> prev = cands;
        cands.clear();
        std::sort(prev.begin(), prev.end(), [](const Cand& a, const Cand& b) {
          return a.size != b.size ? a.size < b.size : a.seq < b.seq;
        });
        for (int j = (int)prev.size() - 1; j >= 0; --j) {
          divide(*prev[j].node, in, kids);
          int dummy = 0;
          push_children(false, dummy);
          nodes.erase(prev[j].node->self);
          if ((int)nodes.size() >= N) break;
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
      }
    }
  }
  result.reserve(reserveHint);
  for (const Node& n : nodes) {
    int best = n.keys[0];
    float maxResp = in[best].response;
    for (size_t k = 1; k < n.keys.size(); ++k)
      if (in[n.keys[k]].response > maxResp) { best = n.keys[k]; maxResp = in[best].response; }
    result.push_back(in[best]);
  }
  return result;
}

// src/ORBextractor.cc:1961-2084 (+ computeOrbDescriptor :177-254)
int Extractor::run(const Img& image, std::vector<KeyPoint>& out, std::vector<uint8_t>& desc) {
  computePyramid(image);
  std::vector<std::vector<KeyPoint>> all;
  computeKeyPoints(all);
  int total = 0;
  for (int l = 0; l < nlevels; ++l) { levelCounts[l] = (int)all[l].size(); total += levelCounts[l]; }
  out.clear(); out.reserve(total);
  desc.assign((size_t)total * 32, 0);
  const bool upstream = flags & ORB_ORACLE_UPSTREAM_PYRAMID;
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  int offset = 0;
  for (int l = 0; l < nlevels; ++l) {
    has_blur[l] = 0;
    std::vector<KeyPoint>& kps = all[l];
    if (kps.empty()) continue;
    // workingMat = clone of mvImagePyramid[l]; blur reflects at ITS border (padded image in this fork)
    Img srcv = pyr[l].view();
    if (upstream) srcv = srcv.roi(kEdge, kEdge, srcv.cols - 2 * kEdge, srcv.rows - 2 * kEdge);
    blurred[l].alloc(srcv.cols, srcv.rows);
    Img bl = blurred[l].view();
    gaussian_blur_7x7_s2(srcv, bl, flags & ORB_ORACLE_LEGACY_BLUR);
    has_blur[l] = 1;
    for (size_t i = 0; i < kps.size(); ++i) {
      const KeyPoint& k = kps[i];
      const float angle = k.angle * factorPI;
      float a, b;
      if (flags & ORB_ORACLE_LIBM_TRIG) { a = cosf(angle); b = sinf(angle); }
      else sincos_canonical(angle, &b, &a);
      const uint8_t* center = bl.ptr(cv_round(k.y)) + cv_round(k.x);
      const int step = bl.step;
      uint8_t* d = &desc[(size_t)(offset + i) * 32];
      for (int byte = 0; byte < 32; ++byte) {
        int val = 0;
        for (int bit = 0; bit < 8; ++bit) {
          const int8_t* p = &kPattern[(byte * 8 + bit) * 4];
          const float x0 = (float)p[0], y0 = (float)p[1], x1 = (float)p[2], y1 = (float)p[3];
          const int t0 = center[cv_round(x0 * b + y0 * a) * step + cv_round(x0 * a - y0 * b)];
          const int t1 = center[cv_round(x1 * b + y1 * a) * step + cv_round(x1 * a - y1 * b)];
          val |= (t0 < t1) << bit;
        }
        d[byte] = (uint8_t)val;
      }
    }
    offset += (int)kps.size();
    if (l != 0) {
      const float s = scale[l];
      for (KeyPoint& k : kps) { k.x *= s; k.y *= s; }
    }
    out.insert(out.end(), kps.begin(), kps.end());
  }
  return total;
}

}  // namespace orb_oracle

// ---------------------------------------------------------------------------------------------
// C API
// ---------------------------------------------------------------------------------------------
using namespace orb_oracle;

struct orb_oracle_extractor {
  Extractor ex;
  std::vector<KeyPoint> kps;
  std::vector<uint8_t> desc;
  orb_oracle_extractor(int nf, float sf, int nl, int ini, int mn, int fl) : ex(nf, sf, nl, ini, mn, fl) {}
};

static_assert(sizeof(orb_oracle_kp) == sizeof(KeyPoint), "kp layout");

extern "C" {

orb_oracle_extractor* orb_oracle_create(int nf, float sf, int nl, int ini, int mn, int flags) {
  if (nf < 0 || nl < 1 || nl > 32 || !(sf > 1.f)) return nullptr;
  return new orb_oracle_extractor(nf, sf, nl, ini, mn, flags);
}
void orb_oracle_destroy(orb_oracle_extractor* e) { delete e; }

void orb_oracle_get_table(const orb_oracle_extractor* e, int what, float* out) {
  const std::vector<float>& t = what == 0 ? e->ex.scale : what == 1 ? e->ex.invScale : what == 2 ? e->ex.sigma2 : e->ex.invSigma2;
  std::copy(t.begin(), t.end(), out);
}
void orb_oracle_get_quota(const orb_oracle_extractor* e, int32_t* out) { std::copy(e->ex.quota.begin(), e->ex.quota.end(), out); }
void orb_oracle_get_umax(const orb_oracle_extractor* e, int32_t* out) { std::copy(e->ex.umax.begin(), e->ex.umax.end(), out); }

int orb_oracle_extract(orb_oracle_extractor* e, const uint8_t* img, int step, int width, int height, orb_oracle_kp* kps,
                       uint8_t* desc, int cap, int32_t* level_counts) {
  if (!e || !img || width <= 0 || height <= 0) return -1;
  Img im{const_cast<uint8_t*>(img), width, height, step};
  const int n = e->ex.run(im, e->kps, e->desc);
  if (level_counts) std::copy(e->ex.levelCounts.begin(), e->ex.levelCounts.end(), level_counts);
  if (n > cap) return -1;
  if (kps && n) std::memcpy(kps, e->kps.data(), (size_t)n * sizeof(KeyPoint));
  if (desc && n) std::memcpy(desc, e->desc.data(), (size_t)n * 32);
  return n;
}

int orb_oracle_level_dims(const orb_oracle_extractor* e, int l, int* cols, int* rows) {
  if (l < 0 || l >= e->ex.nlevels) return -1;
  *cols = e->ex.pyr[l].cols; *rows = e->ex.pyr[l].rows;
  return 0;
}
const uint8_t* orb_oracle_level_ptr(const orb_oracle_extractor* e, int l) { return e->ex.pyr[l].buf.data(); }
const uint8_t* orb_oracle_blurred_ptr(const orb_oracle_extractor* e, int l) {
  return e->ex.has_blur[l] ? e->ex.blurred[l].buf.data() : nullptr;
}
int orb_oracle_raw_count(const orb_oracle_extractor* e, int l) { return (int)e->ex.raw[l].size(); }
int orb_oracle_raw_keypoints(const orb_oracle_extractor* e, int l, orb_oracle_kp* out, int cap) {
  const int n = (int)e->ex.raw[l].size();
  if (n > cap) return -1;
  if (n) std::memcpy(out, e->ex.raw[l].data(), (size_t)n * sizeof(KeyPoint));
  return n;
}

void orb_oracle_resize(const uint8_t* src, int sstep, int sw, int sh, uint8_t* dst, int dstep, int dw, int dh) {
  resize_linear_u8(Img{const_cast<uint8_t*>(src), sw, sh, sstep}, Img{dst, dw, dh, dstep});
}
void orb_oracle_border101(const uint8_t* src, int sstep, int sw, int sh, uint8_t* dst, int dstep, int b) {
  copy_make_border_reflect101(Img{const_cast<uint8_t*>(src), sw, sh, sstep}, Img{dst, sw + 2 * b, sh + 2 * b, dstep}, b);
}
void orb_oracle_blur(const uint8_t* src, int sstep, int w, int h, uint8_t* dst, int dstep, int legacy) {
  gaussian_blur_7x7_s2(Img{const_cast<uint8_t*>(src), w, h, sstep}, Img{dst, w, h, dstep}, legacy != 0);
}
int orb_oracle_fast(const uint8_t* img, int step, int w, int h, int threshold, orb_oracle_kp* out, int cap) {
  std::vector<KeyPoint> k;
  fast_detect(Img{const_cast<uint8_t*>(img), w, h, step}, threshold, true, k);
  if ((int)k.size() > cap) return -1;
  if (!k.empty()) std::memcpy(out, k.data(), k.size() * sizeof(KeyPoint));
  return (int)k.size();
}
void orb_oracle_fast_score_map(const uint8_t* img, int step, int w, int h, uint8_t* out) {
  std::memset(out, 0, (size_t)w * h);
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) out[(size_t)y * w + x] = (uint8_t)fast_corner_score(img + (size_t)y * step + x, step);
}
float orb_oracle_fast_atan2(float y, float x) { return fast_atan2(y, x); }
void orb_oracle_sincos(float a, int libm, float* s, float* c) {
  if (libm) { *s = sinf(a); *c = cosf(a); }
  else sincos_canonical(a, s, c);
}
int orb_oracle_distribute(const orb_oracle_kp* in, int n, int minx, int maxx, int miny, int maxy, int nfeat,
                          orb_oracle_kp* out, int cap) {
  std::vector<KeyPoint> v(n);
  if (n) std::memcpy(v.data(), in, (size_t)n * sizeof(KeyPoint));
  std::vector<KeyPoint> r = distribute(v, minx, maxx, miny, maxy, nfeat, nfeat);
  if ((int)r.size() > cap) return -1;
  if (!r.empty()) std::memcpy(out, r.data(), r.size() * sizeof(KeyPoint));
  return (int)r.size();
}

// src/ORBmatcher.cc:2073-2093: 8 x 32-bit words, SWAR popcount.
int orb_oracle_descriptor_distance(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; ++i) {
    uint32_t x, y;
    std::memcpy(&x, a + 4 * i, 4);
    std::memcpy(&y, b + 4 * i, 4);
    uint32_t v = x ^ y;
    v = v - ((v >> 1) & 0x55555555u);
    v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
    dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
  }
  return dist;
}

void orb_oracle_hamming_knn(const uint8_t* q, int Q, const uint8_t* db, int64_t D, int64_t base, int sentinel,
                            orb_oracle_knn2* out, int nthreads) {
  auto work = [&](int lo, int hi) {
    for (int qi = lo; qi < hi; ++qi) {
      int best = sentinel, second = sentinel;
      int64_t i1 = -1, i2 = -1;
      const uint8_t* qd = q + (size_t)qi * 32;
      for (int64_t j = 0; j < D; ++j) {
        const int d = orb_oracle_descriptor_distance(qd, db + (size_t)j * 32);
        if (d < best) { second = best; i2 = i1; best = d; i1 = base + j; }
        else if (d < second) { second = d; i2 = base + j; }
      }
      out[qi] = orb_oracle_knn2{best, second, i1, i2};
    }
  };
  nthreads = std::max(1, std::min(nthreads, Q));
  if (nthreads == 1) { work(0, Q); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back(work, (int)((int64_t)Q * t / nthreads), (int)((int64_t)Q * (t + 1) / nthreads));
  for (auto& t : th) t.join();
}

// Frame::ComputeStereoMatches (src/Frame.cc:880-1176) for one rectified pair, on the keypoints / descriptors / pyramids the two
// extractors produced. bf = mbf, b = mb. uright / depth have nL entries (-1 = no match).
// Canonicalisations: rows outside [0, nRows) are ignored when the right keypoints are binned (the reference indexes
// vRowIndices with them, which is out of bounds for this fork's padded coordinates at high octaves); an empty match list skips the
// median filter (the reference reads vDistIdx[0] of an empty vector); float expressions are not contracted.
int orb_oracle_stereo_match(const orb_oracle_extractor* left, const orb_oracle_extractor* right, const orb_oracle_kp* kl,
                            const uint8_t* dl, int nL, const orb_oracle_kp* kr, const uint8_t* dr, int nR, float bf, float b,
                            float* uright, float* depth) {
  const int TH_HIGH = 100, TH_LOW = 50;
  const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
  const Extractor& EL = left->ex;
  const Extractor& ER = right->ex;
  // UPSTREAM_PYRAMID: mvImagePyramid[l] is the interior of the stored padded buffer
  const int off = (EL.flags & ORB_ORACLE_UPSTREAM_PYRAMID) ? kEdge : 0;
  const int nRows = EL.pyr[0].rows - 2 * off;
  for (int i = 0; i < nL; ++i) { uright[i] = -1.0f; depth[i] = -1.0f; }
  std::vector<std::vector<int>> rowIdx(nRows);
  for (int iR = 0; iR < nR; ++iR) {
    const float kpY = kr[iR].y;
    const float r = 2.0f * ER.scale[kr[iR].octave];
    const int maxr = (int)std::ceil(kpY + r);
    const int minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; ++yi)
      if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR);
  }
  const float minZ = b;
  const float minD = 0;
  const float maxD = bf / minZ;
  std::vector<std::pair<int, int>> distIdx;
  for (int iL = 0; iL < nL; ++iL) {
    const orb_oracle_kp& kpL = kl[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int row = (int)vL;
    if (row < 0 || row >= nRows) continue;
    const std::vector<int>& cand = rowIdx[row];
    if (cand.empty()) continue;
    const float minU = uL - maxD;
    const float maxU = uL - minD;
    if (maxU < 0) continue;
    int bestDist = TH_HIGH;
    int bestIdxR = 0;
    for (int iR : cand) {
      const orb_oracle_kp& kpR = kr[iR];
      if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
      const float uR = kpR.x;
      if (uR >= minU && uR <= maxU) {
        const int dist = orb_oracle_descriptor_distance(dl + (size_t)iL * 32, dr + (size_t)iR * 32);
        if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
      }
    }
    if (bestDist >= thOrbDist) continue;
    const float uR0 = kr[bestIdxR].x;
    const float sf = EL.invScale[kpL.octave];
    const float scaleduL = std::round(kpL.x * sf);
    const float scaledvL = std::round(kpL.y * sf);
    const float scaleduR0 = std::round(uR0 * sf);
    const int w = 5, L = 5;
    const Plane& PL = EL.pyr[kpL.octave];
    const Plane& PR = ER.pyr[kpL.octave];
    const int cy = (int)scaledvL + off, cxl = (int)scaleduL + off, cxr = (int)scaleduR0 + off;
    const float iniu = scaleduR0 - L - w;
    const float endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= PR.cols - 2 * off) continue;
    int bestSad = INT_MAX, bestinc = 0;
    float vd[2 * 5 + 1];
    const float cl = (float)PL.buf[(size_t)cy * PL.cols + cxl];
    for (int inc = -L; inc <= L; ++inc) {
      const float cr = (float)PR.buf[(size_t)cy * PR.cols + cxr + inc];
      double acc = 0.0;   // cv::norm(NORM_L1) of two CV_32F patches accumulates in double; every term is an integer
      for (int dy = -w; dy <= w; ++dy)
        for (int dx = -w; dx <= w; ++dx) {
          const float a = (float)PL.buf[(size_t)(cy + dy) * PL.cols + cxl + dx] - cl;
          const float c = (float)PR.buf[(size_t)(cy + dy) * PR.cols + cxr + inc + dx] - cr;
          acc += std::fabs((double)(a - c));
        }
      const float dist = (float)acc;
      if (dist < (float)bestSad) { bestSad = (int)dist; bestinc = inc; }
      vd[L + inc] = dist;
    }
    if (bestinc == -L || bestinc == L) continue;
    const float d1 = vd[L + bestinc - 1], d2 = vd[L + bestinc], d3 = vd[L + bestinc + 1];
    const float deltaR = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));
    if (deltaR < -1 || deltaR > 1) continue;
    float bestuR = EL.scale[kpL.octave] * ((float)scaleduR0 + (float)bestinc + deltaR);
    float disparity = uL - bestuR;
    if (disparity >= minD && disparity < maxD) {
      if (disparity <= 0) {
        disparity = 0.01;
        bestuR = uL - 0.01;
      }
      depth[iL] = bf / disparity;
      uright[iL] = bestuR;
      distIdx.push_back(std::make_pair(bestSad, iL));
    }
  }
  if (distIdx.empty()) return 0;
  std::sort(distIdx.begin(), distIdx.end());
  const float median = distIdx[distIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  int kept = (int)distIdx.size();
  for (int i = (int)distIdx.size() - 1; i >= 0; --i) {
    if (distIdx[i].first < thDist) break;
    uright[distIdx[i].second] = -1;
    depth[distIdx[i].second] = -1;
    --kept;
  }
  return kept;
}

// ---- guided matching: Frame grid + SearchByProjection candidate loops + ComputeThreeMaxima ------------------------------------
namespace {
constexpr int GRID_COLS = 64, GRID_ROWS = 48, HISTO_LENGTH = 30;   // include/Frame.h FRAME_GRID_COLS / ROWS, include/ORBmatcher.h:54

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2026-2068) on the bin sizes: indices of the three fullest bins (first bin wins ties),
// the 2nd / 3rd dropped when below 10 % of the fullest (float compare as written). Pinned against the reference's own function by
// tests/test_reference_binary.py::test_three_maxima_equals_reference.
static void three_maxima(const int* size, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; ++i) {
    const int s = size[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}
extern "C" void orb_oracle_three_maxima(const int32_t* sizes, int L, int32_t* out3) {
  int i1 = -1, i2 = -1, i3 = -1;
  three_maxima(sizes, L, i1, i2, i3);
  out3[0] = i1; out3[1] = i2; out3[2] = i3;
}

struct FrameGrid {   // mGrid, mfGridElementWidthInv / HeightInv (src/Frame.cc:200-202, 432-460)
  float min_x, min_y, inv_w, inv_h;
  std::vector<int> cell[GRID_COLS][GRID_ROWS];
  FrameGrid(const orb_oracle_kp* k, int n, const orb_oracle_search_params& p) : min_x(p.min_x), min_y(p.min_y) {
    inv_w = static_cast<float>(GRID_COLS) / static_cast<float>(p.max_x - p.min_x);
    inv_h = static_cast<float>(GRID_ROWS) / static_cast<float>(p.max_y - p.min_y);
    for (int i = 0; i < n; ++i) {   // AssignFeaturesToGrid + PosInGrid (src/Frame.cc:447-459, 734-744)
      const int px = (int)std::round((k[i].x - min_x) * inv_w);
      const int py = (int)std::round((k[i].y - min_y) * inv_h);
      if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
      cell[px][py].push_back(i);
    }
  }
  // GetFeaturesInArea (src/Frame.cc:633-717)
  void in_area(const orb_oracle_kp* k, float x, float y, float r, int min_level, int max_level, std::vector<int>& out) const {
    out.clear();
    const int cx0 = std::max(0, (int)std::floor((x - min_x - r) * inv_w));
    if (cx0 >= GRID_COLS) return;
    const int cx1 = std::min(GRID_COLS - 1, (int)std::ceil((x - min_x + r) * inv_w));
    if (cx1 < 0) return;
    const int cy0 = std::max(0, (int)std::floor((y - min_y - r) * inv_h));
    if (cy0 >= GRID_ROWS) return;
    const int cy1 = std::min(GRID_ROWS - 1, (int)std::ceil((y - min_y + r) * inv_h));
    if (cy1 < 0) return;
    const bool check_levels = (min_level > 0) || (max_level >= 0);
    for (int ix = cx0; ix <= cx1; ++ix)
      for (int iy = cy0; iy <= cy1; ++iy)
        for (int i : cell[ix][iy]) {
          if (check_levels) {
            if (k[i].octave < min_level) continue;
            if (max_level >= 0 && k[i].octave > max_level) continue;
          }
          const float dx = k[i].x - x, dy = k[i].y - y;
          if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(i);
        }
  }
};
}  // namespace

int orb_oracle_features_in_area(const orb_oracle_kp* kps, int n, const orb_oracle_search_params* prm, float x, float y, float r,
                                int min_level, int max_level, int32_t* out) {
  FrameGrid grid(kps, n, *prm);
  std::vector<int> v;
  grid.in_area(kps, x, y, r, min_level, max_level, v);
  std::copy(v.begin(), v.end(), out);
  return (int)v.size();
}

// The candidate loop of ORBmatcher::Fuse (src/ORBmatcher.cc:1166-1232; without the chi-square test: the Scw form :1380-1405 and the
// projection passes of SearchBySim3 :1540-1580, :1640-1680), one query at a time, no state between queries.
void orb_oracle_fuse_search(const orb_oracle_kp* kps, const uint8_t* desc, const float* uright, int n, const orb_oracle_query* q,
                            const uint8_t* qdesc, int m, const orb_oracle_search_params* prm, const float* inv_level_sigma2, int32_t* best_idx,
                            int32_t* best_dist) {
  FrameGrid grid(kps, n, *prm);
  std::vector<int> vIndices;
  for (int j = 0; j < m; ++j) {
    best_idx[j] = -1;
    best_dist[j] = 256;
    if (q[j].flags & 1) continue;
    const float u = q[j].u, v = q[j].v, ur = q[j].ur;
    grid.in_area(kps, u, v, q[j].radius, -1, -1, vIndices);   // pKF->GetFeaturesInArea(u, v, radius) (:1166)
    int bestDist = 256, bestIdx = -1;
    for (int idx : vIndices) {
      const orb_oracle_kp& kp = kps[idx];
      const int kpLevel = kp.octave;
      if (kpLevel < q[j].min_level || kpLevel > q[j].max_level) continue;   // nPredictedLevel - 1 .. nPredictedLevel (:1188)
      if (inv_level_sigma2) {
        if (uright && uright[idx] >= 0) {   // :1193-1206
          const float ex = u - kp.x;
          const float ey = v - kp.y;
          const float er = ur - uright[idx];
          const float e2 = ex * ex + ey * ey + er * er;
          if (e2 * inv_level_sigma2[kpLevel] > 7.8) continue;
        } else {                            // :1207-1218
          const float ex = u - kp.x;
          const float ey = v - kp.y;
          const float e2 = ex * ex + ey * ey;
          if (e2 * inv_level_sigma2[kpLevel] > 5.99) continue;
        }
      }
      const int dist = orb_oracle_descriptor_distance(qdesc + (size_t)j * 32, desc + (size_t)idx * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    best_dist[j] = bestDist;
    if (bestDist <= prm->th_dist) best_idx[j] = bestIdx;   // :1232 (TH_LOW), :1575 (TH_HIGH)
  }
}

int orb_oracle_search_by_projection(const orb_oracle_kp* kps, const uint8_t* desc, const float* uright, const uint8_t* claimed_in,
                                    int n, const orb_oracle_query* q, const uint8_t* qdesc, int m,
                                    const orb_oracle_search_params* prm, int32_t* assigned, int32_t* query_match) {
  FrameGrid grid(kps, n, *prm);
  std::vector<uint8_t> claimed(n, 0);   // "F.mvpMapPoints[idx] && F.mvpMapPoints[idx]->Observations() > 0" (src/ORBmatcher.cc:127-130)
  if (claimed_in) std::copy(claimed_in, claimed_in + n, claimed.begin());
  std::fill(assigned, assigned + n, -1);
  if (query_match) std::fill(query_match, query_match + m, -1);
  std::vector<std::pair<int, int>> rot_hist[HISTO_LENGTH];   // (query, keypoint) per accepted match (:1839)
  const float factor = HISTO_LENGTH / 360.0f;                // this fork's corrected factor (:1713)
  int nmatches = 0;
  std::vector<int> cand;
  for (int iq = 0; iq < m; ++iq) {
    const orb_oracle_query& Q = q[iq];
    if (Q.flags & 1) continue;
    grid.in_area(kps, Q.u, Q.v, Q.radius, Q.min_level, Q.max_level, cand);
    if (cand.empty()) continue;
    int best = 256, best2 = 256, best_level = -1, best_level2 = -1, best_idx = -1;   // :112-116, :1791-1792
    for (int idx : cand) {
      if (claimed[idx]) continue;
      if (uright && uright[idx] > 0) {   // :134-141, :1804-1811
        const float er = std::fabs(Q.ur - uright[idx]);
        if (er > Q.radius) continue;
      }
      const int dist = orb_oracle_descriptor_distance(qdesc + (size_t)iq * 32, desc + (size_t)idx * 32);
      if (dist < best) {
        best2 = best; best = dist;
        best_level2 = best_level; best_level = kps[idx].octave;
        best_idx = idx;
      } else if (dist < best2) {
        best_level2 = kps[idx].octave;
        best2 = dist;
      }
    }
    if (best <= prm->th_dist) {   // :170, :1824 (best_idx >= 0 is implied for th_dist < 256)
      if (best_idx < 0) continue;
      if (prm->nn_ratio > 0.f && best_level == best_level2 && best > prm->nn_ratio * best2) continue;   // :174
      assigned[best_idx] = iq;    // F.mvpMapPoints[bestIdx] = pMP (:177, :1826)
      if (Q.flags & 2) claimed[best_idx] = 1;
      if (query_match) query_match[iq] = best_idx;
      ++nmatches;
      if (prm->check_orientation) {   // :1828-1839
        float rot = Q.angle - kps[best_idx].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)std::round(rot * factor);
        if (bin == HISTO_LENGTH) bin = 0;
        if (bin >= 0 && bin < HISTO_LENGTH) rot_hist[bin].push_back({iq, best_idx});   // the reference asserts the range
      }
    }
  }
  if (prm->check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hsize[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; ++i) hsize[i] = (int)rot_hist[i].size();
    three_maxima(hsize, HISTO_LENGTH, ind1, ind2, ind3);   // ComputeThreeMaxima (:2026-2068)
    for (int i = 0; i < HISTO_LENGTH; ++i) {   // :1849-1860
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (const auto& e : rot_hist[i]) {
        assigned[e.second] = -1;
        if (query_match) query_match[e.first] = -1;
        --nmatches;
      }
    }
  }
  return nmatches;
}

int orb_oracle_search_for_initialization(const orb_oracle_kp* kps, const uint8_t* desc, int n, const orb_oracle_query* q, const uint8_t* qdesc,
                                         int m, const orb_oracle_search_params* prm, int32_t* vnMatches12, int32_t* matches21) {
  FrameGrid grid(kps, n, *prm);
  int nmatches = 0;
  std::fill(vnMatches12, vnMatches12 + m, -1);                       // :574
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = HISTO_LENGTH / 360.0f;                        // :583
  std::vector<int> vMatchedDistance(n, INT_MAX), vnMatches21(n, -1); // :586-588
  std::vector<int> vIndices2;
  for (int i1 = 0; i1 < m; ++i1) {
    if (q[i1].flags & 1) continue;                                   // level1 > 0 (:596-597)
    grid.in_area(kps, q[i1].u, q[i1].v, q[i1].radius, q[i1].min_level, q[i1].max_level, vIndices2);
    if (vIndices2.empty()) continue;
    const uint8_t* d1 = qdesc + (size_t)i1 * 32;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;      // :611-613
    for (int i2 : vIndices2) {
      const int dist = orb_oracle_descriptor_distance(d1, desc + (size_t)i2 * 32);
      if (vMatchedDistance[i2] <= dist) continue;                    // :624
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= prm->th_dist) {                                  // :641
      if (bestDist < (float)bestDist2 * prm->nn_ratio) {             // :644
        if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }   // :647-651
        vnMatches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (prm->check_orientation) {
          float rot = q[i1].angle - kps[bestIdx2].angle;
          if (rot < 0.0) rot += 360.0f;
          int bin = (int)std::round(rot * factor);
          if (bin == HISTO_LENGTH) bin = 0;
          if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(i1);   // :673
        }
      }
    }
  }
  if (prm->check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hsize[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; ++i) hsize[i] = (int)rotHist[i].size();
    three_maxima(hsize, HISTO_LENGTH, ind1, ind2, ind3);   // ComputeThreeMaxima (:2026-2068)
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i])
        if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }   // :696-700
    }
  }
  if (matches21) {
    std::fill(matches21, matches21 + n, -1);
    for (int i1 = 0; i1 < m; ++i1)
      if (vnMatches12[i1] >= 0) matches21[vnMatches12[i1]] = i1;
  }
  return nmatches;
}

// ORBmatcher::CheckDistEpipolarLine (src/ORBmatcher.cc:206-245): float32, one rounding per operation (no contraction)
static bool check_dist_epipolar_line(const orb_oracle_kp& kp1, const orb_oracle_kp& kp2, const float* F12, const float* level_sigma2) {
  const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
  const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
  const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
  const float num = a * kp2.x + b * kp2.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * level_sigma2[kp2.octave];   // double compare, as written in the reference
}

int orb_oracle_search_for_triangulation(const uint8_t* desc1, const orb_oracle_kp* kps1_un, const uint8_t* kf1_has_mp, const float* uright1, int n1,
                                        const int32_t* node1, const int32_t* start1, const int32_t* feature1, int count1,
                                        const uint8_t* desc2, const orb_oracle_kp* kps2_un, const uint8_t* kf2_has_mp, const float* uright2, int n2,
                                        const int32_t* node2, const int32_t* start2, const int32_t* feature2, int count2,
                                        const float* f12, float ex, float ey, const float* level_sigma2, const float* scale_factor,
                                        int only_stereo, int th_low, int check_orientation, int32_t* matches12) {
  typedef std::map<unsigned, std::vector<unsigned>> FeatureVector;
  FeatureVector vFeatVec1, vFeatVec2;
  for (int j = 0; j < count1; ++j) vFeatVec1[node1[j]].assign(feature1 + start1[j], feature1 + start1[j + 1]);
  for (int j = 0; j < count2; ++j) vFeatVec2[node2[j]].assign(feature2 + start2[j], feature2 + start2[j + 1]);
  int nmatches = 0;
  std::vector<bool> vbMatched2(n2, false);          // :899
  std::vector<int> vMatches12(n1, -1);              // :900
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = HISTO_LENGTH / 360.0f;       // :906
  FeatureVector::const_iterator f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin(), f1end = vFeatVec1.end(), f2end = vFeatVec2.end();
  while (f1it != f1end && f2it != f2end) {
    if (f1it->first == f2it->first) {
      for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
        const size_t idx1 = f1it->second[i1];
        if (kf1_has_mp[idx1]) continue;                                  // :925-929
        const bool bStereo1 = uright1[idx1] >= 0;
        if (only_stereo && !bStereo1) continue;
        const orb_oracle_kp& kp1 = kps1_un[idx1];
        const uint8_t* d1 = desc1 + idx1 * 32;
        int bestDist = th_low;                                           // :945
        int bestIdx2 = -1;
        for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
          const size_t idx2 = f2it->second[i2];
          if (vbMatched2[idx2] || kf2_has_mp[idx2]) continue;            // :957
          const bool bStereo2 = uright2[idx2] >= 0;
          if (only_stereo && !bStereo2) continue;
          const int dist = orb_oracle_descriptor_distance(d1, desc2 + idx2 * 32);
          if (dist > th_low || dist > bestDist) continue;                // :975
          const orb_oracle_kp& kp2 = kps2_un[idx2];
          if (!bStereo1 && !bStereo2) {                                  // :1005-1013
            const float distex = ex - kp2.x;
            const float distey = ey - kp2.y;
            if (distex * distex + distey * distey < 100 * scale_factor[kp2.octave]) continue;
          }
          if (check_dist_epipolar_line(kp1, kp2, f12, level_sigma2)) {
            bestIdx2 = (int)idx2;
            bestDist = dist;
          }
        }
        if (bestIdx2 >= 0) {
          const orb_oracle_kp& kp2 = kps2_un[bestIdx2];
          vMatches12[idx1] = bestIdx2;
          vbMatched2[bestIdx2] = true;
          nmatches++;
          if (check_orientation) {
            float rot = kp1.angle - kp2.angle;
            if (rot < 0.0) rot += 360.0f;
            int bin = (int)std::round(rot * factor);
            if (bin == HISTO_LENGTH) bin = 0;
            if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back((int)idx1);
          }
        }
      }
      f1it++; f2it++;
    } else if (f1it->first < f2it->first) {
      f1it = vFeatVec1.lower_bound(f2it->first);
    } else {
      f2it = vFeatVec2.lower_bound(f1it->first);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hsize[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; ++i) hsize[i] = (int)rotHist[i].size();
    three_maxima(hsize, HISTO_LENGTH, ind1, ind2, ind3);   // ComputeThreeMaxima (:2026-2068)
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i]) {
        vbMatched2[vMatches12[idx1]] = false;
        vMatches12[idx1] = -1;
        nmatches--;
      }
    }
  }
  for (int i = 0; i < n1; ++i) matches12[i] = vMatches12[i];
  return nmatches;
}

int orb_oracle_search_by_bow(const uint8_t* kf_desc, const orb_oracle_kp* kf_kps_un, const uint8_t* kf_valid, const int32_t* kf_node,
                             const int32_t* kf_start, const int32_t* kf_feature, int kf_count, const uint8_t* f_desc,
                             const orb_oracle_kp* f_kps, int n_f, const int32_t* f_node, const int32_t* f_start, const int32_t* f_feature,
                             int f_count, int th_dist, float nn_ratio, int check_orientation, int32_t* assigned,
                             const uint8_t* f_blocked, int32_t* kf_match, int n_kf) {
  typedef std::map<unsigned, std::vector<unsigned>> FeatureVector;
  FeatureVector fvKF, fvF;
  for (int j = 0; j < kf_count; ++j) fvKF[kf_node[j]].assign(kf_feature + kf_start[j], kf_feature + kf_start[j + 1]);
  for (int j = 0; j < f_count; ++j) fvF[f_node[j]].assign(f_feature + f_start[j], f_feature + f_start[j + 1]);
  std::fill(assigned, assigned + n_f, -1);   // vpMapPointMatches = vector<MapPoint*>(F.N, NULL) (:255)
  if (kf_match) std::fill(kf_match, kf_match + n_kf, -1);   // vpMatches12 of the KeyFrame-KeyFrame form (:735)
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  const float factor = HISTO_LENGTH / 360.0f;   // :268
  FeatureVector::const_iterator KFit = fvKF.begin(), Fit = fvF.begin(), KFend = fvKF.end(), Fend = fvF.end();
  while (KFit != KFend && Fit != Fend) {
    if (KFit->first == Fit->first) {
      const std::vector<unsigned>& vKF = KFit->second;
      const std::vector<unsigned>& vF = Fit->second;
      for (size_t iKF = 0; iKF < vKF.size(); ++iKF) {
        const unsigned realIdxKF = vKF[iKF];
        if (!kf_valid[realIdxKF]) continue;   // !pMP || pMP->isBad() (:283-289)
        const uint8_t* dKF = kf_desc + (size_t)realIdxKF * 32;
        int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
        for (size_t iF = 0; iF < vF.size(); ++iF) {
          const unsigned realIdxF = vF[iF];
          if (assigned[realIdxF] >= 0) continue;   // :300-301 (vbMatched2[idx2], :771)
          if (f_blocked && f_blocked[realIdxF]) continue;   // !pMP2 || pMP2->isBad() (:771-776)
          const int dist = orb_oracle_descriptor_distance(dKF, f_desc + (size_t)realIdxF * 32);
          if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = (int)realIdxF; }
          else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist1 <= th_dist) {   // :339
          if (static_cast<float>(bestDist1) < nn_ratio * static_cast<float>(bestDist2)) {   // :342
            assigned[bestIdxF] = (int)realIdxKF;
            if (kf_match) kf_match[realIdxKF] = bestIdxF;
            if (check_orientation) {
              float rot = kf_kps_un[realIdxKF].angle - f_kps[bestIdxF].angle;
              if (rot < 0.0) rot += 360.0f;
              int bin = (int)std::round(rot * factor);
              if (bin == HISTO_LENGTH) bin = 0;
              if (bin >= 0 && bin < HISTO_LENGTH) rotHist[bin].push_back(bestIdxF);
            }
            ++nmatches;
          }
        }
      }
      ++KFit; ++Fit;
    } else if (KFit->first < Fit->first) {
      KFit = fvKF.lower_bound(Fit->first);
    } else {
      Fit = fvF.lower_bound(KFit->first);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1, hsize[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; ++i) hsize[i] = (int)rotHist[i].size();
    three_maxima(hsize, HISTO_LENGTH, ind1, ind2, ind3);   // ComputeThreeMaxima (:2026-2068)
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx : rotHist[i]) {
        if (kf_match && assigned[idx] >= 0) kf_match[assigned[idx]] = -1;
        assigned[idx] = -1;
        --nmatches;
      }
    }
  }
  return nmatches;
}

// ---- cv::undistortPoints as Frame::UndistortKeyPoints calls it (src/Frame.cc:787-793) -----------------------------------------
void orb_oracle_undistort_points(float ffx, float ffy, float fcx, float fcy, const float* dist, int n_dist, float* xy, int n) {
  if (dist[0] == 0.0f) return;   // :772-776
  const double fx = ffx, fy = ffy, cx = fcx, cy = fcy;
  double k[14] = {0};
  for (int i = 0; i < n_dist && i < 14; ++i) k[i] = dist[i];
  const double ifx = 1. / fx, ify = 1. / fy;
  for (int i = 0; i < n; ++i) {
    const double u = xy[2 * i], v = xy[2 * i + 1];
    double x = (u - cx) * ifx, y = (v - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {   // TermCriteria(MAX_ITER, 5, 0.01)
      const double r2 = x * x + y * y;
      const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
      const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
      const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    // RR = P * I with P = K: rows (fx 0 cx), (0 fy cy), (0 0 1)
    const double xx = fx * x + 0.0 * y + cx, yy = 0.0 * x + fy * y + cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
    xy[2 * i] = (float)(xx * ww);
    xy[2 * i + 1] = (float)(yy * ww);
  }
}

// ---- DBoW2 vocabulary transform ----------------------------------------------------------------------------------------------
struct orb_oracle_vocabulary {
  struct Node {   // TemplatedVocabulary::Node (TemplatedVocabulary.h:270-310)
    int parent = 0;
    std::vector<int> children;
    uint8_t descriptor[32] = {0};
    double weight = 0;
    int word_id = -1;
    bool is_leaf() const { return children.empty(); }
  };
  int k, L;
  std::vector<Node> nodes;
  int n_words = 0;
};

orb_oracle_vocabulary* orb_oracle_vocabulary_create(int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf,
                                                    const uint8_t* desc, const double* weight) {
  auto* v = new orb_oracle_vocabulary;
  v->k = k; v->L = L;
  v->nodes.resize(n_nodes);
  for (int nid = 1; nid < n_nodes; ++nid) {   // loadFromTextFile (TemplatedVocabulary.h:1380-1420): one node per line, ids in line order
    auto& nd = v->nodes[nid];
    nd.parent = parent[nid];
    v->nodes[parent[nid]].children.push_back(nid);
    std::memcpy(nd.descriptor, desc + (size_t)nid * 32, 32);
    nd.weight = weight[nid];
    if (is_leaf[nid]) nd.word_id = v->n_words++;
  }
  return v;
}
void orb_oracle_vocabulary_destroy(orb_oracle_vocabulary* v) { delete v; }

void orb_oracle_bow_transform(const orb_oracle_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* word, int32_t* node,
                              int32_t* bow_word, double* bow_value, int32_t* bow_count, int32_t* fv_node, int32_t* fv_start,
                              int32_t* fv_feature, int32_t* fv_count) {
  std::map<unsigned, double> bow;                       // BowVector
  std::map<unsigned, std::vector<unsigned>> fv;         // FeatureVector
  for (int i = 0; i < n; ++i) {
    // transform(feature, word_id, weight, nid, levelsup) (TemplatedVocabulary.h:1240-1280)
    const uint8_t* f = desc + (size_t)i * 32;
    const int nid_level = v->L - levelsup;
    int nid = 0, final_id = 0, current_level = 0;
    do {
      ++current_level;
      const std::vector<int>& ch = v->nodes[final_id].children;
      final_id = ch[0];
      double best_d = orb_oracle_descriptor_distance(f, v->nodes[final_id].descriptor);
      for (size_t c = 1; c < ch.size(); ++c) {
        const double d = orb_oracle_descriptor_distance(f, v->nodes[ch[c]].descriptor);
        if (d < best_d) { best_d = d; final_id = ch[c]; }
      }
      if (current_level == nid_level) nid = final_id;
    } while (!v->nodes[final_id].is_leaf());
    const int wid = v->nodes[final_id].word_id;
    const double w = v->nodes[final_id].weight;
    if (word) word[i] = wid;
    if (node) node[i] = nid;
    if (w > 0) {   // not stopped (:1163)
      auto it = bow.lower_bound(wid);   // BowVector::addWeight
      if (it != bow.end() && !(bow.key_comp()(wid, it->first))) it->second += w;
      else bow.insert(it, {(unsigned)wid, w});
      fv[nid].push_back(i);             // FeatureVector::addFeature
    }
  }
  // TF_IDF with L1 scoring: mustNormalize -> v.normalize(L1) (BowVector.cpp)
  double norm = 0.0;
  for (auto& e : bow) norm += std::fabs(e.second);
  if (norm > 0.0)
    for (auto& e : bow) e.second /= norm;
  int c = 0;
  for (auto& e : bow) { bow_word[c] = (int)e.first; bow_value[c] = e.second; ++c; }
  *bow_count = c;
  c = 0;
  int pos = 0;
  for (auto& e : fv) {
    fv_node[c] = (int)e.first;
    fv_start[c] = pos;
    for (unsigned i : e.second) fv_feature[pos++] = (int)i;
    ++c;
  }
  fv_start[c] = pos;
  *fv_count = c;
}

void orb_oracle_distinctive_descriptors(const uint8_t* desc, const int64_t* offsets, int points, int32_t* best_idx,
                                        int32_t* best_median, int nthreads) {
  auto work = [&](int p0, int p1) {
    std::vector<std::vector<float>> D;
    for (int p = p0; p < p1; ++p) {
      const uint8_t* d = desc + offsets[p] * 32;
      const size_t N = (size_t)(offsets[p + 1] - offsets[p]);
      if (N == 0) { best_idx[p] = -1; if (best_median) best_median[p] = -1; continue; }   // "observations.empty() -> return" (:434-435)
      D.assign(N, std::vector<float>(N, 0));   // the reference keeps the distances as float (:455-456)
      for (size_t i = 0; i < N; ++i) {
        D[i][i] = 0;
        for (size_t j = i + 1; j < N; ++j) {
          const int dij = orb_oracle_descriptor_distance(d + i * 32, d + j * 32);
          D[i][j] = dij; D[j][i] = dij;
        }
      }
      int best_med = INT_MAX, best = 0;   // :474-475
      for (size_t i = 0; i < N; ++i) {
        std::vector<int> v(D[i].begin(), D[i].end());
        std::sort(v.begin(), v.end());
        const int median = v[0.5 * (N - 1)];   // :483
        if (median < best_med) { best_med = median; best = (int)i; }
      }
      best_idx[p] = best;
      if (best_median) best_median[p] = best_med;
    }
  };
  nthreads = std::max(1, std::min(nthreads, points));
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back(work, (int)((int64_t)points * t / nthreads), (int)((int64_t)points * (t + 1) / nthreads));
  for (auto& t : th) t.join();
}

int64_t orb_oracle_extract_many(int nf, float sf, int nl, int ini, int mn, int flags, const uint8_t* frames, int width,
                                int height, int n, int nthreads) {
  nthreads = std::max(1, std::min(nthreads, n));
  std::vector<int64_t> totals(nthreads, 0);
  auto work = [&](int t) {
    Extractor ex(nf, sf, nl, ini, mn, flags);
    std::vector<KeyPoint> k;
    std::vector<uint8_t> d;
    for (int f = t; f < n; f += nthreads) {
      Img im{const_cast<uint8_t*>(frames) + (size_t)f * width * height, width, height, width};
      totals[t] += ex.run(im, k, d);
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
  for (auto& t : th) t.join();
  int64_t s = 0;
  for (int64_t v : totals) s += v;
  return s;
}

}  // extern "C"
