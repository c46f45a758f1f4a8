"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/liborb_oracle.so (the CPU restatement).

May be imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs. The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
KNN_DTYPE = np.dtype([("d1", "<i4"), ("d2", "<i4"), ("i1", "<i8"), ("i2", "<i8")])

UPSTREAM_PYRAMID, LEGACY_BLUR, LIBM_TRIG = 1, 2, 4


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liborb_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("orb_oracle.cpp", "cv_primitives.cpp", "orb_oracle.h", "cv_primitives.h")]
    if force or not os.path.exists(so) or any(os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liborb_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        u8p, i32p, f32p, vp = C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float), C.c_void_p
        L.orb_oracle_create.restype = vp
        L.orb_oracle_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orb_oracle_destroy.argtypes = [vp]
        L.orb_oracle_get_table.argtypes = [vp, C.c_int, vp]
        L.orb_oracle_get_quota.argtypes = [vp, vp]
        L.orb_oracle_get_umax.argtypes = [vp, vp]
        L.orb_oracle_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, vp]
        L.orb_oracle_level_dims.argtypes = [vp, C.c_int, i32p, i32p]
        L.orb_oracle_level_ptr.restype = vp
        L.orb_oracle_level_ptr.argtypes = [vp, C.c_int]
        L.orb_oracle_blurred_ptr.restype = vp
        L.orb_oracle_blurred_ptr.argtypes = [vp, C.c_int]
        L.orb_oracle_raw_count.argtypes = [vp, C.c_int]
        L.orb_oracle_raw_keypoints.argtypes = [vp, C.c_int, vp, C.c_int]
        L.orb_oracle_resize.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int]
        L.orb_oracle_border101.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int]
        L.orb_oracle_blur.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int]
        L.orb_oracle_fast.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orb_oracle_fast_score_map.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp]
        L.orb_oracle_fast_atan2.restype = C.c_float
        L.orb_oracle_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orb_oracle_sincos.argtypes = [C.c_float, C.c_int, f32p, f32p]
        L.orb_oracle_distribute.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orb_oracle_descriptor_distance.argtypes = [vp, vp]
        L.orb_oracle_hamming_knn.argtypes = [vp, C.c_int, vp, C.c_int64, C.c_int64, C.c_int, vp, C.c_int]
        L.orb_oracle_stereo_match.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, C.c_float, vp, vp]
        L.orb_oracle_extract_many.restype = C.c_int64
        L.orb_oracle_extract_many.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int,
                                              C.c_int, C.c_int, C.c_int]
        _LIB = L
    return _LIB


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class OracleExtractor:
    """Mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:104-185) over the CPU oracle."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, flags=0):
        self.L = lib()
        self.h = self.L.orb_oracle_create(nfeatures, scale_factor, nlevels, ini_th, min_th, flags)
        if not self.h:
            raise ValueError("bad ORB parameters")
        self.nlevels, self.nfeatures, self.flags = nlevels, nfeatures, flags

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orb_oracle_destroy(self.h)
            self.h = None

    def table(self, what: int) -> np.ndarray:
        out = np.zeros(self.nlevels, np.float32)
        self.L.orb_oracle_get_table(self.h, what, _p(out))
        return out

    def quota(self) -> np.ndarray:
        out = np.zeros(self.nlevels, np.int32)
        self.L.orb_oracle_get_quota(self.h, _p(out))
        return out

    def umax(self) -> np.ndarray:
        out = np.zeros(16, np.int32)
        self.L.orb_oracle_get_umax(self.h, _p(out))
        return out

    def __call__(self, img: np.ndarray):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        cap = self.nfeatures + 64 * self.nlevels + 1024
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        lc = np.zeros(self.nlevels, np.int32)
        n = self.L.orb_oracle_extract(self.h, _p(img), img.strides[0], img.shape[1], img.shape[0], _p(kps), _p(desc),
                                      cap, _p(lc))
        if n < 0:
            raise RuntimeError("oracle extract failed")
        return kps[:n].copy(), desc[:n].copy(), lc

    def level(self, l: int, blurred: bool = False):
        c, r = C.c_int32(), C.c_int32()
        self.L.orb_oracle_level_dims(self.h, l, C.byref(c), C.byref(r))
        ptr = (self.L.orb_oracle_blurred_ptr if blurred else self.L.orb_oracle_level_ptr)(self.h, l)
        if not ptr:
            return None
        if blurred and self.flags & UPSTREAM_PYRAMID:   # the blurred copy is made of mvImagePyramid[l] = the interior of the stored padded buffer
            c.value -= 38
            r.value -= 38
        buf = (C.c_uint8 * (c.value * r.value)).from_address(ptr)
        return np.frombuffer(buf, np.uint8).reshape(r.value, c.value).copy()

    def raw_keypoints(self, l: int) -> np.ndarray:
        n = self.L.orb_oracle_raw_count(self.h, l)
        out = np.zeros(max(n, 1), KP_DTYPE)
        self.L.orb_oracle_raw_keypoints(self.h, l, _p(out), max(n, 1))
        return out[:n]


def resize(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    dst = np.zeros((dh, dw), np.uint8)
    lib().orb_oracle_resize(_p(src), src.strides[0], src.shape[1], src.shape[0], _p(dst), dw, dw, dh)
    return dst


def border101(src: np.ndarray, b: int = 19) -> np.ndarray:
    dst = np.zeros((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib().orb_oracle_border101(_p(src), src.strides[0], src.shape[1], src.shape[0], _p(dst), dst.shape[1], b)
    return dst


def blur(src: np.ndarray, legacy: bool = False) -> np.ndarray:
    dst = np.zeros_like(src, order="C")
    lib().orb_oracle_blur(_p(src), src.strides[0], src.shape[1], src.shape[0], _p(dst), dst.shape[1], int(legacy))
    return dst


def fast(img: np.ndarray, threshold: int) -> np.ndarray:
    cap = img.size
    out = np.zeros(max(cap, 1), KP_DTYPE)
    n = lib().orb_oracle_fast(_p(img), img.strides[0], img.shape[1], img.shape[0], threshold, _p(out), cap)
    return out[:n].copy()


def fast_score_map(img: np.ndarray) -> np.ndarray:
    out = np.zeros(img.shape, np.uint8)
    lib().orb_oracle_fast_score_map(_p(img), img.strides[0], img.shape[1], img.shape[0], _p(out))
    return out


def fast_atan2(y: float, x: float) -> float:
    return lib().orb_oracle_fast_atan2(y, x)


def sincos(angle: float, libm: bool = False):
    s, c = C.c_float(), C.c_float()
    lib().orb_oracle_sincos(angle, int(libm), C.byref(s), C.byref(c))
    return s.value, c.value


def distribute(kps: np.ndarray, minx, maxx, miny, maxy, nfeat) -> np.ndarray:
    kps = np.ascontiguousarray(kps)
    out = np.zeros(len(kps) + 8, KP_DTYPE)
    n = lib().orb_oracle_distribute(_p(kps), len(kps), minx, maxx, miny, maxy, nfeat, _p(out), len(out))
    return out[:n].copy()


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().orb_oracle_descriptor_distance(_p(a), _p(b))


def three_maxima(sizes: np.ndarray) -> np.ndarray:
    """ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2026-2068) on bin sizes [L] -> (ind1, ind2, ind3)."""
    L = lib()
    L.orb_oracle_three_maxima.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.orb_oracle_three_maxima.restype = None
    sizes = np.ascontiguousarray(sizes, np.int32)
    out = np.zeros(3, np.int32)
    L.orb_oracle_three_maxima(_p(sizes), len(sizes), _p(out))
    return out


def hamming_knn(q: np.ndarray, db: np.ndarray, index_base: int = 0, sentinel: int = 256, nthreads: int = 1) -> np.ndarray:
    q = np.ascontiguousarray(q, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    out = np.zeros(len(q), KNN_DTYPE)
    lib().orb_oracle_hamming_knn(_p(q), len(q), _p(db), len(db), index_base, sentinel, _p(out), nthreads)
    return out


def extract_many(frames: np.ndarray, nfeatures, nthreads, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, flags=0) -> int:
    frames = np.ascontiguousarray(frames, np.uint8)
    n, h, w = frames.shape
    return lib().orb_oracle_extract_many(nfeatures, scale_factor, nlevels, ini_th, min_th, flags, _p(frames), w, h, n,
                                         nthreads)


def stereo_match(left: "OracleExtractor", right: "OracleExtractor", kl, dl, kr, dr, bf: float, b: float):
    """Frame::ComputeStereoMatches (src/Frame.cc:880-1176) on the outputs and pyramids of the LAST call of `left` / `right`.
    Returns (uright[nL], depth[nL], kept)."""
    kl, dl = np.ascontiguousarray(kl), np.ascontiguousarray(dl)
    kr, dr = np.ascontiguousarray(kr), np.ascontiguousarray(dr)
    ur = np.empty(len(kl), np.float32)
    dp = np.empty(len(kl), np.float32)
    kept = lib().orb_oracle_stereo_match(left.h, right.h, _p(kl), _p(dl), len(kl), _p(kr), _p(dr), len(kr), bf, b, _p(ur), _p(dp))
    return ur, dp, kept


QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("ur", "<f4"), ("angle", "<f4"),
                        ("min_level", "<i2"), ("max_level", "<i2"), ("flags", "<i4"), ("reserved", "<i4")])


class _SearchParams(C.Structure):
    _fields_ = [("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("th_dist", C.c_int32), ("nn_ratio", C.c_float), ("check_orientation", C.c_int32), ("reserved", C.c_int32)]


def search_by_projection(kps_un, desc, queries, qdesc, bounds, th_dist=100, nn_ratio=0.0, check_orientation=True, uright=None,
                         claimed=None):
    """One (query set, target frame) pair through the restated guided search (src/ORBmatcher.cc:69-184, 1702-1866;
    src/Frame.cc:432-460, 633-745). Returns (assigned [n], query_match [m], nmatches)."""
    L = lib()
    L.orb_oracle_search_by_projection.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(_SearchParams),
                                                                  C.c_void_p, C.c_void_p]
    kps_un, desc = np.ascontiguousarray(kps_un), np.ascontiguousarray(desc)
    queries, qdesc = np.ascontiguousarray(queries), np.ascontiguousarray(qdesc)
    assert queries.dtype == QUERY_DTYPE and kps_un.dtype.itemsize == 28
    n, m = len(kps_un), len(queries)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    cl = None if claimed is None else np.ascontiguousarray(claimed, np.uint8)
    prm = _SearchParams(*[float(b) for b in bounds], int(th_dist), float(nn_ratio), int(check_orientation), 0)
    assigned = np.full(max(n, 1), -1, np.int32)
    qmatch = np.full(max(m, 1), -1, np.int32)
    nm = L.orb_oracle_search_by_projection(_p(kps_un), _p(desc), None if ur is None else _p(ur), None if cl is None else _p(cl), n,
                                           _p(queries), _p(qdesc), m, C.byref(prm), _p(assigned), _p(qmatch))
    return assigned[:n], qmatch[:m], nm


def fuse_search(kps_un, desc, queries, qdesc, bounds, th_dist=50, uright=None, inv_level_sigma2=None):
    """Candidate search of ORBmatcher::Fuse / SearchBySim3 (src/ORBmatcher.cc:1100-1700), one keyframe. Returns (best_idx [m], best_dist [m])."""
    L = lib()
    L.orb_oracle_fuse_search.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(_SearchParams), C.c_void_p,
                                                         C.c_void_p, C.c_void_p]
    L.orb_oracle_fuse_search.restype = None
    kps_un, desc = np.ascontiguousarray(kps_un), np.ascontiguousarray(desc)
    queries, qdesc = np.ascontiguousarray(queries), np.ascontiguousarray(qdesc)
    assert queries.dtype == QUERY_DTYPE and kps_un.dtype.itemsize == 28
    n, m = len(kps_un), len(queries)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    sig = None
    if inv_level_sigma2 is not None:
        sig = np.zeros(16, np.float32)
        sig[:len(inv_level_sigma2)] = inv_level_sigma2
    prm = _SearchParams(*[float(b) for b in bounds], int(th_dist), 0.0, 0, 0)
    bi, bd = np.full(max(m, 1), -1, np.int32), np.full(max(m, 1), 256, np.int32)
    L.orb_oracle_fuse_search(_p(kps_un), _p(desc), None if ur is None else _p(ur), n, _p(queries), _p(qdesc), m, C.byref(prm),
                             None if sig is None else _p(sig), _p(bi), _p(bd))
    return bi[:m], bd[:m]


def features_in_area(kps_un, bounds, x, y, r, min_level=-1, max_level=-1):
    """Frame::GetFeaturesInArea (src/Frame.cc:633-717) with its scan order."""
    L = lib()
    L.orb_oracle_features_in_area.argtypes = [C.c_void_p, C.c_int, C.POINTER(_SearchParams), C.c_float, C.c_float, C.c_float, C.c_int,
                                              C.c_int, C.c_void_p]
    kps_un = np.ascontiguousarray(kps_un)
    prm = _SearchParams(*[float(b) for b in bounds], 0, 0.0, 0, 0)
    out = np.empty(max(len(kps_un), 1), np.int32)
    n = L.orb_oracle_features_in_area(_p(kps_un), len(kps_un), C.byref(prm), x, y, r, min_level, max_level, _p(out))
    return out[:n].copy()


def distinctive_descriptors(desc, offsets, nthreads=1):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517), batched. Returns (best_idx, best_median)."""
    L = lib()
    L.orb_oracle_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    L.orb_oracle_distinctive_descriptors.restype = None
    desc = np.ascontiguousarray(desc, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.int64)
    P = len(offsets) - 1
    bi, bm = np.empty(max(P, 1), np.int32), np.empty(max(P, 1), np.int32)
    if P > 0:
        L.orb_oracle_distinctive_descriptors(_p(desc), _p(offsets), P, _p(bi), _p(bm), nthreads)
    return bi[:P], bm[:P]


class OracleVocabulary:
    """DBoW2 vocabulary restatement (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) — transform() for one descriptor set."""

    def __init__(self, k, L, parent, is_leaf, desc, weight):
        L_ = lib()
        L_.orb_oracle_vocabulary_create.restype = C.c_void_p
        L_.orb_oracle_vocabulary_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L_.orb_oracle_vocabulary_destroy.argtypes = [C.c_void_p]
        L_.orb_oracle_vocabulary_destroy.restype = None
        L_.orb_oracle_bow_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 9
        L_.orb_oracle_bow_transform.restype = None
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8); weight = np.ascontiguousarray(weight, np.float64)
        self.h = C.c_void_p(L_.orb_oracle_vocabulary_create(k, L, len(parent), _p(parent), _p(is_leaf), _p(desc), _p(weight)))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orb_oracle_vocabulary_destroy(self.h)
            self.h = None

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        m = max(n, 1)
        word, node = np.empty(m, np.int32), np.empty(m, np.int32)
        bw, bv, bc = np.empty(m, np.int32), np.empty(m, np.float64), np.zeros(1, np.int32)
        fn, fs, ff, fc = np.empty(m, np.int32), np.empty(m + 1, np.int32), np.empty(m, np.int32), np.zeros(1, np.int32)
        lib().orb_oracle_bow_transform(self.h, _p(desc), n, levelsup, _p(word), _p(node), _p(bw), _p(bv), _p(bc), _p(fn), _p(fs), _p(ff), _p(fc))
        b, f = int(bc[0]), int(fc[0])
        bow = {int(w): float(v) for w, v in zip(bw[:b], bv[:b])}
        fv = {int(fn[j]): ff[fs[j]:fs[j + 1]].tolist() for j in range(f)}
        return word[:n], node[:n], bow, fv


def undistort_points(xy, fx, fy, cx, cy, dist):
    """cv::undistortPoints(pts, pts, K, dist, Mat(), K) as Frame::UndistortKeyPoints calls it (src/Frame.cc:787-793). xy [n,2] float32."""
    L = lib()
    L.orb_oracle_undistort_points.argtypes = [C.c_float] * 4 + [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    L.orb_oracle_undistort_points.restype = None
    out = np.ascontiguousarray(xy, np.float32).copy()
    d = np.ascontiguousarray(dist, np.float32)
    L.orb_oracle_undistort_points(fx, fy, cx, cy, _p(d), len(d), _p(out), len(out))
    return out


def search_by_bow(kf_desc, kf_kps_un, kf_valid, kf_fv, f_desc, f_kps, f_fv, th_dist=50, nn_ratio=0.7, check_orientation=True, f_blocked=None):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (src/ORBmatcher.cc:248-410); with f_blocked and th_dist = TH_LOW - 1 the
    KeyFrame-KeyFrame form (:722-880). kf_fv / f_fv = (node, start, feature, count) CSR FeatureVectors.
    Returns (assigned [n_f], nmatches, kf_match [n_kf])."""
    L = lib()
    vp = C.c_void_p
    L.orb_oracle_search_by_bow.argtypes = [vp] * 6 + [C.c_int, vp, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_float, C.c_int, vp, vp, vp, C.c_int]
    c = lambda a, dt: np.ascontiguousarray(a, dt)
    kf_desc, f_desc = c(kf_desc, np.uint8), c(f_desc, np.uint8)
    kf_kps_un, f_kps = np.ascontiguousarray(kf_kps_un), np.ascontiguousarray(f_kps)
    kf_valid = c(kf_valid, np.uint8)
    kn, ks, kfe, kc = c(kf_fv[0], np.int32), c(kf_fv[1], np.int32), c(kf_fv[2], np.int32), int(kf_fv[3])
    fn, fs, ffe, fc = c(f_fv[0], np.int32), c(f_fv[1], np.int32), c(f_fv[2], np.int32), int(f_fv[3])
    n_f, n_kf = len(f_kps), len(kf_kps_un)
    fb = None if f_blocked is None else c(f_blocked, np.uint8)
    assigned = np.full(max(n_f, 1), -1, np.int32)
    kf_match = np.full(max(n_kf, 1), -1, np.int32)
    nm = L.orb_oracle_search_by_bow(_p(kf_desc), _p(kf_kps_un), _p(kf_valid), _p(kn), _p(ks), _p(kfe), kc, _p(f_desc), _p(f_kps), n_f,
                                    _p(fn), _p(fs), _p(ffe), fc, th_dist, nn_ratio, int(check_orientation), _p(assigned),
                                    None if fb is None else _p(fb), _p(kf_match), n_kf)
    return assigned[:n_f], nm, kf_match[:n_kf]


def search_for_triangulation(desc1, kps1_un, has_mp1, uright1, fv1, desc2, kps2_un, has_mp2, uright2, fv2, f12, ex, ey, level_sigma2, scale_factor,
                             only_stereo=False, th_low=50, check_orientation=True):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:879-1090), one keyframe pair; fv? = (node, start, feature, count) CSR.
    Returns (vMatches12 [n1], nmatches)."""
    L = lib()
    vp, ci, cf = C.c_void_p, C.c_int, C.c_float
    L.orb_oracle_search_for_triangulation.argtypes = ([vp] * 4 + [ci] + [vp] * 3 + [ci]) * 2 + [vp, cf, cf, vp, vp, ci, ci, ci, vp]
    c = lambda a, dt: np.ascontiguousarray(a, dt)
    d1, d2 = c(desc1, np.uint8), c(desc2, np.uint8)
    k1, k2 = np.ascontiguousarray(kps1_un), np.ascontiguousarray(kps2_un)
    h1, h2, u1, u2 = c(has_mp1, np.uint8), c(has_mp2, np.uint8), c(uright1, np.float32), c(uright2, np.float32)
    a1 = [c(fv1[0], np.int32), c(fv1[1], np.int32), c(fv1[2], np.int32)]
    a2 = [c(fv2[0], np.int32), c(fv2[1], np.int32), c(fv2[2], np.int32)]
    F = c(np.asarray(f12).reshape(9), np.float32)
    ls, sf = c(level_sigma2, np.float32), c(scale_factor, np.float32)
    n1, n2 = len(k1), len(k2)
    m12 = np.full(max(n1, 1), -1, np.int32)
    nm = L.orb_oracle_search_for_triangulation(_p(d1), _p(k1), _p(h1), _p(u1), n1, _p(a1[0]), _p(a1[1]), _p(a1[2]), int(fv1[3]),
                                               _p(d2), _p(k2), _p(h2), _p(u2), n2, _p(a2[0]), _p(a2[1]), _p(a2[2]), int(fv2[3]),
                                               _p(F), float(ex), float(ey), _p(ls), _p(sf), int(only_stereo), int(th_low), int(check_orientation),
                                               _p(m12))
    return m12[:n1], nm


def search_for_initialization(kps_un, desc, queries, qdesc, bounds, th_dist=50, nn_ratio=0.9, check_orientation=True):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:570-712), one frame pair. Returns (matches12 [m], matches21 [n], nmatches)."""
    L = lib()
    L.orb_oracle_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(_SearchParams),
                                                       C.c_void_p, C.c_void_p]
    kps_un, desc = np.ascontiguousarray(kps_un), np.ascontiguousarray(desc)
    queries, qdesc = np.ascontiguousarray(queries), np.ascontiguousarray(qdesc)
    n, m = len(kps_un), len(queries)
    prm = _SearchParams(*[float(b) for b in bounds], int(th_dist), float(nn_ratio), int(check_orientation), 0)
    m12, m21 = np.full(max(m, 1), -1, np.int32), np.full(max(n, 1), -1, np.int32)
    nm = L.orb_oracle_search_for_initialization(_p(kps_un), _p(desc), n, _p(queries), _p(qdesc), m, C.byref(prm), _p(m12), _p(m21))
    return m12[:m], m21[:n], nm
