"""Runs oracle/_ref/orb_ref_extract: the reference's own src/ORBextractor.cc compiled unmodified against oracle/cvshim
(built by `make -C oracle _ref` where /root/reference exists; the binary is git-ignored and travels with the snapshot).
Test infrastructure only."""
import os
import struct
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "orb_ref_extract")
BIN_UPSTREAM = os.path.join(ROOT, "oracle", "_ref", "orb_ref_extract_upstream")   # same source minus src/ORBextractor.cc:2166
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
                     ("class_id", "<i4")])


def available() -> bool:
    return os.access(BIN, os.X_OK)


def upstream_available() -> bool:
    return os.access(BIN_UPSTREAM, os.X_OK)


def ref_extract(frames: np.ndarray, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, legacy_blur=False, upstream=False):
    """frames: [n, H, W] u8 -> (tables [4, nlevels] f32, list of (kps, desc, [mvImagePyramid level images])).
    upstream=True runs the build without the fork's `mvImagePyramid[level] = temp;` (raulmur's un-padded pyramid)."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, H, W = frames.shape
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<8if", 0x4F524231, n, W, H, nfeatures, nlevels, ini_th, min_th, scale_factor))
            f.write(frames.tobytes())
        env = dict(os.environ)
        env.pop("ORB_REF_LEGACY_BLUR", None)
        if legacy_blur:
            env["ORB_REF_LEGACY_BLUR"] = "1"
        subprocess.run([BIN_UPSTREAM if upstream else BIN, fin, fout], check=True, env=env, timeout=600)
        buf = open(fout, "rb").read()
    pos = 0
    nl = struct.unpack_from("<i", buf, pos)[0]; pos += 4
    tables = np.frombuffer(buf, "<f4", 4 * nl, pos).reshape(4, nl).copy(); pos += 16 * nl
    out = []
    for _ in range(n):
        N = struct.unpack_from("<i", buf, pos)[0]; pos += 4
        kps = np.frombuffer(buf, KP_DTYPE, N, pos).copy(); pos += 28 * N
        desc = np.frombuffer(buf, np.uint8, 32 * N, pos).reshape(N, 32).copy(); pos += 32 * N
        levels = []
        for _l in range(nl):
            c, r = struct.unpack_from("<2i", buf, pos); pos += 8
            levels.append(np.frombuffer(buf, np.uint8, r * c, pos).reshape(r, c).copy()); pos += r * c
        out.append((kps, desc, levels))
    assert pos == len(buf)
    return tables, out


def write_input(path, frames, nfeatures, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, H, W = frames.shape
    with open(path, "wb") as f:
        f.write(struct.pack("<8if", 0x4F524231, n, W, H, nfeatures, nlevels, ini_th, min_th, scale_factor))
        f.write(frames.tobytes())


BOW_BIN = os.path.join(ROOT, "oracle", "_ref", "orb_ref_bow")


def bow_available() -> bool:
    return os.access(BOW_BIN, os.X_OK)


def ref_bow(voc_path: str, frames, levelsup: int = 4):
    """The reference's own DBoW2: loadFromTextFile(voc_path) + transform(). frames: list of [n, 32] u8 arrays ->
    list of (BowVector {word: value}, FeatureVector {node: [feature indices]})."""
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "desc.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<2i", len(frames), levelsup))
            for a in frames:
                a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
                f.write(struct.pack("<i", len(a)))
                f.write(a.tobytes())
        subprocess.run([BOW_BIN, voc_path, fin, fout], check=True, timeout=600)
        buf = open(fout, "rb").read()
    pos, out = 0, []
    for _ in frames:
        nb = struct.unpack_from("<i", buf, pos)[0]; pos += 4
        bow = {}
        for _j in range(nb):
            w, v = struct.unpack_from("<Id", buf, pos); pos += 12
            bow[w] = v
        nf = struct.unpack_from("<i", buf, pos)[0]; pos += 4
        fv = {}
        for _j in range(nf):
            node, m = struct.unpack_from("<Ii", buf, pos); pos += 8
            fv[node] = list(struct.unpack_from(f"<{m}I", buf, pos)); pos += 4 * m
        out.append((bow, fv))
    assert pos == len(buf)
    return out


# ---- oracle/_ref/orb_ref_match: function bodies cut out of src/ORBmatcher.cc, src/MapPoint.cc, src/Frame.cc by line range at build time
# (oracle/Makefile `_ref/gen_stamp`) and compiled inside oracle/ref_match_main.cpp, plus DBoW2's FORB::distance compiled unmodified.
MATCH_BIN = os.path.join(ROOT, "oracle", "_ref", "orb_ref_match")


def match_available() -> bool:
    return os.access(MATCH_BIN, os.X_OK)


def _run_match(mode: str, payload: bytes) -> bytes:
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(payload)
        subprocess.run([MATCH_BIN, mode, fin, fout], check=True, timeout=600)
        return open(fout, "rb").read()


def ref_distance(a: np.ndarray, b: np.ndarray):
    """n descriptor pairs -> (FORB::distance [n], ORBmatcher::DescriptorDistance [n]) by the reference's own code."""
    a, b = np.ascontiguousarray(a, np.uint8).reshape(-1, 32), np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    n = len(a)
    inter = np.concatenate([a, b], axis=1)   # a[i], b[i] interleaved
    out = np.frombuffer(_run_match("distance", struct.pack("<i", n) + inter.tobytes()), "<i4").reshape(n, 2)
    return out[:, 0].copy(), out[:, 1].copy()


def ref_scan(q: np.ndarray, db: np.ndarray, ratio: float = 0.6, matched_distance=None):
    """The reference's best / second-best loop over every database row in index order. matched_distance=None: SearchByBoW form
    (src/ORBmatcher.cc:304-343, sentinel 256, accept `(float)best < ratio * (float)second`); otherwise the SearchForInitialization form
    (:611-645, sentinel INT_MAX, rows with vMatchedDistance[i] <= dist skipped, accept `best < (float)second * ratio`).
    Returns int32 [Q, 4]: best, second, best index (-1 = none), accepted."""
    q, db = np.ascontiguousarray(q, np.uint8).reshape(-1, 32), np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
    payload = struct.pack("<iif", len(q), len(db), ratio) + q.tobytes() + db.tobytes()
    mode = "scan"
    if matched_distance is not None:
        payload += np.ascontiguousarray(matched_distance, "<i4").tobytes()
        mode = "scan_init"
    return np.frombuffer(_run_match(mode, payload), "<i4").reshape(len(q), 4).copy()


def ref_three_maxima(sizes: np.ndarray) -> np.ndarray:
    """sizes int32 [n, L] (bin populations) -> int32 [n, 3] (ind1, ind2, ind3) by ORBmatcher::ComputeThreeMaxima."""
    sizes = np.ascontiguousarray(sizes, "<i4")
    n, L = sizes.shape
    return np.frombuffer(_run_match("three_maxima", struct.pack("<ii", n, L) + sizes.tobytes()), "<i4").reshape(n, 3).copy()


def ref_distinctive(desc: np.ndarray, offsets: np.ndarray) -> np.ndarray:
    """CSR of observation descriptors -> int32 [P, 2] (BestIdx, BestMedian) by the loop of MapPoint::ComputeDistinctiveDescriptors."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    P = len(offsets) - 1
    payload = struct.pack("<i", P)
    for p in range(P):
        rows = desc[offsets[p]:offsets[p + 1]]
        payload += struct.pack("<i", len(rows)) + rows.tobytes()
    return np.frombuffer(_run_match("distinctive", payload), "<i4").reshape(P, 2).copy()


def ref_stereo(scale, levels_left, levels_right, kl, dl, kr, dr, bf: float, b: float):
    """Frame::ComputeStereoMatches (src/Frame.cc:880-1176) itself on padded pyramid levels, keypoints (28-byte cv::KeyPoint records) and
    descriptors of both images -> (mvuRight [nL], mvDepth [nL])."""
    nl = len(scale)
    payload = struct.pack("<iff", nl, bf, b) + np.ascontiguousarray(scale, "<f4").tobytes()
    for levels, k, d in ((levels_left, kl, dl), (levels_right, kr, dr)):
        for im in levels:
            im = np.ascontiguousarray(im, np.uint8)
            payload += struct.pack("<ii", im.shape[1], im.shape[0]) + im.tobytes()
        k = np.ascontiguousarray(k)
        assert k.dtype.itemsize == 28
        payload += struct.pack("<i", len(k)) + k.tobytes() + np.ascontiguousarray(d, np.uint8).tobytes()
    out = np.frombuffer(_run_match("stereo", payload), "<f4")
    n = len(kl)
    return out[:n].copy(), out[n:2 * n].copy()


def ref_search_by_projection(kps_un, desc, uright, claimed, bounds, scale, th, nnratio, proj, qdesc):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) (src/ORBmatcher.cc:69-184) with RadiusByViewingCos, on a Frame whose grid is
    built by the reference's AssignFeaturesToGrid / PosInGrid and searched by its GetFeaturesInArea (src/Frame.cc:432-460, 729-745, 633-717).
    proj: structured array (projx, projy, projxr, viewcos: f4; level, flags: i4; flags 1 = !mbTrackInView, 2 = Observations() > 0).
    Returns (nmatches, assigned [n]: index of the map point this call assigned to the keypoint, -1 none)."""
    n, m = len(kps_un), len(proj)
    kps_un = np.ascontiguousarray(kps_un)
    assert kps_un.dtype.itemsize == 28 and proj.dtype.itemsize == 24
    payload = struct.pack("<4f2fi", *[float(b) for b in bounds], th, nnratio, len(scale)) + np.ascontiguousarray(scale, "<f4").tobytes()
    payload += struct.pack("<i", n) + kps_un.tobytes() + np.ascontiguousarray(desc, np.uint8).tobytes()
    payload += np.ascontiguousarray(uright if uright is not None else np.full(n, -1, np.float32), "<f4").tobytes()
    payload += np.ascontiguousarray(claimed if claimed is not None else np.zeros(n, np.uint8), np.uint8).tobytes()
    payload += struct.pack("<i", m) + np.ascontiguousarray(proj).tobytes() + np.ascontiguousarray(qdesc, np.uint8).tobytes()
    out = np.frombuffer(_run_match("sbp", payload), "<i4")
    return int(out[0]), out[1:1 + n].copy()


LAST_DTYPE = np.dtype([("X", "<f4"), ("Y", "<f4"), ("Z", "<f4"), ("angle", "<f4"), ("octave", "<i4"), ("flags", "<i4")])


def ref_search_by_projection_frames(kps_un, desc, uright, claimed, bounds, scale, th, cam, mb, mbf, last_tz, mono, check_orientation, last, qdesc):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1702-1871, the TrackWithMotionModel matcher) with
    CurrentFrame.mTcw = identity and LastFrame.mTcw = translation (0, 0, last_tz). cam = (fx, fy, cx, cy). last: LAST_DTYPE per last-frame keypoint
    (world point, angle of its keypoint, octave, flags 1 = no map point, 2 = Observations() > 0, 4 = outlier).
    Returns (nmatches, assigned [n]: last-frame index whose map point the current keypoint got by this call, -1 none)."""
    n, m = len(kps_un), len(last)
    kps_un = np.ascontiguousarray(kps_un)
    assert kps_un.dtype.itemsize == 28 and last.dtype == LAST_DTYPE
    payload = struct.pack("<4f2f2f4ff3i", *[float(b) for b in bounds], th, 0.0, mb, mbf, *[float(c) for c in cam], last_tz, int(mono),
                          int(check_orientation), len(scale)) + np.ascontiguousarray(scale, "<f4").tobytes()
    payload += struct.pack("<i", n) + kps_un.tobytes() + np.ascontiguousarray(desc, np.uint8).tobytes()
    payload += np.ascontiguousarray(uright if uright is not None else np.full(n, -1, np.float32), "<f4").tobytes()
    payload += np.ascontiguousarray(claimed if claimed is not None else np.zeros(n, np.uint8), np.uint8).tobytes()
    payload += struct.pack("<i", m) + np.ascontiguousarray(last).tobytes() + np.ascontiguousarray(qdesc, np.uint8).tobytes()
    out = np.frombuffer(_run_match("sbpf", payload), "<i4")
    return int(out[0]), out[1:1 + n].copy()


def _fv_bytes(fv: dict) -> bytes:
    b = struct.pack("<i", len(fv))
    for node in sorted(fv):
        idx = np.asarray(fv[node], "<u4")
        b += struct.pack("<Ii", node, len(idx)) + idx.tobytes()
    return b


def ref_search_by_bow(kf_desc, kf_kps_un, kf_valid, kf_fv: dict, f_desc, f_kps, f_fv: dict, nnratio: float, check_orientation: bool):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (src/ORBmatcher.cc:248-410). kf_fv / f_fv: {node: [feature indices]}.
    Returns (nmatches, assigned [n_f]: keyframe feature whose map point the frame keypoint got, -1 none)."""
    nkf, nf = len(kf_kps_un), len(f_kps)
    payload = struct.pack("<fi", nnratio, int(check_orientation))
    payload += struct.pack("<i", nkf) + np.ascontiguousarray(kf_desc, np.uint8).tobytes() + np.ascontiguousarray(kf_kps_un).tobytes()
    payload += np.ascontiguousarray(kf_valid, np.uint8).tobytes() + _fv_bytes(kf_fv)
    payload += struct.pack("<i", nf) + np.ascontiguousarray(f_desc, np.uint8).tobytes() + np.ascontiguousarray(f_kps).tobytes() + _fv_bytes(f_fv)
    out = np.frombuffer(_run_match("sbow", payload), "<i4")
    return int(out[0]), out[1:1 + nf].copy()


def ref_search_by_bow_keyframes(desc1, kps1_un, valid1, fv1: dict, desc2, kps2_un, valid2, fv2: dict, nnratio: float, check_orientation: bool):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (src/ORBmatcher.cc:724-867). Returns (nmatches, vpMatches12 [n1] as the index of the
    keyframe-2 feature whose map point was matched, -1 none)."""
    payload = struct.pack("<fi", nnratio, int(check_orientation))
    for d, k, v, fv in ((desc1, kps1_un, valid1, fv1), (desc2, kps2_un, valid2, fv2)):
        payload += struct.pack("<i", len(k)) + np.ascontiguousarray(d, np.uint8).tobytes() + np.ascontiguousarray(k).tobytes()
        payload += np.ascontiguousarray(v, np.uint8).tobytes() + _fv_bytes(fv)
    out = np.frombuffer(_run_match("sbow2", payload), "<i4")
    return int(out[0]), out[1:1 + len(kps1_un)].copy()


FUSE_DTYPE = np.dtype([("X", "<f4"), ("Y", "<f4"), ("Z", "<f4"), ("nz", "<f4"), ("level", "<i4"), ("flags", "<i4")])


def ref_fuse(kps_un, desc, uright, held, bounds, scale, inv_sigma2, th, cam, mbf, points, qdesc, sim3_form=False):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:1100-1270) on a keyframe with identity pose. points: FUSE_DTYPE (world point, z of the
    viewing normal (0, 0, nz), PredictScale result, flags 1 = null, 2 = isBad(), 4 = IsInKeyFrame(pKF)). Returns (nFused, bestIdx [m], -1 = not fused).
    sim3_form: Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1282-1419) with Scw = identity instead (flag 4 = among the keyframe's own map points)."""
    n, m = len(kps_un), len(points)
    kps_un = np.ascontiguousarray(kps_un)
    assert kps_un.dtype.itemsize == 28 and points.dtype == FUSE_DTYPE
    payload = struct.pack("<4ff5fi", *[float(b) for b in bounds], th, mbf, *[float(c) for c in cam], len(scale))
    payload += np.ascontiguousarray(scale, "<f4").tobytes() + np.ascontiguousarray(inv_sigma2, "<f4").tobytes()
    payload += struct.pack("<i", n) + kps_un.tobytes() + np.ascontiguousarray(desc, np.uint8).tobytes()
    payload += np.ascontiguousarray(uright if uright is not None else np.full(n, -1, np.float32), "<f4").tobytes()
    payload += np.ascontiguousarray(held if held is not None else np.zeros(n, np.uint8), np.uint8).tobytes()
    payload += struct.pack("<i", m) + np.ascontiguousarray(points).tobytes() + np.ascontiguousarray(qdesc, np.uint8).tobytes()
    out = np.frombuffer(_run_match("fuse2" if sim3_form else "fuse", payload), "<i4")
    return int(out[0]), out[1:1 + m].copy()


def ref_search_for_triangulation(desc1, kps1_un, has1, ur1, fv1: dict, desc2, kps2_un, has2, ur2, fv2: dict, f12, cw, cam2, level_sigma2, scale, only_stereo,
                                 check_orientation):
    """ORBmatcher::SearchForTriangulation + CheckDistEpipolarLine (src/ORBmatcher.cc:879-1088, 206-233); keyframe 2 has the identity pose, cw = camera
    centre of keyframe 1 (the reference derives the epipole from it). Returns (nmatches, vMatches12 [n1])."""
    payload = struct.pack("<3i", int(only_stereo), int(check_orientation), len(scale))
    payload += np.ascontiguousarray(scale, "<f4").tobytes() + np.ascontiguousarray(level_sigma2, "<f4")[:len(scale)].tobytes()
    payload += np.ascontiguousarray(f12, "<f4").reshape(9).tobytes() + np.ascontiguousarray(cw, "<f4").tobytes() + np.ascontiguousarray(cam2, "<f4").tobytes()
    for d, k, h, u, fv in ((desc1, kps1_un, has1, ur1, fv1), (desc2, kps2_un, has2, ur2, fv2)):
        payload += struct.pack("<i", len(k)) + np.ascontiguousarray(d, np.uint8).tobytes() + np.ascontiguousarray(k).tobytes()
        payload += np.ascontiguousarray(h, np.uint8).tobytes() + np.ascontiguousarray(u, "<f4").tobytes() + _fv_bytes(fv)
    out = np.frombuffer(_run_match("stri", payload), "<i4")
    return int(out[0]), out[1:1 + len(kps1_un)].copy()


KFPT_DTYPE = np.dtype([("X", "<f4"), ("Y", "<f4"), ("Z", "<f4"), ("angle", "<f4"), ("level", "<i4"), ("flags", "<i4")])


def ref_search_by_projection_keyframe(kps_un, desc, held, bounds, scale, th, cam, orbdist, check_orientation, points, qdesc, loop_form=False):
    """loop_form=False: SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1883-2015, relocalisation; mTcw = identity);
    loop_form=True: SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:422-551, loop closing; Scw = identity, int th, TH_LOW).
    points: KFPT_DTYPE (world point, angle of its keypoint in the source keyframe, PredictScale result, flags 1 null, 2 bad, 4 already found).
    Returns (nmatches, assigned [n]: point index per target keypoint, -1 none)."""
    n, m = len(kps_un), len(points)
    kps_un = np.ascontiguousarray(kps_un)
    assert kps_un.dtype.itemsize == 28 and points.dtype == KFPT_DTYPE
    payload = struct.pack("<4ff4f3i", *[float(b) for b in bounds], th, *[float(c) for c in cam], int(orbdist), int(check_orientation), len(scale))
    payload += np.ascontiguousarray(scale, "<f4").tobytes()
    payload += struct.pack("<i", n) + kps_un.tobytes() + np.ascontiguousarray(desc, np.uint8).tobytes()
    payload += np.ascontiguousarray(held if held is not None else np.zeros(n, np.uint8), np.uint8).tobytes()
    payload += struct.pack("<i", m) + np.ascontiguousarray(points).tobytes() + np.ascontiguousarray(qdesc, np.uint8).tobytes()
    out = np.frombuffer(_run_match("sbps" if loop_form else "sbpk", payload), "<i4")
    return int(out[0]), out[1:1 + n].copy()


SIM3_DTYPE = np.dtype([("X", "<f4"), ("Y", "<f4"), ("Z", "<f4"), ("level", "<i4"), ("flags", "<i4"), ("prematch", "<i4")])


def ref_search_by_sim3(bounds, th, cam, scale, kps1, desc1, pts1, kps2, desc2, pts2):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1433-1684) with identity poses, s12 = 1, R12 = I, t12 = 0. pts?: SIM3_DTYPE per feature (world point of
    the map point it holds, PredictScale result, flags 1 = no map point, 2 = bad; prematch (keyframe 1 only) = feature of keyframe 2 already matched).
    Returns (nFound, newly matched feature of keyframe 2 per feature of keyframe 1, -1 none)."""
    payload = struct.pack("<4ff4fi", *[float(b) for b in bounds], th, *[float(c) for c in cam], len(scale)) + np.ascontiguousarray(scale, "<f4").tobytes()
    for k, d, p in ((kps1, desc1, pts1), (kps2, desc2, pts2)):
        assert p.dtype == SIM3_DTYPE and len(p) == len(k)
        payload += struct.pack("<i", len(k)) + np.ascontiguousarray(k).tobytes() + np.ascontiguousarray(d, np.uint8).tobytes() + np.ascontiguousarray(p).tobytes()
    out = np.frombuffer(_run_match("sim3", payload), "<i4")
    return int(out[0]), out[1:1 + len(kps1)].copy()


def ref_search_for_initialization(kps2_un, desc2, kps1_un, desc1, prev_matched, bounds, nnratio, check_orientation, window):
    """ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:570-712) on F2's reference-built grid.
    Returns (nmatches, vnMatches12 [m])."""
    n, m = len(kps2_un), len(kps1_un)
    payload = struct.pack("<4ffii", *[float(b) for b in bounds], nnratio, int(check_orientation), int(window))
    payload += struct.pack("<i", n) + np.ascontiguousarray(kps2_un).tobytes() + np.ascontiguousarray(desc2, np.uint8).tobytes()
    payload += struct.pack("<i", m) + np.ascontiguousarray(kps1_un).tobytes() + np.ascontiguousarray(desc1, np.uint8).tobytes()
    payload += np.ascontiguousarray(prev_matched, "<f4").tobytes()
    out = np.frombuffer(_run_match("sinit", payload), "<i4")
    return int(out[0]), out[1:1 + m].copy()
