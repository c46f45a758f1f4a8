"""Pins the CPU oracle (oracle/) against the committed golden vectors under tests/golden/.

The goldens were produced by tests/golden/make_golden.py with the REAL OpenCV (cv2 4.13): every arithmetic primitive the
reference delegates to OpenCV (cv::resize, cv::copyMakeBorder, cv::FAST, cv::GaussianBlur, cv::fastAtan2; call sites
src/ORBextractor.cc:160, 1510, 1521, 2042, 2119, 2138, 2159) ran inside OpenCV itself; only the reference's control flow was
restated. cv2 is NOT needed to run these tests. Bar: bit-exact everywhere."""
import glob
import hashlib
import os

import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PIPELINES = sorted(p for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                   if os.path.basename(p) not in ("fast_atan2.npz", "primitives_333x257.npz") and not os.path.basename(p).startswith(("stereo_", "undistort_")))
STEREO = sorted(glob.glob(os.path.join(GOLDEN, "stereo_*.npz")))   # tests/golden/make_golden_stereo.py


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_golden_files_present():
    assert len(PIPELINES) >= 5
    assert os.path.exists(os.path.join(GOLDEN, "fast_atan2.npz"))
    assert os.path.exists(os.path.join(GOLDEN, "primitives_333x257.npz"))


@pytest.mark.parametrize("path", PIPELINES, ids=[os.path.basename(p)[:-4] for p in PIPELINES])
def test_pipeline_matches_cv2_composition(oracle, path):
    g = np.load(path)
    w, h, nf, nl, ini, mn, seed = [int(v) for v in g["params"]]
    img = synth.synth_frame(w, h, seed)
    assert sha(img) == str(g["image_sha"]), "synthetic generator drifted: regenerate goldens"
    oe = oracle.OracleExtractor(nf, float(g["scale"]), nl, ini, mn)
    kps, desc, lc = oe(img)
    for l in range(nl):
        assert sha(oe.level(l)) == str(g["level_sha"][l]), f"pyramid level {l}"
        raw = oe.raw_keypoints(l)
        assert len(raw) == int(g["raw_counts"][l]), f"raw FAST count level {l}"
        r3 = np.stack([raw["x"], raw["y"], raw["response"]], 1).astype(np.float32) if len(raw) else np.zeros((0, 3), np.float32)
        assert sha(r3) == str(g["raw_sha"][l]), f"raw FAST keypoints level {l}"
        b = oe.level(l, blurred=True)
        assert (sha(b) if b is not None else "") == str(g["blur_sha"][l]), f"blurred level {l}"
    gk = g["kps"]
    assert len(kps) == len(gk) and int(lc.sum()) == len(gk)
    got = np.stack([kps["x"], kps["y"], kps["size"], kps["angle"], kps["response"], kps["octave"].astype(np.float32)], 1)
    assert np.array_equal(got.view(np.uint32), gk.view(np.uint32)), "keypoints (x, y, size, angle, response, octave)"
    assert (kps["class_id"] == -1).all()
    assert np.array_equal(desc, g["desc"])


def test_fast_atan2_vectors(oracle):
    g = np.load(os.path.join(GOLDEN, "fast_atan2.npz"))
    got = np.array([oracle.fast_atan2(float(y), float(x)) for y, x in zip(g["y"], g["x"])], np.float32)
    assert np.array_equal(got.view(np.uint32), g["angle"].view(np.uint32))


def test_primitive_vectors(oracle):
    g = np.load(os.path.join(GOLDEN, "primitives_333x257.npz"))
    src = g["src"]
    assert np.array_equal(oracle.resize(src, 278, 214), g["resize"])
    assert np.array_equal(oracle.border101(src, 19), g["border"])
    assert np.array_equal(oracle.blur(src), g["blur"])
    for th, key in ((20, "fast20"), (7, "fast7")):
        k = oracle.fast(src, th)
        got = np.stack([k["x"], k["y"], k["response"]], 1).astype(np.float32)
        assert np.array_equal(got, g[key]), key
        # F5: corner at T <=> threshold-free score >= T
        sm = oracle.fast_score_map(src)
        assert (sm[k["y"].astype(int), k["x"].astype(int)] == k["response"]).all()


def test_constructor_tables(oracle):
    """src/ORBextractor.cc:776-925 — values listed in SURVEY.md section 8."""
    oe = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    assert list(oe.quota()) == [217, 181, 151, 126, 105, 87, 73, 60]
    assert list(oracle.OracleExtractor(2000).quota()) == [434, 362, 302, 251, 209, 175, 145, 122]
    assert list(oracle.OracleExtractor(1200).quota()) == [261, 217, 181, 151, 126, 105, 87, 72]
    assert list(oracle.OracleExtractor(8000).quota()) == [1737, 1448, 1207, 1005, 838, 698, 582, 485]
    assert list(oe.umax()) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    sc = oe.table(0)
    want = np.array([1, 1.2000000477, 1.4400000572, 1.7280001640, 2.0736002922, 2.4883203506, 2.9859845638, 3.5831816196], np.float32)
    assert np.array_equal(sc, want)
    assert np.array_equal(oe.table(1), np.float32(1) / sc)
    assert np.array_equal(oe.table(2), sc * sc)
    assert np.array_equal(oe.table(3), np.float32(1) / (sc * sc))


def test_descriptor_distance_and_knn_rule(oracle):
    """src/ORBmatcher.cc:2073-2093 (SWAR popcount) and the strict-< best/second update (:304-335)."""
    rng = np.random.RandomState(0)
    a = rng.randint(0, 256, (200, 32), dtype=np.uint8)
    b = rng.randint(0, 256, (200, 32), dtype=np.uint8)
    want = np.unpackbits(a ^ b, axis=1).sum(1)
    got = np.array([oracle.descriptor_distance(a[i], b[i]) for i in range(200)])
    assert np.array_equal(got, want)
    # plain-Python statement of the scan
    q, db = synth.synth_descriptors(40, 300, seed=3)
    db = np.concatenate([db, db[:5]])
    d = np.unpackbits(q[:, None, :] ^ db[None, :, :], axis=2).sum(2)
    r = oracle.hamming_knn(q, db)
    for i in range(len(q)):
        best, second, idx, idx2 = 256, 256, -1, -1
        for j in range(len(db)):
            if d[i, j] < best:
                second, idx2 = best, idx
                best, idx = int(d[i, j]), j
            elif d[i, j] < second:
                second, idx2 = int(d[i, j]), j
        assert (r["d1"][i], r["d2"][i], r["i1"][i]) == (best, second, idx)
        assert r["i2"][i] == (idx2 if second < 256 else -1)
    # threaded scan == single-threaded
    r4 = oracle.hamming_knn(q, db, nthreads=4)
    assert np.array_equal(r.view(np.uint8), r4.view(np.uint8))


def test_quadtree_properties(oracle):
    """DistributeOctTree (:1050-1417): result is a subset, one key per node, count bounded by N+3 / 4*nIni."""
    rng = np.random.RandomState(9)
    n = 3000
    kps = np.zeros(n, oracle.KP_DTYPE)
    kps["x"] = rng.randint(0, 600, n)
    kps["y"] = rng.randint(0, 440, n)
    kps["response"] = rng.randint(7, 120, n)
    for N in (1, 50, 217, 1000, 5000):
        out = oracle.distribute(kps, 0, 600, 0, 440, N)
        assert 0 < len(out) <= max(N + 3, 4)
        keys = set(zip(kps["x"].tolist(), kps["y"].tolist(), kps["response"].tolist()))
        assert all((float(a), float(b), float(c)) in keys for a, b, c in zip(out["x"], out["y"], out["response"]))
    assert len(oracle.distribute(kps[:0], 0, 600, 0, 440, 100)) == 0
    one = oracle.distribute(kps[:1], 0, 600, 0, 440, 100)
    assert len(one) == 1 and one["x"][0] == kps["x"][0]


def test_oracle_is_deterministic_and_reentrant(oracle):
    img = synth.synth_frame(320, 240, 42)
    a = oracle.OracleExtractor(500)(img)
    b = oracle.OracleExtractor(500)(img)
    assert np.array_equal(a[0].view(np.uint8), b[0].view(np.uint8)) and np.array_equal(a[1], b[1])
    frames = np.stack([img, img[::-1].copy()])
    assert oracle.extract_many(frames, 500, 2) == oracle.extract_many(frames, 500, 1) > 0


@pytest.mark.parametrize("path", STEREO, ids=[os.path.basename(p)[:-4] for p in STEREO])
def test_stereo_matches_cv2_composition(oracle, path):
    """orb_oracle_stereo_match (restatement of src/Frame.cc:880-1176) == the real-OpenCV composition: mvuRight and mvDepth
    bit for bit, on the oracle's own extraction of both images."""
    g = np.load(path)
    w, h, nf, seed, shift = (int(v) for v in g["params"])
    left, right = synth.synth_stereo_pair(w, h, seed, shift)
    oL, oR = oracle.OracleExtractor(nf), oracle.OracleExtractor(nf)
    kl, dl, _ = oL(left)
    kr, dr, _ = oR(right)
    assert len(kl) == int(g["n_left"]) and len(kr) == int(g["n_right"])
    ur, dp, kept = oracle.stereo_match(oL, oR, kl, dl, kr, dr, float(g["bf"]), float(g["b"]))
    assert np.array_equal(ur.view(np.uint32), g["uright"].view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), g["depth"].view(np.uint32))
    assert kept == int((g["uright"] >= 0).sum()) and kept > 200


def test_libm_trig_descriptor_bit_agreement(oracle):
    """DESIGN.md §2 canonicalisation 2: cosf / sinf (src/ORBextractor.cc:184-186) are libm calls, not correctly rounded by contract; the
    canonical sincos (double reduction + kernels, one rounding) is what oracle and CUDA share. This diagnostic reports how many descriptor
    bits change when the oracle calls glibc's cosf / sinf instead (flag LIBM_TRIG) — north_star's "descriptor bit agreement reported"."""
    total = diff = rows = rows_diff = 0
    for name, seed in (("tum", 1234), ("kitti", 1236), ("euroc", 1237)):
        w, h, nf = synth.CONFIGS[name]
        img = synth.synth_frame(w, h, seed)
        k0, d0, _ = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)(img)
        k1, d1, _ = oracle.OracleExtractor(nf, 1.2, 8, 20, 7, flags=oracle.LIBM_TRIG)(img)
        assert np.array_equal(k0.view(np.uint8), k1.view(np.uint8))          # keypoints and angles do not depend on cosf / sinf
        x = np.unpackbits(d0 ^ d1, axis=1).sum(axis=1)
        total += d0.size * 8; diff += int(x.sum()); rows += len(d0); rows_diff += int((x > 0).sum())
    agreement = 1.0 - diff / total
    print(f"descriptor bit agreement canonical vs glibc cosf/sinf: {agreement:.7f} ({diff} of {total} bits, {rows_diff} of {rows} rows)")
    assert agreement >= 0.9999
