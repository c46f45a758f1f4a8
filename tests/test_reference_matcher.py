"""Pins the matcher-side oracles — and on the GPU box the CUDA kernels directly — against the REFERENCE'S OWN CODE for
M1 DescriptorDistance, M2 the best / second-best scan with both accept tests, ComputeThreeMaxima, N4 the least-median loop of
MapPoint::ComputeDistinctiveDescriptors, N1 Frame::ComputeStereoMatches and N2: the frame grid, SearchByProjection (local-map form), SearchByBoW and
SearchForInitialization.

How: src/ORBmatcher.cc, src/MapPoint.cc and src/Frame.cc cannot be compiled whole in this image (they include the whole SLAM system, Eigen,
g2o, Pangolin), but these function bodies touch none of it. `make -C oracle _ref` cuts them out of /root/reference BY LINE RANGE into
generated, git-ignored includes (oracle/_ref/gen_*.inc; each range's first line is grep-asserted) and compiles them inside
oracle/ref_match_main.cpp, which only declares the members they name, over oracle/cvshim. DBoW2's FORB::distance — the verbatim twin of
DescriptorDistance — is linked in compiled unmodified. The binary travels to the GPU box with the snapshot."""
import numpy as np
import pytest

import ref_binary
from orb_slam2_detailed_comments_b200 import synth

needs_ref = pytest.mark.skipif(not ref_binary.match_available(), reason="oracle/_ref/orb_ref_match not built (needs /root/reference at build time)")
INT_MAX = 2**31 - 1


def _special_descriptors():
    rows = [np.zeros(32, np.uint8), np.full(32, 255, np.uint8), np.full(32, 0x55, np.uint8), np.full(32, 0xAA, np.uint8)]
    for bit in (0, 7, 8, 31, 32, 127, 128, 255):
        r = np.zeros(32, np.uint8)
        r[bit >> 3] = 1 << (bit & 7)
        rows.append(r)
    return np.stack(rows)


def _popcount_pairs(a, b):
    return np.unpackbits(a ^ b, axis=1).sum(axis=1).astype(np.int32)


# ---- M1 ----------------------------------------------------------------------------------------------------------------------------------
@needs_ref
def test_descriptor_distance_equals_reference(oracle):
    rng = np.random.RandomState(3)
    n = 120_000
    a = rng.randint(0, 256, (n, 32), dtype=np.uint8)
    b = rng.randint(0, 256, (n, 32), dtype=np.uint8)
    near = rng.rand(n) < 0.3                        # close pairs: a few flipped bits, like real matches
    b[near] = a[near]
    flips = rng.randint(0, 256, (int(near.sum()), 6))
    for j in range(6):
        b[np.flatnonzero(near), flips[:, j] >> 3] ^= (1 << (flips[:, j] & 7)).astype(np.uint8)
    sp = _special_descriptors()
    ia, ib = np.meshgrid(np.arange(len(sp)), np.arange(len(sp)))
    a = np.concatenate([a, sp[ia.ravel()]])
    b = np.concatenate([b, sp[ib.ravel()]])
    forb, dd = ref_binary.ref_distance(a, b)
    assert np.array_equal(forb, dd), "FORB::distance and ORBmatcher::DescriptorDistance are verbatim twins"
    assert np.array_equal(dd, _popcount_pairs(a, b))
    assert dd.min() == 0 and dd.max() == 256
    got = np.array([oracle.descriptor_distance(a[i], b[i]) for i in range(len(a))], np.int32)
    assert np.array_equal(got, dd), "oracle DescriptorDistance != reference DescriptorDistance"


# ---- M2 ----------------------------------------------------------------------------------------------------------------------------------
def _scan_cases():
    cases = []
    q, db = synth.synth_descriptors(300, 4000, seed=21)              # 10 % planted near-duplicates: accepts exist
    cases.append(("planted", q, db))
    q2, db2 = synth.synth_descriptors(64, 500, seed=22)
    db2 = np.concatenate([db2, db2[:40], db2[:40]])                    # ties on the minimum: earliest index wins, duplicate becomes second
    cases.append(("ties", np.concatenate([q2, db2[:20]]), db2))
    cases.append(("one_row", q2[:9], db2[:1]))                         # second stays at the sentinel
    cases.append(("empty", q2[:5], db2[:0]))                           # best stays at the sentinel, index -1
    cases.append(("dist256", np.full((3, 32), 255, np.uint8), np.zeros((6, 32), np.uint8)))   # 256 is never < 256
    sp = _special_descriptors()
    cases.append(("special", sp, np.concatenate([sp[::-1], sp])))
    return cases


@needs_ref
@pytest.mark.parametrize("name,q,db", _scan_cases(), ids=[c[0] for c in _scan_cases()])
def test_scan_equals_reference(oracle, name, q, db):
    """oracle.hamming_knn == the loop of src/ORBmatcher.cc:304-335 (sentinel 256) and of :611-637 (sentinel INT_MAX), and the product's accept
    expression == :339-343 evaluated by the reference's own compiled compare."""
    from orb_slam2_detailed_comments_b200.matcher import ORBmatcher
    for ratio in (0.6, 0.7, 0.9):
        ref = ref_binary.ref_scan(q, db, ratio)
        got = oracle.hamming_knn(q, db, sentinel=256)
        assert np.array_equal(got["d1"], ref[:, 0]), "best distance"
        assert np.array_equal(got["d2"], ref[:, 1]), "second distance"
        assert np.array_equal(got["i1"], ref[:, 2].astype(np.int64)), "best index"
        acc = ORBmatcher(ratio).accept(got)
        assert np.array_equal(acc.astype(np.int32), ref[:, 3]), "accept test (TH_LOW + float ratio compare)"
    if name == "planted":
        assert ref[:, 3].sum() > 0 and ref[:, 3].sum() < len(q)
    # SearchForInitialization's form: INT_MAX sentinels; with no keypoint matched yet (vMatchedDistance = INT_MAX) nothing is skipped
    ref = ref_binary.ref_scan(q, db, 0.9, matched_distance=np.full(len(db), INT_MAX, np.int32))
    got = oracle.hamming_knn(q, db, sentinel=INT_MAX)
    assert np.array_equal(got["d1"], ref[:, 0]) and np.array_equal(got["d2"], ref[:, 1]) and np.array_equal(got["i1"], ref[:, 2].astype(np.int64))


@needs_ref
def test_scan_init_skip_rule_equals_reference(oracle):
    """`if (vMatchedDistance[i2] <= dist) continue;` (src/ORBmatcher.cc:624): rows whose recorded distance is not larger are invisible —
    the same as removing them from the database. Checked by removing them for the oracle."""
    rng = np.random.RandomState(5)
    q, db = synth.synth_descriptors(80, 600, seed=23)
    md = rng.randint(60, 200, len(db)).astype(np.int32)
    md[rng.rand(len(db)) < 0.5] = INT_MAX
    ref = ref_binary.ref_scan(q, db, 0.9, matched_distance=md)
    dist = np.unpackbits(q[:, None, :] ^ db[None, :, :], axis=2).sum(axis=2).astype(np.int32)
    for i in range(len(q)):
        keep = np.flatnonzero(md > dist[i])
        got = oracle.hamming_knn(q[i:i + 1], db[keep], sentinel=INT_MAX)[0]
        assert got["d1"] == ref[i, 0] and got["d2"] == ref[i, 1]
        assert (keep[got["i1"]] if got["i1"] >= 0 else -1) == ref[i, 2]


# ---- ComputeThreeMaxima --------------------------------------------------------------------------------------------------------------------
@needs_ref
def test_three_maxima_equals_reference(oracle):
    rng = np.random.RandomState(7)
    rows = [rng.randint(0, 40, 30) for _ in range(400)]
    rows += [rng.randint(0, 3, 30) for _ in range(200)]                       # many ties
    rows += [np.zeros(30, np.int64), np.full(30, 5), np.eye(30, dtype=np.int64)[4] * 9]
    for m1 in (10, 20, 100, 101):                                              # the 10 % rule at and around the boundary (float compare)
        for m2 in (m1 // 10 - 1, m1 // 10, m1 // 10 + 1):
            for m3 in (0, max(m2 - 1, 0), m1 // 10):
                r = np.zeros(30, np.int64)
                r[[3, 17, 29]] = [m1, max(m2, 0), m3]
                rows.append(r)
                rows.append(r[::-1].copy())
    sizes = np.stack(rows).astype(np.int32)
    ref = ref_binary.ref_three_maxima(sizes)
    got = np.stack([oracle.three_maxima(s) for s in sizes])
    assert np.array_equal(got, ref)
    assert (ref[:, 1] == -1).any() and (ref[:, 2] == -1).any() and (ref[:, 2] >= 0).any()


# ---- N4 ----------------------------------------------------------------------------------------------------------------------------------
def _distinctive_case(seed=31, points=300):
    rng = np.random.RandomState(seed)
    sizes = rng.randint(1, 24, points)
    sizes[:6] = [1, 2, 2, 3, 40, 97]
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    desc = rng.randint(0, 256, (int(offsets[-1]), 32), dtype=np.uint8)
    for p in range(points):                                                    # observations of one point resemble each other; some are equal
        base = desc[offsets[p]].copy()
        for r in range(offsets[p] + 1, offsets[p + 1]):
            d = base.copy()
            for bit in rng.randint(0, 256, rng.randint(0, 30)):
                d[bit >> 3] ^= 1 << (bit & 7)
            desc[r] = d
    return desc, offsets


@needs_ref
def test_distinctive_descriptors_equal_reference(oracle):
    desc, offsets = _distinctive_case()
    ref = ref_binary.ref_distinctive(desc, offsets)
    bi, bm = oracle.distinctive_descriptors(desc, offsets)
    assert np.array_equal(bi, ref[:, 0]), "BestIdx"
    assert np.array_equal(bm, ref[:, 1]), "BestMedian"


# ---- N1 ----------------------------------------------------------------------------------------------------------------------------------
def _safe_rows(k, scale, n_rows, right):
    """Keypoints the reference can take without undefined behaviour: with this fork's padded coordinates a high-octave keypoint near the bottom
    bins into vRowIndices[yi >= nRows] (src/Frame.cc:925-930, DESIGN.md canonicalisation 5); those are left out on both sides of the test."""
    y = k["y"]
    if right:
        r = np.float32(2.0) * scale[k["octave"]]
        return (np.floor(y - r) >= 0) & (np.ceil(y + r) <= n_rows - 1)
    return (y >= 0) & (y.astype(np.int64) < n_rows)


@needs_ref
@pytest.mark.parametrize("w,h,nf,shift,bf,b", [(640, 480, 1000, 9, 40.0, 0.08), (752, 480, 1200, 14, 47.9, 0.11), (1241, 376, 2000, 21, 386.1448, 0.5372)])
def test_stereo_matches_equal_reference(oracle, w, h, nf, shift, bf, b):
    left, right = synth.synth_stereo_pair(w, h, 4321 + w, shift)
    oL, oR = oracle.OracleExtractor(nf, 1.2, 8, 20, 7), oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    kl, dl, _ = oL(left)
    kr, dr, _ = oR(right)
    scale = oL.table(0)
    n_rows = oL.level(0).shape[0]
    ml, mr = _safe_rows(kl, scale, n_rows, False), _safe_rows(kr, scale, n_rows, True)
    assert ml.sum() > 0.8 * len(kl) and mr.sum() > 0.8 * len(kr)
    kl, dl, kr, dr = kl[ml], dl[ml], kr[mr], dr[mr]
    ur, dp, kept = oracle.stereo_match(oL, oR, kl, dl, kr, dr, bf, b)
    rur, rdp = ref_binary.ref_stereo(scale, [oL.level(l) for l in range(8)], [oR.level(l) for l in range(8)], kl, dl, kr, dr, bf, b)
    assert kept > 50, kept
    assert np.array_equal(ur.view(np.uint32), rur.view(np.uint32)), f"mvuRight differs at {np.flatnonzero(ur != rur)[:8]}"
    assert np.array_equal(dp.view(np.uint32), rdp.view(np.uint32)), "mvDepth"
    assert int((rur >= 0).sum()) == kept


@needs_ref
def test_stereo_matches_equal_reference_upstream_pyramid(oracle):
    """The oracle's ComputeStereoMatches under UPSTREAM_PYRAMID (mvImagePyramid[l] = the un-padded interior) against the reference's own
    function body fed with those interior levels. With un-padded coordinates every keypoint row is inside vRowIndices: nothing is left out."""
    w, h, nf, shift, bf, b = 752, 480, 1200, 14, 47.9, 0.11
    left, right = synth.synth_stereo_pair(w, h, 977, shift)
    oL = oracle.OracleExtractor(nf, 1.2, 8, 20, 7, flags=oracle.UPSTREAM_PYRAMID)
    oR = oracle.OracleExtractor(nf, 1.2, 8, 20, 7, flags=oracle.UPSTREAM_PYRAMID)
    kl, dl, _ = oL(left)
    kr, dr, _ = oR(right)
    scale = oL.table(0)
    ml, mr = _safe_rows(kl, scale, h, False), _safe_rows(kr, scale, h, True)
    kl, dl, kr, dr = kl[ml], dl[ml], kr[mr], dr[mr]
    ur, dp, kept = oracle.stereo_match(oL, oR, kl, dl, kr, dr, bf, b)
    core = lambda a: np.ascontiguousarray(a[19:-19, 19:-19])
    rur, rdp = ref_binary.ref_stereo(scale, [core(oL.level(l)) for l in range(8)], [core(oR.level(l)) for l in range(8)], kl, dl, kr, dr, bf, b)
    assert kept > 50, kept
    assert np.array_equal(ur.view(np.uint32), rur.view(np.uint32)), f"mvuRight differs at {np.flatnonzero(ur != rur)[:8]}"
    assert np.array_equal(dp.view(np.uint32), rdp.view(np.uint32)), "mvDepth"


# ---- CUDA == reference code, directly ----------------------------------------------------------------------------------------------------------
@needs_ref
@pytest.mark.gpu
def test_cuda_knn_equals_reference_scan():
    from orb_slam2_detailed_comments_b200 import hamming_knn
    from orb_slam2_detailed_comments_b200.matcher import ORBmatcher
    for name, q, db in _scan_cases():
        ref = ref_binary.ref_scan(q, db, 0.6)
        got = hamming_knn(q, db)
        assert np.array_equal(got["d1"], ref[:, 0]) and np.array_equal(got["d2"], ref[:, 1]), name
        assert np.array_equal(got["i1"], ref[:, 2].astype(np.int64)), name
        assert np.array_equal(ORBmatcher(0.6).accept(got).astype(np.int32), ref[:, 3]), name
        ref = ref_binary.ref_scan(q, db, 0.9, matched_distance=np.full(len(db), INT_MAX, np.int32))
        got = hamming_knn(q, db, sentinel=INT_MAX)
        assert np.array_equal(got["d1"], ref[:, 0]) and np.array_equal(got["d2"], ref[:, 1]) and np.array_equal(got["i1"], ref[:, 2].astype(np.int64)), name


@needs_ref
@pytest.mark.gpu
def test_cuda_distance_and_distinctive_equal_reference():
    import torch
    from orb_slam2_detailed_comments_b200.matcher import ORBmatcher
    rng = np.random.RandomState(9)
    a = np.concatenate([rng.randint(0, 256, (50_000, 32), dtype=np.uint8), _special_descriptors()])
    b = np.concatenate([rng.randint(0, 256, (50_000, 32), dtype=np.uint8), _special_descriptors()[::-1]])
    _, dd = ref_binary.ref_distance(a, b)
    got = ORBmatcher.DescriptorDistanceBatch(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()).cpu().numpy()
    assert np.array_equal(got, dd)
    desc, offsets = _distinctive_case(seed=33)
    ref = ref_binary.ref_distinctive(desc, offsets)
    bi, bm = ORBmatcher.ComputeDistinctiveDescriptors(desc, offsets)
    assert np.array_equal(bi, ref[:, 0]) and np.array_equal(bm, ref[:, 1])


# ---- N2: guided matchers --------------------------------------------------------------------------------------------------------------------
PROJ_DTYPE = np.dtype([("projx", "<f4"), ("projy", "<f4"), ("projxr", "<f4"), ("viewcos", "<f4"), ("level", "<i4"), ("flags", "<i4")])
SCALE = np.cumprod(np.concatenate([[1.0], np.full(7, 1.2)])).astype(np.float32)


def _local_map_case(seed, n=900, m=700, w=752, h=480, stereo=True, th=1.0, bounds=None):
    """Keypoints / map points for the local-map SearchByProjection: the reference's inputs (projection, predicted level, viewing cosine per map
    point) and the C ABI's query records derived from them the way the host code derives them: radius = RadiusByViewingCos(cos) [* th] *
    mvScaleFactors[level] in float32, level window [level - 1, level] (src/ORBmatcher.cc:88-105)."""
    from orb_slam2_detailed_comments_b200._lib import QUERY_DTYPE
    c = synth.synth_guided_case(seed, n=n, m=m, w=w, h=h, stereo=stereo, bounds=bounds)
    rng = np.random.default_rng(seed + 1000)
    q = c["queries"]
    proj = np.zeros(m, PROJ_DTYPE)
    proj["projx"], proj["projy"], proj["projxr"] = q["u"], q["v"], q["ur"]
    proj["viewcos"] = np.where(rng.random(m) < 0.5, 0.9995, 0.97).astype(np.float32)
    proj["level"] = np.clip(q["max_level"], 0, 7)
    proj["flags"] = q["flags"]
    r = np.where(proj["viewcos"] > np.float32(0.998), np.float32(2.5), np.float32(4.0)).astype(np.float32)   # RadiusByViewingCos (:187-194)
    if th != 1.0:
        r = (r * np.float32(th)).astype(np.float32)
    q2 = np.zeros(m, QUERY_DTYPE)
    q2["u"], q2["v"], q2["ur"] = proj["projx"], proj["projy"], proj["projxr"]
    q2["radius"] = (r * SCALE[proj["level"]]).astype(np.float32)
    q2["min_level"], q2["max_level"] = proj["level"] - 1, proj["level"]
    q2["flags"] = proj["flags"]
    return c, proj, q2


@needs_ref
@pytest.mark.parametrize("seed,stereo,th,ratio,bounds", [(1, True, 1.0, 0.8, None), (2, False, 3.0, 0.8, None), (3, True, 5.0, 0.6, (-11.5, -7.25, 770.0, 493.5)),
                                                         (4, True, 1.0, 0.9, None)])
def test_search_by_projection_equals_reference(oracle, seed, stereo, th, ratio, bounds):
    """oracle.search_by_projection == the reference's own SearchByProjection(Frame&, vector<MapPoint*>&, th) + RadiusByViewingCos + frame grid
    (AssignFeaturesToGrid, PosInGrid, GetFeaturesInArea), compiled from line-range cuts: window lookup and scan order, level window, stereo check,
    claims through Observations(), level-conditional ratio test, TH_HIGH."""
    c, proj, q2 = _local_map_case(seed, stereo=stereo, th=th, bounds=bounds)
    nm_ref, as_ref = ref_binary.ref_search_by_projection(c["kps"], c["desc"], c["uright"], c["claimed"], c["bounds"], SCALE, th, ratio, proj, c["qdesc"])
    assigned, qmatch, nm = oracle.search_by_projection(c["kps"], c["desc"], q2, c["qdesc"], c["bounds"], th_dist=100, nn_ratio=ratio,
                                                       check_orientation=False, uright=c["uright"], claimed=c["claimed"])
    assert nm == nm_ref and nm_ref > 100, (nm, nm_ref)
    assert np.array_equal(assigned, as_ref), f"assignment differs at keypoints {np.flatnonzero(assigned != as_ref)[:10]}"


CAM = (512.0, 256.0, 320.0, 240.0)   # fx, fy, cx, cy: powers of two / small integers, so that the reference's projection is exact (see below)


def _frame_to_frame_case(seed, mode, n=900, m=800, stereo=True, th=15.0, mb=0.1, mbf=40.0):
    """Inputs of SearchByProjection(CurrentFrame, LastFrame, th, bMono) and the query records the host derives from them (src/ORBmatcher.cc:1744-1783).
    The pose algebra in front of the scan is NOT part of the device op (the host projects); it is made exact here so that both sides see the same
    u, v: identity rotation, zero translation of the current frame, depths that are powers of two, u and v with two fractional bits."""
    from orb_slam2_detailed_comments_b200._lib import QUERY_DTYPE
    c = synth.synth_guided_case(seed, n=n, m=m, stereo=stereo)
    rng = np.random.default_rng(seed + 77)
    q = c["queries"]
    fx, fy, cx, cy = (np.float32(v) for v in CAM)
    u = (np.round(q["u"] * 4) / 4).astype(np.float32)
    v = (np.round(q["v"] * 4) / 4).astype(np.float32)
    z = rng.choice(np.array([0.5, 1.0, 2.0, 4.0, -1.0], np.float32), m, p=[0.2, 0.3, 0.25, 0.2, 0.05])
    last = np.zeros(m, ref_binary.LAST_DTYPE)
    last["X"], last["Y"], last["Z"] = (u - cx) / fx * z, (v - cy) / fy * z, z
    last["angle"] = q["angle"]
    last["octave"] = np.clip(q["max_level"], 0, 7)
    last["flags"] = (rng.random(m) < 0.05) * 1 + (rng.random(m) < 0.8) * 2 + (rng.random(m) < 0.05) * 4
    invz = (np.float64(1.0) / z.astype(np.float64)).astype(np.float32)               # const float invzc = 1.0 / x3Dc.at<float>(2)
    uu = fx * last["X"] * invz + cx                                                   # float32 throughout, as the reference's expression
    vv = fy * last["Y"] * invz + cy
    ok = z > 0
    assert np.array_equal(uu[ok], u[ok]) and np.array_equal(vv[ok], v[ok])
    mnx, mny, mxx, mxy = (np.float32(b) for b in c["bounds"])
    skip = (last["flags"] & 1).astype(bool) | (last["flags"] & 4).astype(bool) | (invz < 0) | (uu < mnx) | (uu > mxx) | (vv < mny) | (vv > mxy)
    last_tz, mono = {"forward": (1.0, False), "backward": (-1.0, False), "still": (0.0, False), "mono": (1.0, True)}[mode]
    q2 = np.zeros(m, QUERY_DTYPE)
    q2["u"], q2["v"] = uu, vv
    q2["radius"] = (np.float32(th) * SCALE[last["octave"]]).astype(np.float32)
    q2["ur"] = uu - np.float32(mbf) * invz
    q2["angle"] = last["angle"]
    o = last["octave"].astype(np.int16)
    if mode == "forward":
        q2["min_level"], q2["max_level"] = o, -1
    elif mode == "backward":
        q2["min_level"], q2["max_level"] = 0, o
    else:
        q2["min_level"], q2["max_level"] = o - 1, o + 1
    q2["flags"] = np.where(skip, 1, np.where(last["flags"] & 2, 2, 0))
    return c, last, q2, dict(th=th, mb=mb, mbf=mbf, last_tz=last_tz, mono=mono)


@needs_ref
@pytest.mark.parametrize("seed,mode,stereo,ori", [(1, "still", True, True), (2, "forward", True, True), (3, "backward", True, False), (4, "mono", False, True),
                                                   (5, "still", False, False)])
def test_frame_to_frame_projection_equals_reference(oracle, seed, mode, stereo, ori):
    """oracle.search_by_projection == the reference's own SearchByProjection(CurrentFrame, LastFrame, th, bMono), the matcher of
    TrackWithMotionModel: forward / backward / neither level windows, outlier / null / behind-the-camera / out-of-bounds skips, stereo check with
    u - mbf * invzc, best-only accept at TH_HIGH, claims through Observations(), rotation histogram + three maxima."""
    c, last, q2, p = _frame_to_frame_case(seed, mode, stereo=stereo)
    nm_ref, as_ref = ref_binary.ref_search_by_projection_frames(c["kps"], c["desc"], c["uright"], c["claimed"], c["bounds"], SCALE, p["th"], CAM, p["mb"],
                                                                p["mbf"], p["last_tz"], p["mono"], ori, last, c["qdesc"])
    assigned, qmatch, nm = oracle.search_by_projection(c["kps"], c["desc"], q2, c["qdesc"], c["bounds"], th_dist=100, nn_ratio=0.0,
                                                       check_orientation=ori, uright=c["uright"], claimed=c["claimed"])
    assert nm == nm_ref and nm_ref > 100, (nm, nm_ref)
    assert np.array_equal(assigned, as_ref), f"assignment differs at keypoints {np.flatnonzero(assigned != as_ref)[:10]}"


@needs_ref
@pytest.mark.gpu
def test_cuda_frame_to_frame_projection_equals_reference():
    from orb_slam2_detailed_comments_b200.matcher import ORBmatcher
    for seed, mode, stereo, ori in ((21, "still", True, True), (22, "forward", True, True), (23, "backward", False, True), (24, "mono", False, False)):
        c, last, q2, p = _frame_to_frame_case(seed, mode, stereo=stereo)
        nm_ref, as_ref = ref_binary.ref_search_by_projection_frames(c["kps"], c["desc"], c["uright"], c["claimed"], c["bounds"], SCALE, p["th"], CAM,
                                                                    p["mb"], p["mbf"], p["last_tz"], p["mono"], ori, last, c["qdesc"])
        n, m = len(c["kps"]), len(q2)
        assigned, qmatch, nm = ORBmatcher(0.9, ori).SearchByProjection(c["kps"][None], c["desc"][None], np.array([n], np.int32), q2[None], c["qdesc"][None],
                                                                       np.array([m], np.int32), c["bounds"], use_ratio=False,
                                                                       uright=None if c["uright"] is None else c["uright"][None],
                                                                       claimed=c["claimed"][None])
        assert nm[0] == nm_ref and np.array_equal(assigned[0], as_ref), (seed, mode, nm[0], nm_ref)


@needs_ref
@pytest.mark.parametrize("seed,ori,ratio", [(1, True, 0.7), (2, False, 0.6), (3, True, 0.9)])
def test_search_by_bow_equals_reference(oracle, seed, ori, ratio):
    """oracle.search_by_bow == the reference's own SearchByBoW(KeyFrame*, Frame&, ...): FeatureVector walk with lower_bound, per-node scan, TH_LOW
    + ratio accept, claims through vpMapPointMatches, rotation histogram + ComputeThreeMaxima."""
    import test_search_by_bow as tb
    voc = synth.synth_vocabulary(10, 3, seed=11)
    ov = oracle.OracleVocabulary(*voc)
    kf, kk, valid, f, fk = tb._pair(voc, seed, 500, 560)
    kfv, ffv = ov.transform(kf, 2)[3], ov.transform(f, 2)[3]
    # FeatureVectors that do not share every node: exercises both lower_bound branches
    kfv = {n: v for n, v in kfv.items() if n % 5 != 1}
    ffv = {n: v for n, v in ffv.items() if n % 7 != 2}
    nm_ref, as_ref = ref_binary.ref_search_by_bow(kf, kk, valid, kfv, f, fk, ffv, ratio, ori)
    got = oracle.search_by_bow(kf, kk, valid, tb._csr(kfv, len(kf)), f, fk, tb._csr(ffv, len(f)), 50, ratio, ori)
    assert got[1] == nm_ref and nm_ref > 40, (got[1], nm_ref)
    assert np.array_equal(got[0], as_ref)


@needs_ref
@pytest.mark.parametrize("seed,ori,ratio", [(4, True, 0.75), (5, False, 0.6), (6, True, 0.9)])
def test_search_by_bow_keyframes_equals_reference(oracle, seed, ori, ratio):
    """The KeyFrame-KeyFrame form (loop closing, src/ORBmatcher.cc:724-867) == oracle.search_by_bow with the keyframe-2 features that hold no good
    map point blocked and th_dist = TH_LOW - 1 (this form accepts on `bestDist1 < TH_LOW`): vbMatched2 claims, strict ratio test, histogram over idx1."""
    import test_search_by_bow as tb
    voc = synth.synth_vocabulary(10, 3, seed=11)
    ov = oracle.OracleVocabulary(*voc)
    d1, k1, valid1, d2, k2 = tb._pair(voc, seed, 520, 540)
    rng = np.random.default_rng(seed)
    valid2 = (rng.random(len(d2)) < 0.85).astype(np.uint8)
    fv1, fv2 = ov.transform(d1, 2)[3], ov.transform(d2, 2)[3]
    fv1 = {n: v for n, v in fv1.items() if n % 6 != 1}
    fv2 = {n: v for n, v in fv2.items() if n % 5 != 3}
    nm_ref, m12_ref = ref_binary.ref_search_by_bow_keyframes(d1, k1, valid1, fv1, d2, k2, valid2, fv2, ratio, ori)
    got = oracle.search_by_bow(d1, k1, valid1, tb._csr(fv1, len(d1)), d2, k2, tb._csr(fv2, len(d2)), 49, ratio, ori, f_blocked=1 - valid2)
    assert got[1] == nm_ref and nm_ref > 40, (got[1], nm_ref)
    assert np.array_equal(got[2], m12_ref)


def _fuse_case(seed, n=900, m=800, stereo=True, th=3.0, mbf=40.0, null_points=True):
    """Inputs of Fuse(pKF, vpMapPoints, th) and the query records the host derives from them (src/ORBmatcher.cc:1117-1166): projection exact by
    construction (identity pose, power-of-two depths, u / v with two fractional bits), PredictScale and the viewing checks are host-side values."""
    from orb_slam2_detailed_comments_b200._lib import QUERY_DTYPE
    c = synth.synth_guided_case(seed, n=n, m=m, stereo=stereo)
    rng = np.random.default_rng(seed + 99)
    q = c["queries"]
    fx, fy, cx, cy = (np.float32(v) for v in CAM)
    u = (np.round(q["u"] * 4) / 4).astype(np.float32)
    v = (np.round(q["v"] * 4) / 4).astype(np.float32)
    z = rng.choice(np.array([0.5, 1.0, 2.0, 4.0, -1.0], np.float32), m, p=[0.2, 0.3, 0.25, 0.2, 0.05])
    pts = np.zeros(m, ref_binary.FUSE_DTYPE)
    pts["X"], pts["Y"], pts["Z"] = (u - cx) / fx * z, (v - cy) / fy * z, z
    pts["nz"] = np.where(rng.random(m) < 0.05, -1.0, 1.0)           # a normal pointing away fails PO.dot(Pn) < 0.5 * dist3D (:1153)
    pts["level"] = np.clip(q["max_level"], 0, 7)
    pts["flags"] = (rng.random(m) < 0.04) * (1 if null_points else 0) + (rng.random(m) < 0.04) * 2 + (rng.random(m) < 0.04) * 4
    ok = z > 0
    invz = (np.float32(1.0) / z).astype(np.float32)                  # const float invz = 1 / p3Dc.at<float>(2)
    uu = fx * (pts["X"] * invz) + cx
    vv = fy * (pts["Y"] * invz) + cy
    assert np.array_equal(uu[ok], u[ok]) and np.array_equal(vv[ok], v[ok])
    mnx, mny, mxx, mxy = (np.float32(b) for b in c["bounds"])
    in_image = (uu >= mnx) & (uu < mxx) & (vv >= mny) & (vv < mxy)   # KeyFrame::IsInImage (src/KeyFrame.cc:926-929)
    skip = (pts["flags"] != 0) | (z < 0) | ~in_image | (pts["nz"] < 0)
    q2 = np.zeros(m, QUERY_DTYPE)
    q2["u"], q2["v"] = uu, vv
    q2["radius"] = (np.float32(th) * SCALE[pts["level"]]).astype(np.float32)
    q2["ur"] = uu - np.float32(mbf) * invz
    q2["min_level"], q2["max_level"] = pts["level"] - 1, pts["level"]
    q2["flags"] = np.where(skip, 1, 0)
    inv_sigma2 = (np.float32(1.0) / (SCALE * SCALE)).astype(np.float32)
    return c, pts, q2, inv_sigma2, dict(th=th, mbf=mbf)


@needs_ref
@pytest.mark.parametrize("seed,stereo,th", [(1, True, 3.0), (2, False, 3.0), (3, True, 6.0)])
def test_fuse_equals_reference(oracle, seed, stereo, th):
    """oracle.fuse_search == the candidate loop of the reference's own Fuse(pKF, vpMapPoints, th) over KeyFrame::GetFeaturesInArea: level window
    [predicted - 1, predicted], chi-square gates 7.8 (stereo keypoints) / 5.99 (monocular) with mvInvLevelSigma2, best at TH_LOW."""
    c, pts, q2, inv_sigma2, p = _fuse_case(seed, stereo=stereo, th=th)
    held = c["claimed"]
    nf_ref, best_ref = ref_binary.ref_fuse(c["kps"], c["desc"], c["uright"], held, c["bounds"], SCALE, inv_sigma2, th, CAM, p["mbf"], pts, c["qdesc"])
    bi, bd = oracle.fuse_search(c["kps"], c["desc"], q2, c["qdesc"], c["bounds"], th_dist=50, uright=c["uright"], inv_level_sigma2=inv_sigma2)
    assert int((bi >= 0).sum()) == nf_ref and nf_ref > 60, (int((bi >= 0).sum()), nf_ref)
    assert np.array_equal(bi, best_ref), f"bestIdx differs at points {np.flatnonzero(bi != best_ref)[:10]}"


@needs_ref
@pytest.mark.parametrize("seed,stereo,th", [(4, True, 4.0), (5, False, 3.0)])
def test_fuse_sim3_form_equals_reference(oracle, seed, stereo, th):
    """The loop-closing form Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (Scw = identity: scale 1): same candidate loop without the chi-square gates,
    INT_MAX sentinel, spAlreadyFound skip == oracle.fuse_search without inv_level_sigma2."""
    c, pts, q2, inv_sigma2, p = _fuse_case(seed, stereo=stereo, th=th, null_points=False)   # this form dereferences every entry
    held = c["claimed"].copy()
    held[:64] = 1                                          # enough occupied slots for the points the keyframe already holds
    nf_ref, best_ref = ref_binary.ref_fuse(c["kps"], c["desc"], c["uright"], held, c["bounds"], SCALE, inv_sigma2, th, CAM, p["mbf"], pts, c["qdesc"],
                                           sim3_form=True)
    bi, bd = oracle.fuse_search(c["kps"], c["desc"], q2, c["qdesc"], c["bounds"], th_dist=50, uright=c["uright"], inv_level_sigma2=None)
    assert int((bi >= 0).sum()) == nf_ref and nf_ref > 60, (int((bi >= 0).sum()), nf_ref)
    assert np.array_equal(bi, best_ref), f"bestIdx differs at points {np.flatnonzero(bi != best_ref)[:10]}"


def _keyframe_projection_case(seed, loop_form, n=900, m=800, th=10.0):
    """Inputs of the two keyframe forms of SearchByProjection and the query records the host derives from them; projection exact by construction."""
    from orb_slam2_detailed_comments_b200._lib import QUERY_DTYPE
    c = synth.synth_guided_case(seed, n=n, m=m, stereo=False)
    rng = np.random.default_rng(seed + 55)
    q = c["queries"]
    fx, fy, cx, cy = (np.float32(v) for v in CAM)
    u = (np.round(q["u"] * 4) / 4).astype(np.float32)
    v = (np.round(q["v"] * 4) / 4).astype(np.float32)
    z = rng.choice(np.array([0.5, 1.0, 2.0, 4.0, -1.0], np.float32), m, p=[0.2, 0.3, 0.25, 0.2, 0.05])
    pts = np.zeros(m, ref_binary.KFPT_DTYPE)
    pts["X"], pts["Y"], pts["Z"] = (u - cx) / fx * z, (v - cy) / fy * z, z
    pts["angle"] = q["angle"]
    pts["level"] = np.clip(q["max_level"], 0, 7)
    pts["flags"] = (rng.random(m) < 0.04) * (0 if loop_form else 1) + (rng.random(m) < 0.04) * 2 + (rng.random(m) < 0.04) * 4
    invz = (np.float32(1.0) / z).astype(np.float32)
    if loop_form:                                                    # x = X * invz; u = fx * x + cx (:447-451)
        uu, vv = fx * (pts["X"] * invz) + cx, fy * (pts["Y"] * invz) + cy
    else:                                                            # u = fx * xc * invzc + cx (:1912-1913)
        uu, vv = fx * pts["X"] * invz + cx, fy * pts["Y"] * invz + cy
    assert np.array_equal(uu, u) and np.array_equal(vv, v)           # exact for either sign of z
    mnx, mny, mxx, mxy = (np.float32(b) for b in c["bounds"])
    if loop_form:
        skip = (pts["flags"] != 0) | (z < 0) | ~((uu >= mnx) & (uu < mxx) & (vv >= mny) & (vv < mxy))
    else:                                                            # this form has no `invzc < 0` test
        skip = (pts["flags"] != 0) | (uu < mnx) | (uu > mxx) | (vv < mny) | (vv > mxy)
    q2 = np.zeros(m, QUERY_DTYPE)
    q2["u"], q2["v"] = uu, vv
    q2["radius"] = (np.float32(th) * SCALE[pts["level"]]).astype(np.float32)
    q2["angle"] = pts["angle"]
    q2["min_level"], q2["max_level"] = pts["level"] - 1, pts["level"] + (0 if loop_form else 1)
    q2["flags"] = np.where(skip, 1, 2)                               # every accepted match blocks its keypoint for the later points
    held = c["claimed"].copy()
    held[:64] = 1                                                    # slots for the loop form's already-found points
    return c, pts, q2, held


@needs_ref
@pytest.mark.parametrize("seed,loop_form,ori,orbdist", [(1, False, True, 100), (2, False, False, 64), (3, True, False, 50), (4, True, False, 50)])
def test_keyframe_projection_forms_equal_reference(oracle, seed, loop_form, ori, orbdist):
    """The relocalisation form SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) and the loop-closing form SearchByProjection(pKF,
    Scw, vpPoints, vpMatched, th): oracle.search_by_projection with every query claiming, no stereo check, the form's level window and threshold."""
    th = 10.0
    c, pts, q2, held = _keyframe_projection_case(seed, loop_form, th=th)
    nm_ref, as_ref = ref_binary.ref_search_by_projection_keyframe(c["kps"], c["desc"], held, c["bounds"], SCALE, th, CAM, orbdist, ori, pts, c["qdesc"],
                                                                  loop_form=loop_form)
    assigned, qmatch, nm = oracle.search_by_projection(c["kps"], c["desc"], q2, c["qdesc"], c["bounds"], th_dist=orbdist, nn_ratio=0.0,
                                                       check_orientation=ori and not loop_form, uright=None, claimed=held)
    assert nm == nm_ref and nm_ref > 100, (nm, nm_ref)
    assert np.array_equal(assigned, as_ref), f"assignment differs at keypoints {np.flatnonzero(assigned != as_ref)[:10]}"


@needs_ref
@pytest.mark.parametrize("seed,th", [(1, 7.5), (2, 10.0)])
def test_search_by_sim3_equals_reference(oracle, seed, th):
    """SearchBySim3 == two passes of oracle.fuse_search (TH_HIGH, level window [predicted - 1, predicted], KeyFrame::GetFeaturesInArea / IsInImage) and
    the mutual check, with the pre-matched features of both keyframes left out as the reference does through vbAlreadyMatched1 / 2."""
    from orb_slam2_detailed_comments_b200._lib import QUERY_DTYPE, KP_DTYPE
    c = synth.synth_guided_case(seed + 40, n=850, m=800, stereo=False)
    rng = np.random.default_rng(seed + 5)
    fx, fy, cx, cy = (np.float32(v) for v in CAM)
    quant = lambda a: (np.round(a * 4) / 4).astype(np.float32)
    k2 = c["kps"].copy()
    k2["x"], k2["y"] = quant(k2["x"]), quant(k2["y"])
    d2 = c["desc"]
    q = c["queries"]
    k1 = np.zeros(len(q), KP_DTYPE)
    k1["x"], k1["y"], k1["octave"], k1["angle"] = quant(q["u"]), quant(q["v"]), np.clip(q["max_level"], 0, 7), q["angle"]
    d1 = c["qdesc"]
    mnx, mny, mxx, mxy = (np.float32(b) for b in c["bounds"])

    def points(k, other_n, with_prematch):
        n = len(k)
        p = np.zeros(n, ref_binary.SIM3_DTYPE)
        z = rng.choice(np.array([0.5, 1.0, 2.0, 4.0, -1.0], np.float32), n, p=[0.2, 0.3, 0.25, 0.2, 0.05])
        p["X"], p["Y"], p["Z"] = (k["x"] - cx) / fx * z, (k["y"] - cy) / fy * z, z
        p["level"] = k["octave"]
        p["flags"] = (rng.random(n) < 0.05) * 1 + (rng.random(n) < 0.03) * 2
        p["prematch"] = -1
        invz = (np.float32(1.0) / z).astype(np.float32)
        u, v = fx * (p["X"] * invz) + cx, fy * (p["Y"] * invz) + cy          # x = X * invz; u = fx * x + cx (:1496-1502)
        assert np.array_equal(u, k["x"]) and np.array_equal(v, k["y"])
        in_image = (u >= mnx) & (u < mxx) & (v >= mny) & (v < mxy)
        return p, (p["flags"] != 0) | (z < 0) | ~in_image

    p1, skip1 = points(k1, len(k2), True)
    p2, skip2 = points(k2, len(k1), False)
    pre = rng.choice(len(k1), 40, replace=False)
    tgt = rng.choice(np.flatnonzero(p2["flags"] == 0), 40, replace=False)
    pre = pre[p1["flags"][pre] == 0][:30]                                     # vpMatches12[i] != NULL only makes sense for features with a point
    p1["prematch"][pre] = tgt[:len(pre)]
    matched1 = p1["prematch"] >= 0
    matched2 = np.zeros(len(k2), bool)
    matched2[p1["prematch"][matched1]] = True

    def queries(k, p, skip):
        qq = np.zeros(len(k), QUERY_DTYPE)
        qq["u"], qq["v"] = k["x"], k["y"]
        qq["radius"] = (np.float32(th) * SCALE[p["level"]]).astype(np.float32)
        qq["min_level"], qq["max_level"] = p["level"] - 1, p["level"]
        qq["flags"] = np.where(skip, 1, 0)
        return qq

    nf_ref, m12_ref = ref_binary.ref_search_by_sim3(c["bounds"], th, CAM, SCALE, k1, d1, p1, k2, d2, p2)
    bi1, _ = oracle.fuse_search(k2, d2, queries(k1, p1, skip1 | matched1), d1, c["bounds"], th_dist=100)
    bi2, _ = oracle.fuse_search(k1, d1, queries(k2, p2, skip2 | matched2), d2, c["bounds"], th_dist=100)
    m12 = np.full(len(k1), -1, np.int32)
    for i1 in range(len(k1)):
        if bi1[i1] >= 0 and bi2[bi1[i1]] == i1:
            m12[i1] = bi1[i1]
    assert int((m12 >= 0).sum()) == nf_ref and nf_ref > 100, (int((m12 >= 0).sum()), nf_ref)
    assert np.array_equal(m12, m12_ref)


@needs_ref
@pytest.mark.gpu
def test_cuda_fuse_equals_reference():
    from orb_slam2_detailed_comments_b200.matcher import ORBmatcher
    for seed, stereo, th in ((31, True, 3.0), (32, False, 4.0)):
        c, pts, q2, inv_sigma2, p = _fuse_case(seed, stereo=stereo, th=th)
        nf_ref, best_ref = ref_binary.ref_fuse(c["kps"], c["desc"], c["uright"], c["claimed"], c["bounds"], SCALE, inv_sigma2, th, CAM, p["mbf"], pts,
                                               c["qdesc"])
        n, m = len(c["kps"]), len(q2)
        bi, bd = ORBmatcher().FuseSearch(c["kps"][None], c["desc"][None], np.array([n], np.int32), q2[None], c["qdesc"][None], np.array([m], np.int32),
                                         c["bounds"], uright=None if c["uright"] is None else c["uright"][None], inv_level_sigma2=inv_sigma2)
        assert int((bi[0] >= 0).sum()) == nf_ref and np.array_equal(bi[0], best_ref), (seed, nf_ref)


@needs_ref
@pytest.mark.parametrize("seed,ori,only_stereo,general_f,wide", [(1, True, False, False, False), (2, False, False, True, False), (3, True, True, False, False),
                                                                   (4, True, False, True, True)])
def test_search_for_triangulation_equals_reference(oracle, seed, ori, only_stereo, general_f, wide):
    """oracle.search_for_triangulation == the reference's own SearchForTriangulation + CheckDistEpipolarLine: FeatureVector walk, map-point / stereo
    skips, `dist > TH_LOW || dist > bestDist` (the later of equal distances wins), epipole exclusion for monocular pairs, the float32 epipolar distance
    against 3.84 * mvLevelSigma2, this fork's vbMatched2 bookkeeping (set on accept, cleared by the orientation filter), rotation histogram."""
    import test_search_triangulation as tt
    voc = synth.synth_vocabulary(10, 3, seed=11)
    ov = oracle.OracleVocabulary(*voc)
    d1, k1, has1, ur1, d2, k2, has2, ur2, epi = tt._pair(voc, seed, 320, 350, general_f=general_f, wide=wide)
    fv1, fv2 = ov.transform(d1, 2)[3], ov.transform(d2, 2)[3]
    fx, fy, cx, cy = (np.float32(v) for v in CAM)
    z = np.float32(2.0)
    cw = np.array([(epi["ex"] - cx) / fx * z, (epi["ey"] - cy) / fy * z, z], np.float32)
    assert fx * cw[0] * (np.float32(1.0) / z) + cx == epi["ex"] and fy * cw[1] * (np.float32(1.0) / z) + cy == epi["ey"]   # the reference's expression, exact here
    nm_ref, m12_ref = ref_binary.ref_search_for_triangulation(d1, k1, has1, ur1, fv1, d2, k2, has2, ur2, fv2, epi["f12"], cw, CAM, epi["level_sigma2"],
                                                              epi["scale_factor"][:8], only_stereo, ori)
    got = oracle.search_for_triangulation(d1, k1, has1, ur1, tt._csr(fv1, len(d1)), d2, k2, has2, ur2, tt._csr(fv2, len(d2)), epi["f12"], epi["ex"],
                                          epi["ey"], epi["level_sigma2"], epi["scale_factor"], only_stereo, 50, ori)
    assert got[1] == nm_ref and nm_ref > (5 if only_stereo else 25), (got[1], nm_ref)
    assert np.array_equal(got[0], m12_ref)


@needs_ref
@pytest.mark.gpu
def test_cuda_search_by_projection_equals_reference():
    from orb_slam2_detailed_comments_b200.matcher import ORBmatcher
    for seed, stereo, th, ratio in ((11, True, 1.0, 0.8), (12, False, 3.0, 0.8), (13, True, 5.0, 0.6)):
        c, proj, q2 = _local_map_case(seed, stereo=stereo, th=th)
        nm_ref, as_ref = ref_binary.ref_search_by_projection(c["kps"], c["desc"], c["uright"], c["claimed"], c["bounds"], SCALE, th, ratio, proj, c["qdesc"])
        n, m = len(c["kps"]), len(q2)
        assigned, qmatch, nm = ORBmatcher(ratio).SearchByProjection(c["kps"][None], c["desc"][None], np.array([n], np.int32), q2[None], c["qdesc"][None],
                                                                    np.array([m], np.int32), c["bounds"], use_ratio=True,
                                                                    uright=None if c["uright"] is None else c["uright"][None], claimed=c["claimed"][None])
        assert nm[0] == nm_ref and np.array_equal(assigned[0], as_ref), (seed, nm[0], nm_ref)


@needs_ref
@pytest.mark.parametrize("seed,ori", [(1, True), (2, False), (3, True)])
def test_search_for_initialization_equals_reference(oracle, seed, ori):
    """oracle.search_for_initialization == the reference's own SearchForInitialization: level-0 filter, window 100 on F2's grid, INT_MAX sentinels,
    vMatchedDistance skip, eviction of an earlier match by a strictly closer one, rotation histogram over ALL accepts, three maxima."""
    import test_search_init as ti
    from orb_slam2_detailed_comments_b200._lib import KP_DTYPE
    c = ti._case(seed, n=420, m=400)
    q = c["queries"]
    k1 = np.zeros(len(q), KP_DTYPE)
    k1["angle"] = q["angle"]
    k1["octave"] = np.where(q["flags"] & 1, 1, 0)        # `if (level1 > 0) continue;` (:597-599)
    prev = np.stack([q["u"], q["v"]], axis=1)
    nm_ref, m12_ref = ref_binary.ref_search_for_initialization(c["kps"], c["desc"], k1, c["qdesc"], prev, c["bounds"], 0.9, ori, 100)
    m12, m21, nm = oracle.search_for_initialization(c["kps"], c["desc"], q, c["qdesc"], c["bounds"], 50, 0.9, ori)
    assert nm == nm_ref and nm_ref > 40, (nm, nm_ref)
    assert np.array_equal(m12, m12_ref)


# ---- seed sweep: every pinned matcher again on further seeds with randomised switches (CPU, oracle against the reference's own code) ----

@needs_ref
@pytest.mark.parametrize("seed", range(20, 36))
def test_matcher_pins_seed_sweep(oracle, seed):
    rng = np.random.default_rng(1000 + seed)
    pick = lambda *a: a[int(rng.integers(len(a)))]
    flag = lambda: bool(rng.integers(2))
    test_search_by_projection_equals_reference(oracle, seed, flag(), pick(1.0, 3.0, 5.0), pick(0.6, 0.8, 0.9), None)
    mode = pick("still", "forward", "backward", "mono")
    test_frame_to_frame_projection_equals_reference(oracle, seed, mode, mode != "mono" and flag(), flag())
    test_search_by_bow_equals_reference(oracle, seed, flag(), pick(0.6, 0.7, 0.9))
    test_search_by_bow_keyframes_equals_reference(oracle, seed, flag(), pick(0.6, 0.75, 0.9))
    test_fuse_equals_reference(oracle, seed, flag(), pick(3.0, 4.0, 6.0))
    test_fuse_sim3_form_equals_reference(oracle, seed, flag(), pick(3.0, 4.0))
    loop_form = flag()
    test_keyframe_projection_forms_equal_reference(oracle, seed, loop_form, flag(), pick(50, 64, 100))
    test_search_by_sim3_equals_reference(oracle, seed, pick(7.5, 10.0))
    test_search_for_triangulation_equals_reference(oracle, seed, flag(), False, flag(), flag())
    test_search_for_initialization_equals_reference(oracle, seed, flag())
