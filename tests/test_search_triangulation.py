"""ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:879-1090) + CheckDistEpipolarLine (:206-245): oracle vs an independent
Python statement (CPU), orb_search_for_triangulation vs oracle (GPU). FeatureVectors come from the vocabulary transform."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth
from orb_slam2_detailed_comments_b200._lib import KP_DTYPE, EPI_DTYPE
from test_search_by_bow import _csr

f32 = np.float32
SCALE = (f32(1.2) ** np.arange(8)).astype(np.float32)
SIGMA2 = (SCALE * SCALE).astype(np.float32)


def _pair(voc, seed, n1=600, n2=650, shared=0.7, general_f=False, wide=False):
    """Two keyframes seeing the same scene under a sideways translation: a true match has (almost) the same image row."""
    k, L, parent, is_leaf, desc, weight = voc
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(is_leaf)
    d1 = desc[leaves[rng.integers(0, len(leaves), n1)]].copy()
    d1 ^= np.packbits(rng.random((n1, 256)) < 0.04, axis=1)
    d2 = desc[leaves[rng.integers(0, len(leaves), n2)]].copy()
    src = rng.integers(0, n1, n2)
    sh = rng.random(n2) < shared
    d2[sh] = d1[src[sh]]
    d2 ^= np.packbits(rng.random((n2, 256)) < 0.03, axis=1)
    dup = rng.random(n2) < 0.1                         # exact duplicates of the previous feature: equal distances, the later one must win
    dup[0] = False
    d2[dup] = d2[np.flatnonzero(dup) - 1]
    k1, k2 = np.zeros(n1, KP_DTYPE), np.zeros(n2, KP_DTYPE)
    for kk, n in ((k1, n1), (k2, n2)):
        kk["x"], kk["y"] = rng.uniform(0, 752, n).astype(f32), rng.uniform(0, 480, n).astype(f32)
        kk["octave"] = rng.integers(0, 8, n)
        kk["angle"] = rng.uniform(0, 360, n).astype(f32)
    k2["x"][sh] = k1["x"][src[sh]] - rng.uniform(5, 60, int(sh.sum())).astype(f32)
    k2["y"][sh] = k1["y"][src[sh]] + rng.normal(0, 1.2, int(sh.sum())).astype(f32)
    k2["y"][dup] = k2["y"][np.flatnonzero(dup) - 1]
    k2["angle"][sh] = (k1["angle"][src[sh]] - rng.normal(10, 3, int(sh.sum()))) % 360
    has1, has2 = (rng.random(n1) < 0.2).astype(np.uint8), (rng.random(n2) < 0.2).astype(np.uint8)
    ur1 = np.where(rng.random(n1) < 0.3, k1["x"] - 20, -1).astype(f32)
    ur2 = np.where(rng.random(n2) < 0.3, k2["x"] - 20, -1).astype(f32)
    F = np.array([0, 0, 0, 0, 0, -1, 0, 1, 0], f32)    # x1^T F12 x2 = y2 - y1
    if general_f:
        F = (F + rng.normal(0, [1e-6, 1e-6, 1e-3, 1e-6, 1e-6, 1e-3, 1e-3, 1e-3, 0.5], 9)).astype(f32)
    epi = np.zeros((), EPI_DTYPE)
    epi["f12"], epi["ex"], epi["ey"] = F, f32(300.5), f32(200.25)      # an epipole inside the image: its neighbourhood is excluded
    epi["level_sigma2"][:8], epi["scale_factor"][:8] = SIGMA2 * (f32(4000) if wide else f32(1)), SCALE
    return d1, k1, has1, ur1, d2, k2, has2, ur2, epi


def _epipolar_ok(p1, p2, F, sigma2):
    a = f32(f32(f32(p1["x"] * F[0]) + f32(p1["y"] * F[3])) + F[6])
    b = f32(f32(f32(p1["x"] * F[1]) + f32(p1["y"] * F[4])) + F[7])
    c = f32(f32(f32(p1["x"] * F[2]) + f32(p1["y"] * F[5])) + F[8])
    num = f32(f32(f32(a * p2["x"]) + f32(b * p2["y"])) + c)
    den = f32(f32(a * a) + f32(b * b))
    if den == 0:
        return False
    dsqr = f32(f32(num * num) / den)
    return float(dsqr) < 3.84 * float(sigma2[p2["octave"]])


def _python_rule(d1, k1, has1, ur1, fv1, d2, k2, has2, ur2, fv2, epi, only_stereo, th, ori):
    m12 = np.full(len(d1), -1, np.int32)
    matched2 = np.zeros(len(d2), bool)
    b1, b2 = np.unpackbits(d1, axis=1), np.unpackbits(d2, axis=1)
    F, ex, ey = epi["f12"], f32(epi["ex"]), f32(epi["ey"])
    events = []
    with np.errstate(over="ignore"):
        for node in sorted(set(fv1) & set(fv2)):
            for i1 in fv1[node]:
                if has1[i1] or (only_stereo and not ur1[i1] >= 0):
                    continue
                best, best_i = th, -1
                for i2 in fv2[node]:
                    if matched2[i2] or has2[i2] or (only_stereo and not ur2[i2] >= 0):
                        continue
                    dist = int(np.count_nonzero(b1[i1] != b2[i2]))
                    if dist > th or dist > best:
                        continue
                    if not ur1[i1] >= 0 and not ur2[i2] >= 0:
                        dx, dy = f32(ex - k2["x"][i2]), f32(ey - k2["y"][i2])
                        if f32(f32(dx * dx) + f32(dy * dy)) < f32(f32(100) * epi["scale_factor"][k2["octave"][i2]]):
                            continue
                    if _epipolar_ok(k1[i1], k2[i2], F, epi["level_sigma2"]):
                        best, best_i = dist, i2
                if best_i >= 0:
                    m12[i1] = best_i
                    matched2[best_i] = True
                    rot = f32(k1["angle"][i1]) - f32(k2["angle"][best_i])
                    if rot < 0:
                        rot = f32(rot + f32(360))
                    v = float(f32(rot * (f32(30) / f32(360))))
                    b = int(np.floor(abs(v) + 0.5))
                    events.append((0 if b == 30 else b, i1))
    nm = len(events)
    if ori:
        import guided_ref
        keep = guided_ref.three_maxima([sum(1 for b, _ in events if b == i) for i in range(30)])
        for b, i1 in events:
            if b not in keep:
                m12[i1] = -1
                nm -= 1
    return m12, nm


@pytest.mark.parametrize("seed,ori,only_stereo,general_f,wide", [(1, True, False, False, False), (2, False, False, True, False),
                                                                   (3, True, True, False, False), (4, True, False, True, True)])
def test_oracle_equals_python_statement(oracle, seed, ori, only_stereo, general_f, wide):
    voc = synth.synth_vocabulary(10, 3, seed=11)
    ov = oracle.OracleVocabulary(*voc)
    d1, k1, has1, ur1, d2, k2, has2, ur2, epi = _pair(voc, seed, 300, 340, general_f=general_f, wide=wide)
    fv1, fv2 = ov.transform(d1, 2)[3], ov.transform(d2, 2)[3]
    got = oracle.search_for_triangulation(d1, k1, has1, ur1, _csr(fv1, len(d1)), d2, k2, has2, ur2, _csr(fv2, len(d2)), epi["f12"], epi["ex"],
                                          epi["ey"], epi["level_sigma2"], epi["scale_factor"], only_stereo, 50, ori)
    want = _python_rule(d1, k1, has1, ur1, fv1, d2, k2, has2, ur2, fv2, epi, only_stereo, 50, ori)
    assert got[1] == want[1] and np.array_equal(got[0], want[0])
    assert want[1] > (5 if only_stereo else 25)


@pytest.mark.gpu
@pytest.mark.parametrize("levelsup,ori,only_stereo,general_f,wide", [(1, True, False, False, False), (2, True, False, True, False),
                                                                       (0, False, True, False, False), (2, True, False, False, True)])
def test_gpu_equals_oracle(oracle, levelsup, ori, only_stereo, general_f, wide):
    import orb_slam2_detailed_comments_b200 as pkg
    voc = synth.synth_vocabulary(10, 3, seed=11)
    gv, ov = pkg.ORBVocabulary(*voc), oracle.OracleVocabulary(*voc)
    sizes = [(700, 800), (1500, 1400), (1, 5), (300, 0), (0, 300), (900, 901)]
    P, qcap, cap = len(sizes), 1503, 1407
    rng = np.random.default_rng(0)
    D1 = rng.integers(0, 256, (P, qcap, 32), dtype=np.uint8); D2 = rng.integers(0, 256, (P, cap, 32), dtype=np.uint8)
    K1, K2 = np.zeros((P, qcap), KP_DTYPE), np.zeros((P, cap), KP_DTYPE)
    H1, H2 = np.zeros((P, qcap), np.uint8), np.zeros((P, cap), np.uint8)
    U1, U2 = np.full((P, qcap), -1, f32), np.full((P, cap), -1, f32)
    E = np.zeros(P, EPI_DTYPE)
    cases = []
    for i, (a, b) in enumerate(sizes):
        d1, k1, has1, ur1, d2, k2, has2, ur2, epi = _pair(voc, 60 + i, max(a, 1), max(b, 1), general_f=general_f, wide=wide)
        d1, k1, has1, ur1, d2, k2, has2, ur2 = d1[:a], k1[:a], has1[:a], ur1[:a], d2[:b], k2[:b], has2[:b], ur2[:b]
        D1[i, :a], K1[i, :a], H1[i, :a], U1[i, :a], D2[i, :b], K2[i, :b], H2[i, :b], U2[i, :b], E[i] = d1, k1, has1, ur1, d2, k2, has2, ur2, epi
        cases.append((d1, k1, has1, ur1, d2, k2, has2, ur2, epi))
    n1, n2 = np.array([s[0] for s in sizes], np.int32), np.array([s[1] for s in sizes], np.int32)
    bow1, bow2 = gv.transform(D1, n1, levelsup), gv.transform(D2, n2, levelsup)
    m12, nm = pkg.ORBmatcher(0.6, ori).SearchForTriangulation(D1, K1, H1, U1, bow1, D2, K2, H2, U2, bow2, E, only_stereo=only_stereo)
    total = 0
    for i, (d1, k1, has1, ur1, d2, k2, has2, ur2, epi) in enumerate(cases):
        fv1 = ov.transform(d1, levelsup)[3] if len(d1) else {}
        fv2 = ov.transform(d2, levelsup)[3] if len(d2) else {}
        want = oracle.search_for_triangulation(d1, k1, has1, ur1, _csr(fv1, max(len(d1), 1)), d2, k2, has2, ur2, _csr(fv2, max(len(d2), 1)),
                                               epi["f12"], epi["ex"], epi["ey"], epi["level_sigma2"], epi["scale_factor"], only_stereo, 50, ori)
        assert int(nm[i]) == want[1], f"pair {i}: nmatches {int(nm[i])} vs {want[1]}"
        assert np.array_equal(m12[i, :len(d1)], want[0]), f"pair {i}: vMatches12"
        assert np.all(m12[i, len(d1):] == -1)
        total += want[1]
    assert total > (20 if only_stereo else 200)
