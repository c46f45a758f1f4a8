"""Candidate search of ORBmatcher::Fuse (src/ORBmatcher.cc:1100-1280, 1282-1430) and of SearchBySim3's projection passes (:1433-1700):
oracle vs an independent Python statement (CPU), orb_fuse_search vs oracle (GPU)."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth
from orb_slam2_detailed_comments_b200._lib import KP_DTYPE, QUERY_DTYPE

f32 = np.float32
INV_SIGMA2 = (1.0 / (f32(1.2) ** np.arange(8)).astype(f32) ** 2).astype(f32)


def _case(seed, n=600, m=500, stereo=True):
    c = synth.synth_guided_case(seed, n=n, m=m, stereo=stereo, radius=9.0)
    q = c["queries"]
    rng = np.random.default_rng(seed + 1000)
    src_lvl = np.where(q["max_level"] >= 0, q["max_level"], q["min_level"]).astype(np.int16)
    pred = np.clip(src_lvl + rng.integers(-1, 2, len(q)), 0, 7).astype(np.int16)     # nPredictedLevel around the source keypoint's level
    q["min_level"], q["max_level"] = pred - 1, pred                                  # :1188
    q["flags"] &= 1                                                                  # no claims in these loops
    if len(q) > 4 and len(c["kps"]) > 4:                                             # exact duplicates in the target: the first in scan order wins
        c["desc"][1::7] = c["desc"][0::7][:len(c["desc"][1::7])]
    return c


def _python_rule(c, th, inv_sigma2):
    import guided_ref
    k, q = c["kps"], c["queries"]
    bits_k, bits_q = np.unpackbits(c["desc"], axis=1), np.unpackbits(c["qdesc"], axis=1)
    grid = guided_ref.Grid(k, c["bounds"])
    bi, bd = np.full(len(q), -1, np.int32), np.full(len(q), 256, np.int32)
    ur = c["uright"]
    for j in range(len(q)):
        if q["flags"][j] & 1:
            continue
        u, v, r = f32(q["u"][j]), f32(q["v"][j]), f32(q["radius"][j])
        best, best_i = 256, -1
        for i in grid.in_area(u, v, r, -1, -1):
            lvl = int(k["octave"][i])
            if lvl < q["min_level"][j] or lvl > q["max_level"][j]:
                continue
            if inv_sigma2 is not None:
                ex, ey = f32(u - k["x"][i]), f32(v - k["y"][i])
                e2 = f32(f32(ex * ex) + f32(ey * ey))
                bound = 5.99
                if ur is not None and ur[i] >= 0:
                    er = f32(f32(q["ur"][j]) - ur[i])
                    e2 = f32(e2 + f32(er * er))
                    bound = 7.8
                if float(f32(e2 * inv_sigma2[lvl])) > bound:
                    continue
            d = int(np.count_nonzero(bits_q[j] != bits_k[i]))
            if d < best:
                best, best_i = d, i
        bd[j] = best
        if best <= th:
            bi[j] = best_i
    return bi, bd


@pytest.mark.parametrize("seed,chi2,stereo,th", [(1, True, True, 50), (2, True, False, 50), (3, False, False, 100), (4, False, True, 50)])
def test_oracle_equals_python_statement(oracle, seed, chi2, stereo, th):
    c = _case(seed, 400, 350, stereo)
    sig = INV_SIGMA2 if chi2 else None
    got = oracle.fuse_search(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], th, c["uright"], sig)
    want = _python_rule(c, th, sig)
    assert np.array_equal(got[1], want[1]) and np.array_equal(got[0], want[0])
    assert (want[0] >= 0).sum() > (20 if chi2 else 100)
    if chi2:   # the chi-square test must actually remove candidates
        loose = oracle.fuse_search(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], th, c["uright"], None)
        assert (loose[0] >= 0).sum() > (want[0] >= 0).sum()


@pytest.mark.gpu
@pytest.mark.parametrize("chi2,stereo,th", [(True, True, 50), (True, False, 50), (False, False, 100), (False, True, 50)])
def test_gpu_equals_oracle(oracle, chi2, stereo, th):
    import orb_slam2_detailed_comments_b200 as pkg
    sizes = [(600, 500), (2000, 1900), (1, 3), (300, 0), (0, 200), (901, 1203)]
    cap, qcap = 2003, 1907
    F = len(sizes)
    rng = np.random.default_rng(5)
    K, D = np.zeros((F, cap), KP_DTYPE), rng.integers(0, 256, (F, cap, 32), dtype=np.uint8)
    Q, QD = np.zeros((F, qcap), QUERY_DTYPE), rng.integers(0, 256, (F, qcap, 32), dtype=np.uint8)
    U = np.full((F, cap), -1, f32)
    cases = []
    for i, (n, m) in enumerate(sizes):
        c = _case(20 + i, max(n, 1), max(m, 1), stereo)
        c = dict(c, kps=c["kps"][:n], desc=c["desc"][:n], queries=c["queries"][:m], qdesc=c["qdesc"][:m],
                 uright=None if c["uright"] is None else c["uright"][:n])
        K[i, :n], D[i, :n], Q[i, :m], QD[i, :m] = c["kps"], c["desc"], c["queries"], c["qdesc"]
        if stereo:
            U[i, :n] = c["uright"]
        cases.append(c)
    nk, nq = np.array([s[0] for s in sizes], np.int32), np.array([s[1] for s in sizes], np.int32)
    sig = INV_SIGMA2 if chi2 else None
    bi, bd = pkg.ORBmatcher(0.6, True).FuseSearch(K, D, nk, Q, QD, nq, cases[0]["bounds"], uright=U if stereo else None, inv_level_sigma2=sig,
                                                  th_dist=th)
    total = 0
    for i, c in enumerate(cases):
        m = len(c["queries"])
        want = oracle.fuse_search(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], th, c["uright"], sig)
        assert np.array_equal(bd[i, :m], want[1]), f"frame {i}: bestDist"
        assert np.array_equal(bi[i, :m], want[0]), f"frame {i}: bestIdx"
        assert np.all(bi[i, m:] == -1)
        total += int((want[0] >= 0).sum())
    assert total > (100 if chi2 else 800)
