"""ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:570-712): oracle vs an independent Python statement (CPU), CUDA path vs
oracle (GPU). Queries: level-0 keypoints of F1 searching a window around their previous match in F2."""
import numpy as np
import pytest

import guided_ref as R
from orb_slam2_detailed_comments_b200 import synth


def _case(seed, n=700, m=650, radius=100.0, **kw):
    c = synth.synth_guided_case(seed, n=n, m=m, radius=radius, levels=3, dup=0.3, skip=0.0, **kw)
    q = c["queries"]
    q["radius"] = radius                        # windowSize (100 in Tracking::MonocularInitialization)
    q["min_level"], q["max_level"] = 0, 0       # GetFeaturesInArea(..., level1, level1) with level1 == 0
    q["flags"] = np.where(np.random.default_rng(seed).random(len(q)) < 0.35, 1, 0)   # F1 keypoints of level > 0 are skipped
    return c


@pytest.mark.parametrize("seed,ori", [(1, True), (2, False), (3, True)])
def test_oracle_equals_python_statement(oracle, seed, ori):
    c = _case(seed, n=260, m=240)
    want = R.search_for_initialization(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], 50, 0.9, ori)
    got = oracle.search_for_initialization(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], 50, 0.9, ori)
    assert got[2] == want[2] and np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]) and want[2] > 30


@pytest.mark.gpu
@pytest.mark.parametrize("ori", [True, False])
def test_gpu_equals_oracle(oracle, ori):
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import KP_DTYPE, QUERY_DTYPE
    cases = [_case(10, 700, 650), _case(11, 2000, 1900, radius=60.0), _case(12, 40, 300, radius=400.0), _case(13, 5, 0), _case(14, 900, 900, radius=15.0)]
    cases[3]["queries"], cases[3]["qdesc"] = cases[3]["queries"][:0], cases[3]["qdesc"][:0]
    F, cap, qcap = len(cases), 2003, 1907
    rng = np.random.default_rng(0)
    kps = np.zeros((F, cap), KP_DTYPE); kps["x"], kps["y"] = rng.uniform(0, 700, (F, cap)), rng.uniform(0, 400, (F, cap))
    desc = rng.integers(0, 256, (F, cap, 32), dtype=np.uint8)
    q = np.zeros((F, qcap), QUERY_DTYPE); q["u"], q["v"], q["radius"] = 100, 100, 50
    qd = rng.integers(0, 256, (F, qcap, 32), dtype=np.uint8)
    nk, nq = np.zeros(F, np.int32), np.zeros(F, np.int32)
    for f, c in enumerate(cases):
        n, m = len(c["kps"]), len(c["queries"])
        nk[f], nq[f] = n, m
        kps[f, :n], desc[f, :n], q[f, :m], qd[f, :m] = c["kps"], c["desc"], c["queries"], c["qdesc"]
    m12, m21, nm = pkg.ORBmatcher(0.9, ori).SearchForInitialization(kps, desc, nk, q, qd, nq, cases[0]["bounds"])
    total = 0
    for f, c in enumerate(cases):
        want = oracle.search_for_initialization(c["kps"], c["desc"], c["queries"], c["qdesc"], cases[0]["bounds"], 50, 0.9, ori)
        n, m = len(c["kps"]), len(c["queries"])
        assert nm[f] == want[2], f"pair {f}: {nm[f]} vs {want[2]}"
        assert np.array_equal(m12[f, :m], want[0]) and np.array_equal(m21[f, :n], want[1]), f"pair {f}"
        assert (m12[f, m:] == -1).all() and (m21[f, n:] == -1).all()
        total += want[2]
    assert total > 300
