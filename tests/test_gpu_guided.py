"""GPU: orb_search_by_projection (grid + masked top-2 + sequential claims + rotation filter) against the CPU oracle,
bit-exact, through the C ABI. Scenarios: tests/guided_cases.py."""
import ctypes as C

import numpy as np
import pytest

from guided_cases import make_case, KP_DTYPE

pytestmark = pytest.mark.gpu


def _batch(cases):
    """Pack ragged cases into the [F,cap] / [F,qcap] arrays of the C ABI (padding filled with junk that must be ignored)."""
    from oracle.oracle_py import QUERY_DTYPE
    F = len(cases)
    cap = max(len(c["kps"]) for c in cases) + 3
    qcap = max(len(c["queries"]) for c in cases) + 5
    rng = np.random.default_rng(99)
    kps = np.zeros((F, cap), KP_DTYPE)
    kps["x"], kps["y"] = rng.uniform(0, 700, (F, cap)), rng.uniform(0, 400, (F, cap))
    desc = rng.integers(0, 256, (F, cap, 32), dtype=np.uint8)
    q = np.zeros((F, qcap), QUERY_DTYPE)
    q["u"], q["v"], q["radius"] = 100, 100, 50
    qd = rng.integers(0, 256, (F, qcap, 32), dtype=np.uint8)
    ur = np.full((F, cap), -1, np.float32)
    cl = np.zeros((F, cap), np.uint8)
    nk, nq = np.zeros(F, np.int32), np.zeros(F, np.int32)
    for f, c in enumerate(cases):
        n, m = len(c["kps"]), len(c["queries"])
        nk[f], nq[f] = n, m
        kps[f, :n], desc[f, :n], q[f, :m], qd[f, :m] = c["kps"], c["desc"], c["queries"], c["qdesc"]
        if c["uright"] is not None:
            ur[f, :n] = c["uright"]
        cl[f, :n] = c["claimed"]
    return kps, desc, ur, cl, nk, q, qd, nq


def _oracle(oracle, c, th, ratio, ori, stereo, claims):
    return oracle.search_by_projection(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], th, ratio, ori,
                                       (c["uright"] if c["uright"] is not None else np.full(len(c["kps"]), -1, np.float32)) if stereo else None,
                                       c["claimed"] if claims else None)


CASES = [
    dict(), dict(radius=40.0, dup=0.5), dict(stereo=True), dict(claims=0.3), dict(n=40, m=300, radius=200.0), dict(rot=0.0),
    dict(n=2100, m=1800, w=1241, h=376), dict(n=0, m=30), dict(n=30, m=0), dict(n=700, m=640, skip=1.0),
]


@pytest.mark.parametrize("use_ratio,ori,th", [(False, True, 100), (True, False, 100), (False, False, 50)])
@pytest.mark.parametrize("stereo,claims", [(False, False), (True, True)])
def test_batch_matches_oracle(oracle, use_ratio, ori, th, stereo, claims):
    import orb_slam2_detailed_comments_b200 as pkg
    cases = []
    for s, kw in enumerate(CASES):
        kw = dict(kw)
        n = kw.pop("n", 600)
        c = make_case(100 + s, n=max(n, 1), **kw)
        if n == 0:
            c["kps"], c["desc"], c["claimed"] = c["kps"][:0], c["desc"][:0], c["claimed"][:0]
            c["uright"] = None
        cases.append(c)
    bounds = cases[0]["bounds"]
    for c in cases:
        c["bounds"] = bounds   # one image box per call
    kps, desc, ur, cl, nk, q, qd, nq = _batch(cases)
    m = pkg.ORBmatcher(0.8, ori)
    assigned, qmatch, nm = m.SearchByProjection(kps, desc, nk, q, qd, nq, bounds, th_dist=th, use_ratio=use_ratio,
                                                uright=ur if stereo else None, claimed=cl if claims else None)
    rescans = 0
    for f, c in enumerate(cases):
        want = _oracle(oracle, c, th, 0.8 if use_ratio else 0.0, ori and not use_ratio, stereo, claims)
        n, mq = len(c["kps"]), len(c["queries"])
        assert nm[f] == want[2], f"frame {f}: nmatches {nm[f]} vs oracle {want[2]}"
        assert np.array_equal(assigned[f, :n], want[0]), f"frame {f}: assigned differs"
        assert np.array_equal(qmatch[f, :mq], want[1]), f"frame {f}: query_match differs"
        assert (assigned[f, n:] == -1).all() and (qmatch[f, mq:] == -1).all()
        rescans += want[2]
    assert rescans > 1000


def test_device_memory_mode_equals_host_mode(oracle):
    import torch
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import lib, check, OrbSearchParams, MEM_DEVICE
    cases = [make_case(200 + s, radius=25.0, dup=0.4, stereo=True) for s in range(6)]
    bounds = cases[0]["bounds"]
    kps, desc, ur, cl, nk, q, qd, nq = _batch(cases)
    want = pkg.ORBmatcher(0.7, True).SearchByProjection(kps, desc, nk, q, qd, nq, bounds, uright=ur, claimed=cl)
    F, cap = kps.shape
    qcap = q.shape[1]
    dev = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1).copy()).cuda()
    t = [dev(a) for a in (kps, desc, ur, cl, q, qd)]
    tnk, tnq = torch.from_numpy(nk).cuda(), torch.from_numpy(nq).cuda()
    a = torch.empty((F, cap), dtype=torch.int32, device="cuda")
    qm = torch.empty((F, qcap), dtype=torch.int32, device="cuda")
    nm = torch.empty((F,), dtype=torch.int32, device="cuda")
    prm = OrbSearchParams(*[float(b) for b in bounds], 100, 0.0, 1, 0)
    st = torch.cuda.current_stream().cuda_stream
    p = lambda x: C.c_void_p(x.data_ptr())
    check(lib().orb_search_by_projection(MEM_DEVICE, p(t[0]), p(t[1]), p(t[2]), p(t[3]), p(tnk), cap, p(t[4]), p(t[5]), p(tnq), qcap, F,
                                         C.byref(prm), p(a), p(qm), p(nm), C.c_void_p(st)))
    torch.cuda.synchronize()
    assert np.array_equal(a.cpu().numpy(), want[0]) and np.array_equal(qm.cpu().numpy(), want[1]) and np.array_equal(nm.cpu().numpy(), want[2])


def test_on_extractor_output(oracle):
    """Frame-to-frame tracking shape: queries = keypoints of frame A (pure image shift), target = extraction of frame B."""
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200 import synth
    from oracle.oracle_py import QUERY_DTYPE
    W, H, shift = 752, 480, 7
    a, b = synth.synth_stereo_pair(W, H, seed=5, shift=shift)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    kps, desc, counts, _ = ex.extract_batch(np.stack([a, b]))
    na, nb = int(counts[0]), int(counts[1])
    q = np.zeros((1, na), QUERY_DTYPE)
    q["u"][0], q["v"][0] = kps["x"][0, :na] - shift, kps["y"][0, :na]
    sf = np.float32(1.2) ** kps["octave"][0, :na]
    q["radius"][0] = np.float32(15) * sf.astype(np.float32)
    q["angle"][0] = kps["angle"][0, :na]
    q["min_level"][0], q["max_level"][0] = kps["octave"][0, :na] - 1, kps["octave"][0, :na] + 1
    q["flags"][0] = 2
    bounds = (0.0, 0.0, float(W), float(H))
    m = pkg.ORBmatcher(0.9, True)
    assigned, qmatch, nm = m.SearchByProjection(kps[1:2], desc[1:2], counts[1:2], q, desc[0:1, :na], np.array([na], np.int32), bounds)
    want = oracle.search_by_projection(kps[1, :nb], desc[1, :nb], q[0], desc[0, :na], bounds, 100, 0.0, True)
    assert nm[0] == want[2] and np.array_equal(assigned[0, :nb], want[0]) and np.array_equal(qmatch[0], want[1])
    assert nm[0] > na // 3   # most keypoints of a shifted frame are tracked


def test_bad_arguments():
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import OrbError
    c = make_case(1, n=10, m=10)
    kps, desc, ur, cl, nk, q, qd, nq = _batch([c])
    with pytest.raises(OrbError):
        pkg.ORBmatcher().SearchByProjection(kps, desc, nk, q, qd, nq, (0.0, 0.0, 0.0, 480.0))
    with pytest.raises(OrbError):
        pkg.ORBmatcher().SearchByProjection(kps, desc, nk, q, qd, nq, c["bounds"], th_dist=300)
