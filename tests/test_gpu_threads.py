"""Thread-safety the reference relies on (SURVEY.md §8(b) "Threading"): the stereo Frame constructor runs mpORBextractorLeft and
mpORBextractorRight from two std::threads at once (src/Frame.cc:158-168), and Tracking, LocalMapping and LoopClosing call the
matchers concurrently. Two / four host threads drive separate handles (extractors) and the shared per-device matcher workspace
(kNN, guided search, BoW transform) at the same time — the ctypes calls release the GIL — and every result must equal the one the
same call gives when it runs alone."""
import threading

import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

pytestmark = pytest.mark.gpu


def _run_threads(fns):
    out, err = [None] * len(fns), []
    start = threading.Barrier(len(fns))

    def work(i):
        try:
            start.wait()
            out[i] = fns[i]()
        except BaseException as e:   # noqa: BLE001 - reported by the assert below
            err.append((i, repr(e)))
    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(fns))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not err, err
    return out


def test_two_extractors_from_two_threads():
    """Left and right extractor of a stereo Frame (different handles, different image sizes on purpose), 30 rounds each."""
    import orb_slam2_detailed_comments_b200 as pkg
    left = synth.synth_batch(752, 480, 6, seed0=4000, distinct=6)
    right = synth.synth_batch(640, 360, 5, seed0=5000, distinct=5)
    exl, exr = pkg.ORBextractor(1200, 1.2, 8, 20, 7), pkg.ORBextractor(900, 1.2, 8, 20, 7)
    want_l = [exl(f) for f in left] + [exl.extract_batch(left)]
    want_r = [exr(f) for f in right] + [exr.extract_batch(right)]

    def loop(ex, frames, want):
        def run():
            for it in range(30):
                i = it % len(frames)
                k, d = ex(frames[i])
                assert np.array_equal(k, want[i][0]) and np.array_equal(d, want[i][1]), f"single image {i}, round {it}"
                if it % 10 == 0:
                    kps, desc, counts, lc = ex.extract_batch(frames)
                    wk, wd, wc, wl = want[-1]
                    assert np.array_equal(counts, wc) and np.array_equal(lc, wl)
                    for j, n in enumerate(counts):
                        assert np.array_equal(kps[j, :n], wk[j, :n]) and np.array_equal(desc[j, :n], wd[j, :n]), f"batch frame {j}, round {it}"
            return True
        return run
    assert _run_threads([loop(exl, left, want_l), loop(exr, right, want_r)]) == [True, True]


def test_one_extractor_shared_by_two_threads_is_serialised():
    """The reference's extractor is not re-entrant; the handle's mutex makes concurrent calls on ONE handle safe (they queue)."""
    import orb_slam2_detailed_comments_b200 as pkg
    frames = synth.synth_batch(640, 480, 4, seed0=6000, distinct=4)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    want = [ex(f) for f in frames]

    def loop(off):
        def run():
            for it in range(20):
                i = (it + off) % 4
                k, d = ex(frames[i])
                assert np.array_equal(k, want[i][0]) and np.array_equal(d, want[i][1]), f"image {i}, round {it}"
            return True
        return run
    assert _run_threads([loop(0), loop(2)]) == [True, True]


def test_matchers_from_four_threads(oracle):
    """kNN, SearchByProjection, the BoW transform and an extractor at the same time (Tracking + LocalMapping + LoopClosing)."""
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200 import matcher
    from guided_cases import make_case
    from test_gpu_guided import _batch
    rng = np.random.default_rng(3)
    db = rng.integers(0, 256, (60000, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    q[::5] = db[rng.integers(0, len(db), len(q[::5]))]
    want_knn = matcher.hamming_knn(q, db)
    cases = [make_case(300 + s, n=500 + 100 * s) for s in range(4)]
    kps, desc, ur, cl, nk, qq, qd, nq = _batch(cases)
    m = pkg.ORBmatcher(0.7, True)
    args = (kps, desc, nk, qq, qd, nq, cases[0]["bounds"])
    want_sp = m.SearchByProjection(*args, th_dist=100)
    voc = synth.synth_vocabulary(10, 3, seed=5)
    gv = pkg.ORBVocabulary(*voc)
    bd = rng.integers(0, 256, (6, 900, 32), dtype=np.uint8)
    bn = np.array([900, 1, 0, 450, 899, 37], np.int32)
    want_bow = gv.transform(bd, bn, 3)
    frames = synth.synth_batch(640, 480, 3, seed0=7000, distinct=3)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    want_ex = [ex(f) for f in frames]

    def eq(a, b):
        if isinstance(a, dict):
            return all(eq(a[k], b[k]) for k in a)
        if isinstance(a, (tuple, list)):
            return all(eq(x, y) for x, y in zip(a, b))
        return np.array_equal(a, b)

    def t_knn():
        return all(eq(matcher.hamming_knn(q, db), want_knn) for _ in range(25))

    def t_sp():
        return all(eq(m.SearchByProjection(*args, th_dist=100), want_sp) for _ in range(25))

    def t_bow():
        # entries past n[f] of the per-feature arrays are not written: compare the defined part
        ok = True
        for _ in range(25):
            r = gv.transform(bd, bn, 3)
            for f, n in enumerate(bn):
                ok = ok and np.array_equal(r["word"][f, :n], want_bow["word"][f, :n]) and r["bow_count"][f] == want_bow["bow_count"][f]
                c = int(r["bow_count"][f])
                ok = ok and np.array_equal(r["bow_word"][f, :c], want_bow["bow_word"][f, :c])
                ok = ok and np.array_equal(r["bow_value"][f, :c], want_bow["bow_value"][f, :c])
        return ok

    def t_ex():
        return all(eq(ex(frames[i % 3]), want_ex[i % 3]) for i in range(25))
    assert _run_threads([t_knn, t_sp, t_bow, t_ex]) == [True] * 4


def test_device_mode_calls_on_two_streams_share_the_scratch_safely(oracle):
    """the matcher entry points keep their intermediates in one per-device scratch; a device-mode call returns while its
    kernels are queued, so a second device-mode call on ANOTHER stream used to overwrite partials that were still being read. The entry points
    now chain the scratch's users with an event (csrc/orb_api_match.cu: WsOrder). A long scan on stream A is followed at once by a short one
    on stream B whose partial keys land in the same words; both must equal the oracle, every round."""
    import torch
    from orb_slam2_detailed_comments_b200 import matcher
    dev = torch.device("cuda:0")
    q1, db1 = synth.synth_descriptors(2000, 300_000, seed=41)
    q2, db2 = synth.synth_descriptors(2000, 3_000, seed=42)
    want1, want2 = oracle.hamming_knn(q1, db1, nthreads=16), oracle.hamming_knn(q2, db2, nthreads=16)
    tq1, tdb1, tq2, tdb2 = (torch.from_numpy(a).to(dev) for a in (q1, db1, q2, db2))
    sa, sb = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    torch.cuda.synchronize(dev)
    for rnd in range(12):
        o1 = torch.zeros((2000, 3), dtype=torch.int64, device=dev)
        o2 = torch.zeros((2000, 3), dtype=torch.int64, device=dev)
        torch.cuda.synchronize(dev)
        if rnd % 2 == 0:
            matcher.hamming_knn_device(tq1, tdb1, out=o1, stream=sa)
            matcher.hamming_knn_device(tq2, tdb2, out=o2, stream=sb)
        else:
            matcher.hamming_knn_device(tq2, tdb2, out=o2, stream=sb)
            matcher.hamming_knn_device(tq1, tdb1, out=o1, stream=sa)
        torch.cuda.synchronize(dev)
        r1, r2 = matcher.knn_records(o1.cpu()), matcher.knn_records(o2.cpu())
        for f in ("d1", "d2", "i1", "i2"):
            assert np.array_equal(r1[f], want1[f]), (rnd, "long scan", f)
            assert np.array_equal(r2[f], want2[f]), (rnd, "short scan", f)
