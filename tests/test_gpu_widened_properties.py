"""Full-size runs (the bench.py shapes: 512 frames x 2000 features) of the widened rows, checked through size-independent
properties instead of the (slow) oracle, plus an oracle spot check on a few frames."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

pytestmark = pytest.mark.gpu
F, N = 512, 2000


def test_guided_search_full_size_properties(oracle):
    import orb_slam2_detailed_comments_b200 as pkg
    cases = [synth.synth_guided_case(900 + s, n=N, m=N, w=1241, h=376, radius=15.0, stereo=True) for s in range(8)]
    pick = [cases[i % 8] for i in range(F)]
    st = lambda k: np.stack([c[k] for c in pick])
    nk = np.full(F, N, np.int32)
    m = pkg.ORBmatcher(0.9, True)
    assigned, qmatch, nm = m.SearchByProjection(st("kps"), st("desc"), nk, st("queries"), st("qdesc"), nk, cases[0]["bounds"],
                                                uright=st("uright"), claimed=st("claimed"))
    for f in range(F):
        a, q = assigned[f], qmatch[f]
        t = np.flatnonzero(a >= 0)
        assert (q[a[t]] == t).all()                          # every standing assignment is some query's recorded match
        assert not st("claimed")[f][t].any()                 # caller-claimed keypoints are never assigned
        assert nm[f] == np.count_nonzero(q >= 0)             # one standing query record per counted match
        qs = np.flatnonzero(q >= 0)
        flags = pick[f]["queries"]["flags"]
        cl = qs[(flags[qs] & 2) != 0]
        assert len(np.unique(q[cl])) == len(cl)              # a keypoint claimed by one query is never matched by another claiming query
        assert not (flags[qs] & 1).any()                     # skipped queries never match
    assert np.array_equal(assigned[0], assigned[8]) and np.array_equal(qmatch[3], qmatch[11])   # identical inputs, identical outputs
    for f in (0, 5):
        c = pick[f]
        want = oracle.search_by_projection(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], 100, 0.0, True, c["uright"], c["claimed"])
        assert nm[f] == want[2] and np.array_equal(assigned[f], want[0]) and np.array_equal(qmatch[f], want[1])


def test_bow_full_size_properties(oracle):
    import orb_slam2_detailed_comments_b200 as pkg
    voc = synth.synth_vocabulary_full(10, 5, seed=5)   # 10^5 words
    gv = pkg.ORBVocabulary(*voc)
    rng = np.random.default_rng(17)
    leaves = np.flatnonzero(voc[3])
    base = voc[4][leaves[rng.integers(0, len(leaves), (8, N))]]
    base ^= np.packbits(rng.random((8, N, 256)) < 0.05, axis=2)
    desc = np.stack([base[i % 8] for i in range(F)])
    r = gv.transform(desc, np.full(F, N, np.int32), 4)
    assert (r["word"] >= 0).all() and (r["word"] < gv.size()).all()
    for f in range(0, F, 37):
        bc, fc = int(r["bow_count"][f]), int(r["fv_count"][f])
        w, v = r["bow_word"][f, :bc], r["bow_value"][f, :bc]
        assert (np.diff(w) > 0).all() and (v > 0).all() and abs(v.sum() - 1.0) < 1e-12     # ascending words, L1-normalised
        assert set(w.tolist()) == set(r["word"][f].tolist())                                # no stopped words in this vocabulary
        nd, stt = r["fv_node"][f, :fc], r["fv_start"][f, :fc + 1]
        assert (np.diff(nd) > 0).all() and stt[0] == 0 and stt[-1] == N and (np.diff(stt) > 0).all()
        feats = r["fv_feature"][f, :N]
        assert np.array_equal(np.sort(feats), np.arange(N))                                 # every feature in exactly one node
        for j in range(fc):
            seg = feats[stt[j]:stt[j + 1]]
            assert (np.diff(seg) > 0).all() and (r["node"][f, seg] == nd[j]).all()
    assert np.array_equal(r["bow_value"][1].view(np.uint64), r["bow_value"][9].view(np.uint64))
    ov = oracle.OracleVocabulary(*voc)
    w0, n0, bow0, fv0 = ov.transform(desc[3], 4)
    gb, gf = pkg.ORBVocabulary.frame_maps(r, 3)
    assert gb == bow0 and gf == fv0


def test_distinctive_descriptors_many_points(oracle):
    import orb_slam2_detailed_comments_b200 as pkg
    rng = np.random.default_rng(2)
    sizes = rng.integers(1, 40, 100_000)
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    desc = rng.integers(0, 256, (int(offsets[-1]), 32), dtype=np.uint8)
    bi, bm = pkg.ORBmatcher.ComputeDistinctiveDescriptors(desc, offsets)
    assert (bi >= 0).all() and (bi < sizes).all() and (bm[sizes == 1] == 0).all()
    wi, wm = oracle.distinctive_descriptors(desc[:offsets[3000]], offsets[:3001], nthreads=8)
    assert np.array_equal(bi[:3000], wi) and np.array_equal(bm[:3000], wm)
