"""ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (src/ORBmatcher.cc:248-410): oracle vs an independent Python statement (CPU),
orb_search_by_bow vs oracle (GPU). FeatureVectors come from the vocabulary transform of the same descriptors."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth
from orb_slam2_detailed_comments_b200._lib import KP_DTYPE


def _pair(voc, seed, n_kf=700, n_f=800, shared=0.7):
    """Keyframe / frame descriptor sets: `shared` of the frame's features are noisy copies of keyframe features."""
    k, L, parent, is_leaf, desc, weight = voc
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(is_leaf)
    kf = desc[leaves[rng.integers(0, len(leaves), n_kf)]].copy()
    kf ^= np.packbits(rng.random((n_kf, 256)) < 0.04, axis=1)
    f = desc[leaves[rng.integers(0, len(leaves), n_f)]].copy()
    src = rng.integers(0, n_kf, n_f)
    sh = rng.random(n_f) < shared
    f[sh] = kf[src[sh]]
    f ^= np.packbits(rng.random((n_f, 256)) < 0.03, axis=1)
    kk, fk = np.zeros(n_kf, KP_DTYPE), np.zeros(n_f, KP_DTYPE)
    kk["angle"] = rng.uniform(0, 360, n_kf)
    fk["angle"] = rng.uniform(0, 360, n_f)
    fk["angle"][sh] = (kk["angle"][src[sh]] - rng.normal(15, 3, int(sh.sum()))) % 360   # a dominant rotation for the true matches
    valid = (rng.random(n_kf) < 0.85).astype(np.uint8)
    return kf, kk, valid, f, fk


def _csr(fv: dict, cap):
    nodes = sorted(fv)
    node = np.full(cap, -7, np.int32); start = np.zeros(cap + 1, np.int32); feat = np.full(cap, -7, np.int32)
    pos = 0
    for j, nd in enumerate(nodes):
        node[j], start[j] = nd, pos
        feat[pos:pos + len(fv[nd])] = fv[nd]
        pos += len(fv[nd])
    start[len(nodes)] = pos
    return node, start, feat, len(nodes)


def _python_rule(kf, kk, valid, kfv, f, fk, ffv, th, ratio, ori, blocked=None):
    assigned = np.full(len(f), -1, np.int32)
    bits_k, bits_f = np.unpackbits(kf, axis=1), np.unpackbits(f, axis=1)
    events = []
    for node in sorted(set(kfv) & set(ffv)):
        for ik in kfv[node]:
            if not valid[ik]:
                continue
            cand = [i for i in ffv[node] if assigned[i] < 0 and not (blocked is not None and blocked[i])]
            if not cand:
                continue
            d = np.array([int(np.count_nonzero(bits_k[ik] != bits_f[i])) for i in cand])
            o = np.argsort(d, kind="stable")
            b1, b2 = int(d[o[0]]), int(d[o[1]]) if len(o) > 1 else 256
            if b1 <= th and np.float32(b1) < np.float32(ratio) * np.float32(b2):
                t = cand[o[0]]
                assigned[t] = ik
                rot = np.float32(kk["angle"][ik]) - np.float32(fk["angle"][t])
                if rot < 0:
                    rot = np.float32(rot + np.float32(360))
                v = float(np.float32(rot * (np.float32(30) / np.float32(360))))
                b = int(np.floor(abs(v) + 0.5))
                events.append((0 if b == 30 else b, t))
    nm = len(events)
    if ori:
        import guided_ref
        sizes = [sum(1 for b, _ in events if b == i) for i in range(30)]
        keep = guided_ref.three_maxima(sizes)
        for b, t in events:
            if b not in keep:
                assigned[t] = -1
                nm -= 1
    return assigned, nm


@pytest.mark.parametrize("seed,ori,ratio,kfkf", [(1, True, 0.7, False), (2, False, 0.6, False), (3, True, 0.9, False), (4, True, 0.8, True)])
def test_oracle_equals_python_statement(oracle, seed, ori, ratio, kfkf):
    voc = synth.synth_vocabulary(10, 3, seed=11)
    ov = oracle.OracleVocabulary(*voc)
    kf, kk, valid, f, fk = _pair(voc, seed, 300, 340)
    kfv, ffv = ov.transform(kf, 2)[3], ov.transform(f, 2)[3]   # levelsup 2 on L = 3: ten big nodes, long candidate lists
    blocked = (np.random.default_rng(seed).random(len(f)) < 0.2).astype(np.uint8) if kfkf else None   # KeyFrame-KeyFrame form
    th = 49 if kfkf else 50
    got = oracle.search_by_bow(kf, kk, valid, _csr(kfv, len(kf)), f, fk, _csr(ffv, len(f)), th, ratio, ori, blocked)
    want = _python_rule(kf, kk, valid, kfv, f, fk, ffv, th, ratio, ori, blocked)
    assert got[1] == want[1] and np.array_equal(got[0], want[0]) and want[1] > 30
    inv = np.full(len(kf), -1, np.int32)
    inv[want[0][want[0] >= 0]] = np.flatnonzero(want[0] >= 0)
    assert np.array_equal(got[2], inv)   # kf_match is the inverse map of assigned


@pytest.mark.gpu
@pytest.mark.parametrize("levelsup,ori,kfkf", [(1, True, False), (2, True, False), (0, False, False), (2, True, True)])
def test_gpu_equals_oracle(oracle, levelsup, ori, kfkf):
    import orb_slam2_detailed_comments_b200 as pkg
    voc = synth.synth_vocabulary(10, 3, seed=11)
    gv, ov = pkg.ORBVocabulary(*voc), oracle.OracleVocabulary(*voc)
    sizes = [(700, 800), (1500, 1400), (1, 5), (300, 0), (0, 300), (900, 901)]
    qcap, cap = 1503, 1407
    rng = np.random.default_rng(0)
    kf_desc = rng.integers(0, 256, (len(sizes), qcap, 32), dtype=np.uint8); f_desc = rng.integers(0, 256, (len(sizes), cap, 32), dtype=np.uint8)
    kf_kps, f_kps = np.zeros((len(sizes), qcap), KP_DTYPE), np.zeros((len(sizes), cap), KP_DTYPE)
    kf_valid = np.ones((len(sizes), qcap), np.uint8)
    pairs = []
    for i, (a, b) in enumerate(sizes):
        kf, kk, valid, f, fk = _pair(voc, 40 + i, max(a, 1), max(b, 1))
        kf, kk, valid, f, fk = kf[:a], kk[:a], valid[:a], f[:b], fk[:b]
        kf_desc[i, :a], kf_kps[i, :a], kf_valid[i, :a], f_desc[i, :b], f_kps[i, :b] = kf, kk, valid, f, fk
        pairs.append((kf, kk, valid, f, fk))
    nk, nf = np.array([s[0] for s in sizes], np.int32), np.array([s[1] for s in sizes], np.int32)
    kbow, fbow = gv.transform(kf_desc, nk, levelsup), gv.transform(f_desc, nf, levelsup)
    m = pkg.ORBmatcher(0.7, ori)
    blocked = (rng.random((len(sizes), cap)) < 0.2).astype(np.uint8) if kfkf else None
    th = 49 if kfkf else 50
    assigned, kf_match, nm = m.SearchByBoW(kf_desc, kf_kps, kf_valid, kbow, f_desc, f_kps, fbow, th_dist=th, f_blocked=blocked)
    total = 0
    for i, (kf, kk, valid, f, fk) in enumerate(pairs):
        kfv, ffv = ov.transform(kf, levelsup)[3], ov.transform(f, levelsup)[3]
        want = oracle.search_by_bow(kf, kk, valid, _csr(kfv, max(len(kf), 1)), f, fk, _csr(ffv, max(len(f), 1)), th, 0.7, ori,
                                    None if blocked is None else blocked[i, :len(f)])
        assert nm[i] == want[1], f"pair {i}: nmatches {nm[i]} vs {want[1]}"
        assert np.array_equal(assigned[i, :len(f)], want[0]), f"pair {i}: assignment differs"
        assert np.array_equal(kf_match[i, :len(kf)], want[2]) and (kf_match[i, len(kf):] == -1).all(), f"pair {i}: kf_match differs"
        assert (assigned[i, len(f):] == -1).all()
        total += want[1]
    assert total > 500


@pytest.mark.gpu
def test_malformed_feature_vector_rejected(oracle):
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import OrbError
    voc = synth.synth_vocabulary(6, 2, seed=4)
    gv = pkg.ORBVocabulary(*voc)
    kf, kk, valid, f, fk = _pair(voc, 9, 50, 60)
    kbow, fbow = gv.transform(kf[None], np.array([50], np.int32), 1), gv.transform(f[None], np.array([60], np.int32), 1)
    bad = {k: v.copy() for k, v in fbow.items()}
    bad["fv_feature"][0, 0] = 60   # out of range
    with pytest.raises(OrbError):
        pkg.ORBmatcher().SearchByBoW(kf[None], kk[None], valid[None], kbow, f[None], fk[None], bad)
    bad = {k: v.copy() for k, v in fbow.items()}
    bad["fv_start"][0, 1] = -3
    with pytest.raises(OrbError):
        pkg.ORBmatcher().SearchByBoW(kf[None], kk[None], valid[None], kbow, f[None], fk[None], bad)
