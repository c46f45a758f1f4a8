"""Frame::ComputeBoW = DBoW2 TemplatedVocabulary::transform. CPU: the C++ oracle against an independent Python
statement of the same rules and the ORBvoc.txt reader/writer round trip; GPU: orb_bow_transform against the oracle, doubles compared
bit for bit. No upstream fixture (ORBvoc.txt is not shipped with the reference tree): parity unpinned, cross-checked."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth


def _features(voc, n, seed):
    """Descriptors near vocabulary nodes (so different branches are taken) plus pure noise rows."""
    k, L, parent, is_leaf, desc, weight = voc
    rng = np.random.default_rng(seed)
    d = desc[rng.integers(1, len(desc), n)].copy()
    noise = rng.random((n, 256)) < 0.06
    d ^= np.packbits(noise, axis=1)
    d[::7] = rng.integers(0, 256, (len(d[::7]), 32), dtype=np.uint8)
    return d


def _python_transform(voc, feats, levelsup):
    k, L, parent, is_leaf, desc, weight = voc
    children = {}
    word_of, nwords = {}, 0
    for nid in range(1, len(parent)):
        children.setdefault(int(parent[nid]), []).append(nid)
        if is_leaf[nid]:
            word_of[nid] = nwords
            nwords += 1
    bits = np.unpackbits(desc, axis=1)
    fbits = np.unpackbits(feats, axis=1)
    words, nodes, bow, fv = [], [], {}, {}
    for i in range(len(feats)):
        fid, lvl, nid = 0, 0, 0
        while True:
            lvl += 1
            ch = children[fid]
            dist = [(int(np.count_nonzero(bits[c] != fbits[i]))) for c in ch]
            fid = ch[int(np.argmin(dist))]   # argmin = first minimum = "d < best_d"
            if lvl == L - levelsup:
                nid = fid
            if fid not in children:
                break
        w = float(weight[fid])
        words.append(word_of[fid]); nodes.append(nid)
        if w > 0:
            bow[word_of[fid]] = bow[word_of[fid]] + w if word_of[fid] in bow else w
            fv.setdefault(nid, []).append(i)
    norm = 0.0
    for wid in sorted(bow):
        norm += abs(bow[wid])
    if norm > 0:
        bow = {wid: bow[wid] / norm for wid in bow}
    return np.array(words, np.int32), np.array(nodes, np.int32), bow, fv


VOCS = [dict(k=10, L=3, seed=11), dict(k=4, L=5, seed=12, prune=0.15), dict(k=7, L=2, seed=13, stop=0.3)]


@pytest.mark.parametrize("vkw", VOCS)
@pytest.mark.parametrize("levelsup", [0, 1, 2, 4, 9])
def test_oracle_equals_python_statement(oracle, vkw, levelsup):
    voc = synth.synth_vocabulary(**vkw)
    ov = oracle.OracleVocabulary(*voc)
    feats = _features(voc, 260, 3)
    w, nd, bow, fv = ov.transform(feats, levelsup)
    pw, pn, pbow, pfv = _python_transform(voc, feats, levelsup)
    assert np.array_equal(w, pw) and np.array_equal(nd, pn)
    assert bow == pbow and fv == pfv   # float equality: same additions in the same order
    assert len(bow) > 20 and abs(sum(bow.values()) - 1) < 1e-9


def test_text_vocabulary_round_trip(tmp_path):
    from orb_slam2_detailed_comments_b200 import vocabulary
    k, L, parent, is_leaf, desc, weight = synth.synth_vocabulary(5, 3, seed=2)
    p = str(tmp_path / "voc.txt")
    vocabulary.write_text_vocabulary(p, k, L, parent, is_leaf, desc, weight)
    k2, L2, s, wt, parent2, leaf2, desc2, weight2 = vocabulary.read_text_vocabulary(p)
    assert (k2, L2, s, wt) == (k, L, 0, 0)
    assert np.array_equal(parent2, parent) and np.array_equal(leaf2, is_leaf) and np.array_equal(desc2[1:], desc[1:])
    assert np.array_equal(weight2, weight)   # repr() round-trips doubles exactly


@pytest.mark.gpu
@pytest.mark.parametrize("vkw", VOCS)
def test_gpu_equals_oracle(oracle, vkw):
    import orb_slam2_detailed_comments_b200 as pkg
    voc = synth.synth_vocabulary(**vkw)
    gv = pkg.ORBVocabulary(*voc)
    ov = oracle.OracleVocabulary(*voc)
    assert gv.size() == int(voc[3].sum())
    counts = [2000, 1, 0, 777, 2047, 33]
    cap = 2053
    desc = np.random.default_rng(0).integers(0, 256, (len(counts), cap, 32), dtype=np.uint8)   # padding rows must be ignored
    for f, n in enumerate(counts):
        desc[f, :n] = _features(voc, max(n, 1), 20 + f)[:n]
    for levelsup in (4, 1, 0):
        r = gv.transform(desc, np.array(counts, np.int32), levelsup)
        for f, n in enumerate(counts):
            w, nd, bow, fv = ov.transform(desc[f, :n], levelsup)
            assert np.array_equal(r["word"][f, :n], w) and np.array_equal(r["node"][f, :n], nd)
            gbow, gfv = pkg.ORBVocabulary.frame_maps(r, f)
            assert list(gbow.keys()) == list(bow.keys()), f"frame {f}: BowVector words"
            assert np.array_equal(np.array(list(gbow.values())).view(np.uint64), np.array(list(bow.values())).view(np.uint64)), \
                f"frame {f}: BowVector values differ in some bit"
            assert gfv == fv, f"frame {f}: FeatureVector"


@pytest.mark.gpu
def test_gpu_on_extractor_output_and_text_file(oracle, tmp_path):
    """ComputeBoW on real extractor descriptors with a vocabulary that went through the ORBvoc.txt format."""
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200 import vocabulary
    voc = synth.synth_vocabulary(10, 4, seed=21)
    p = str(tmp_path / "voc.txt")
    vocabulary.write_text_vocabulary(p, *voc)
    gv = pkg.ORBVocabulary.loadFromTextFile(p)
    ov = oracle.OracleVocabulary(*voc)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    frames = synth.synth_batch(640, 480, 3, seed0=77, distinct=3)
    kps, desc, counts, _ = ex.extract_batch(frames)
    r = gv.transform(desc, counts, 4)
    for f in range(3):
        w, nd, bow, fv = ov.transform(desc[f, :counts[f]], 4)
        gbow, gfv = pkg.ORBVocabulary.frame_maps(r, f)
        assert gbow == bow and gfv == fv and len(bow) > 100


@pytest.mark.gpu
def test_bad_vocabulary_rejected():
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import OrbError
    k, L, parent, is_leaf, desc, weight = synth.synth_vocabulary(4, 2, seed=3)
    bad = parent.copy(); bad[3] = len(parent) + 5
    with pytest.raises(OrbError):
        pkg.ORBVocabulary(k, L, bad, is_leaf, desc, weight)
    leaf = is_leaf.copy(); leaf[1] = 1   # node 1 has children in this tree
    with pytest.raises(OrbError):
        pkg.ORBVocabulary(k, L, parent, leaf, desc, weight)
