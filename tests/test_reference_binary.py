"""Pins the oracle (and, on the GPU box, the CUDA path) against the REFERENCE'S OWN CODE: oracle/_ref holds
src/ORBextractor.cc and Thirdparty/DBoW2 of the reference, compiled unmodified against oracle/cvshim (`make -C oracle _ref`,
run by __graft_entry__.build() where /root/reference exists; the binaries travel to the GPU box with the snapshot).

What is the reference here and what is not: control flow, float/int arithmetic, containers and ordering of
ORBextractor::operator() / ComputePyramid / ComputeKeyPointsOctTree / DistributeOctTree / IC_Angle / computeOrbDescriptor and of
DBoW2's loadFromTextFile / transform are the reference's; the five OpenCV primitives under them are oracle/cv_primitives.cpp,
which tests/test_oracle_golden.py pins bit-exact against the real cv2 4.13."""
import os

import numpy as np
import pytest

import ref_binary
import test_bow
from orb_slam2_detailed_comments_b200 import synth, vocabulary

needs_ref = pytest.mark.skipif(not ref_binary.available(), reason="oracle/_ref/orb_ref_extract not built (needs /root/reference at build time)")
needs_ref_bow = pytest.mark.skipif(not ref_binary.bow_available(), reason="oracle/_ref/orb_ref_bow not built")

# name, width, height, nfeatures, frames, seed, (scale, levels, ini, min)
CASES = [
    ("tum", 640, 480, 1000, 3, 1234, (1.2, 8, 20, 7)),
    ("tum_init", 640, 480, 2000, 2, 1240, (1.2, 8, 20, 7)),
    ("kitti", 1241, 376, 2000, 2, 1236, (1.2, 8, 20, 7)),
    ("kitti_ini12", 1241, 376, 2000, 1, 1250, (1.2, 8, 12, 7)),
    ("euroc", 752, 480, 1200, 2, 1237, (1.2, 8, 20, 7)),
    ("odd", 333, 257, 700, 2, 1239, (1.2, 8, 20, 7)),
    ("coarse", 480, 360, 300, 2, 77, (1.5, 4, 20, 7)),
    ("strict", 480, 360, 600, 1, 78, (1.2, 8, 40, 25)),
    # tiny quotas (levels with a quota of 0 or 1; N < 4 * nIni in DistributeOctTree) and a 5:1 image (nIni = 5 quadtree roots)
    ("nf5", 640, 480, 5, 1, 90, (1.2, 8, 20, 7)),
    ("nf20", 640, 480, 20, 1, 91, (1.2, 8, 20, 7)),
    ("nf50", 752, 480, 50, 1, 92, (1.2, 8, 20, 7)),
    ("wide5to1", 1000, 200, 500, 1, 93, (1.2, 6, 20, 7)),
]


def _check_frame(ref, kps, desc, levels=None):
    rk, rd, rl = ref
    assert len(kps) == len(rk), f"keypoint count {len(kps)} vs reference {len(rk)}"
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kps[f], rk[f]), f"keypoint field {f}"
    assert np.array_equal(kps["angle"].view(np.uint32), rk["angle"].view(np.uint32)), "angle bits"
    assert np.array_equal(desc, rd), f"descriptor bits differing: {int(np.unpackbits(desc ^ rd).sum())}"
    if levels is not None:
        for l, (a, b) in enumerate(zip(levels, rl)):
            assert np.array_equal(a, b), f"mvImagePyramid[{l}]"


@needs_ref
@pytest.mark.parametrize("name,w,h,nf,n,seed,prm", CASES, ids=[c[0] for c in CASES])
def test_oracle_equals_reference_extractor(oracle, name, w, h, nf, n, seed, prm):
    sf, nl, ini, mn = prm
    frames = synth.synth_batch(w, h, n, seed0=seed, distinct=n)
    tabs, ref = ref_binary.ref_extract(frames, nf, sf, nl, ini, mn)
    oe = oracle.OracleExtractor(nf, sf, nl, ini, mn)
    for t in range(4):
        assert np.array_equal(oe.table(t).view(np.uint32), tabs[t].view(np.uint32)), f"scale table {t}"
    total = 0
    for i in range(n):
        kps, desc, _ = oe(frames[i])
        _check_frame(ref[i], kps, desc, [oe.level(l) for l in range(nl)])
        total += len(kps)
    assert total >= n * nf * 0.8


@needs_ref
def test_oracle_equals_reference_extractor_noise_flat_legacy(oracle):
    rng = np.random.RandomState(3)
    frames = np.stack([rng.randint(0, 256, (240, 320), dtype=np.uint8), np.full((240, 320), 128, np.uint8)])
    _, ref = ref_binary.ref_extract(frames, 500)
    oe = oracle.OracleExtractor(500, 1.2, 8, 20, 7)
    for i in range(2):
        kps, desc, _ = oe(frames[i])
        _check_frame(ref[i], kps, desc)
    assert len(ref[1][0]) == 0
    frames = synth.synth_batch(640, 480, 1, seed0=5)
    _, ref = ref_binary.ref_extract(frames, 1000, legacy_blur=True)
    kps, desc, _ = oracle.OracleExtractor(1000, 1.2, 8, 20, 7, flags=oracle.LEGACY_BLUR)(frames[0])
    _check_frame(ref[0], kps, desc)


needs_ref_upstream = pytest.mark.skipif(not ref_binary.upstream_available(), reason="oracle/_ref/orb_ref_extract_upstream not built")
UPSTREAM_CASES = [c for c in CASES if c[0] in ("tum", "kitti", "euroc", "odd", "coarse", "nf20", "wide5to1")]


@needs_ref_upstream
@pytest.mark.parametrize("name,w,h,nf,n,seed,prm", UPSTREAM_CASES, ids=[c[0] for c in UPSTREAM_CASES])
def test_oracle_upstream_pyramid_equals_reference_without_fork_line(oracle, name, w, h, nf, n, seed, prm):
    """SURVEY F1 switch: the oracle with UPSTREAM_PYRAMID == the reference's source minus `mvImagePyramid[level] = temp;` (src/ORBextractor.cc:2166),
    i.e. raulmur's ComputePyramid: mvImagePyramid[l] is the w_l x h_l interior, levels are resized from it, keypoints are in its coordinates."""
    sf, nl, ini, mn = prm
    frames = synth.synth_batch(w, h, n, seed0=seed, distinct=n)
    _, ref = ref_binary.ref_extract(frames, nf, sf, nl, ini, mn, upstream=True)
    _, ref_fork = ref_binary.ref_extract(frames[:1], nf, sf, nl, ini, mn)
    oe = oracle.OracleExtractor(nf, sf, nl, ini, mn, flags=oracle.UPSTREAM_PYRAMID)
    for i in range(n):
        kps, desc, _ = oe(frames[i])
        _check_frame(ref[i], kps, desc, [oe.level(l)[19:-19, 19:-19] for l in range(nl)])
    assert ref[0][2][0].shape == (h, w) and ref_fork[0][2][0].shape == (h + 38, w + 38)
    # the two pyramids really differ beyond level 0 (otherwise the switch would test nothing)
    assert not np.array_equal(ref[0][2][1], ref_fork[0][2][1][19:-19, 19:-19])


@pytest.mark.gpu
@needs_ref_upstream
@pytest.mark.parametrize("name,w,h,nf,n,seed,prm", UPSTREAM_CASES[:4], ids=[c[0] for c in UPSTREAM_CASES[:4]])
def test_cuda_upstream_pyramid_equals_reference_without_fork_line(name, w, h, nf, n, seed, prm):
    from orb_slam2_detailed_comments_b200 import ORBextractor
    from orb_slam2_detailed_comments_b200._lib import FLAG_UPSTREAM_PYRAMID
    sf, nl, ini, mn = prm
    frames = synth.synth_batch(w, h, n, seed0=seed, distinct=n)
    _, ref = ref_binary.ref_extract(frames, nf, sf, nl, ini, mn, upstream=True)
    ex = ORBextractor(nf, sf, nl, ini, mn, flags=FLAG_UPSTREAM_PYRAMID)
    kps, desc, counts, _ = ex.extract_batch(frames)
    for i in range(n):
        c = int(counts[i])
        _check_frame(ref[i], kps[i, :c], desc[i, :c], [ex.pyramid_level(l, index=i) for l in range(nl)])


def _bow_case(tmp_path, vkw, levelsup, sizes):
    voc = synth.synth_vocabulary(**vkw)
    p = str(tmp_path / "voc.txt")
    vocabulary.write_text_vocabulary(p, *voc, trailing_newline=False)   # see write_text_vocabulary: the reference's eof() loop
    feats = [test_bow._features(voc, max(n, 1), 3 + j)[:n] for j, n in enumerate(sizes)]
    return voc, feats, ref_binary.ref_bow(p, feats, levelsup)


def _check_bow(got_bow, got_fv, node, ref, nid_level):
    rbow, rfv = ref
    assert list(got_bow.keys()) == list(rbow.keys())
    assert np.array_equal(np.array(list(got_bow.values()), np.float64).view(np.uint64),
                          np.array(list(rbow.values()), np.float64).view(np.uint64)), "BowVector values differ in some bit"
    # DESIGN.md §2 canonicalisation 6: a feature whose walk ends in a leaf ABOVE level L - levelsup leaves `nid` uninitialised in
    # the reference (TemplatedVocabulary.h:1162, 1280); canonical value = 0 (root). Such features are left out of the comparison.
    early = set(np.nonzero(np.asarray(node) == 0)[0].tolist()) if nid_level > 0 else set()
    strip = lambda fv: {k: [i for i in v if i not in early] for k, v in fv.items() if any(i not in early for i in v)}
    assert strip(got_fv) == strip(rfv), "FeatureVector"
    return len(early)


@needs_ref_bow
@pytest.mark.parametrize("vkw", test_bow.VOCS + [dict(k=10, L=4, seed=21, prune=0.0)], ids=lambda v: f"k{v['k']}L{v['L']}")
@pytest.mark.parametrize("levelsup", [0, 1, 2, 4, 9])
def test_oracle_equals_reference_dbow2(oracle, tmp_path, vkw, levelsup):
    voc, feats, ref = _bow_case(tmp_path, vkw, levelsup, [260, 50, 0, 1])
    ov = oracle.OracleVocabulary(*voc)
    for f, r in zip(feats, ref):
        _, node, bow, fv = ov.transform(f, levelsup)
        early = _check_bow(bow, fv, node, r, voc[1] - levelsup)
        if vkw.get("prune", 0.05) == 0.0:
            assert early == 0 and fv == r[1]      # balanced tree: the whole FeatureVector is defined and equal


# ---- seeded sweep of random configurations (sizes, scale factors, level counts, quotas, thresholds, image statistics) ----

def _sweep_cases(count=24, seed=20260):
    """Deterministic list of (w, h, nf, sf, nl, ini, mn, kind, seed). The coarsest level keeps at least 64 pixels per side: below
    the 30-pixel FAST cell the reference's own cell arithmetic divides by zero (ORBextractor.cc:1085-1088)."""
    rng = np.random.RandomState(seed)
    out = []
    while len(out) < count:
        w, h = int(rng.randint(150, 900)), int(rng.randint(120, 560))
        sf = float(rng.choice([1.1, 1.2, 1.25, 1.33, 1.5, 2.0]))
        nl = int(rng.randint(1, 9))
        if min(w, h) / sf ** (nl - 1) < 64:
            continue
        ini = int(rng.randint(8, 61))
        out.append((w, h, int(rng.choice([30, 120, 500, 1000, 1500])), sf, nl, ini, int(rng.randint(1, ini + 1)),
                    str(rng.choice(["synth", "salted", "lowcontrast"])), int(rng.randint(1, 10000))))
    return out


def _sweep_frame(w, h, kind, seed):
    img = synth.synth_frame(w, h, seed)
    rng = np.random.RandomState(seed + 1)
    if kind == "salted":       # isolated extreme pixels: very high FAST scores, ties in the quadtree's response ordering
        img = np.where(rng.rand(h, w) < 0.01, rng.choice([0, 255], size=(h, w)), img).astype(np.uint8)
    elif kind == "lowcontrast":   # most cells fall through to minThFAST
        img = (96 + (img.astype(np.int32) - 96) // 6).astype(np.uint8)
    return np.ascontiguousarray(img)


SWEEP = _sweep_cases()


@needs_ref
@pytest.mark.parametrize("case", SWEEP, ids=[f"{c[0]}x{c[1]}_s{c[3]}_l{c[4]}_{c[7]}" for c in SWEEP])
def test_oracle_equals_reference_extractor_sweep(oracle, case):
    w, h, nf, sf, nl, ini, mn, kind, seed = case
    img = _sweep_frame(w, h, kind, seed)
    tabs, ref = ref_binary.ref_extract(img[None], nf, sf, nl, ini, mn)
    oe = oracle.OracleExtractor(nf, sf, nl, ini, mn)
    for t in range(4):
        assert np.array_equal(oe.table(t).view(np.uint32), tabs[t].view(np.uint32)), f"scale table {t}"
    kps, desc, _ = oe(img)
    _check_frame(ref[0], kps, desc, [oe.level(l) for l in range(nl)])


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("case", SWEEP, ids=[f"{c[0]}x{c[1]}_s{c[3]}_l{c[4]}_{c[7]}" for c in SWEEP])
def test_cuda_equals_reference_extractor_sweep(case):
    from orb_slam2_detailed_comments_b200 import ORBextractor
    w, h, nf, sf, nl, ini, mn, kind, seed = case
    img = _sweep_frame(w, h, kind, seed)
    _, ref = ref_binary.ref_extract(img[None], nf, sf, nl, ini, mn)
    ex = ORBextractor(nf, sf, nl, ini, mn)
    kps, desc, counts, _ = ex.extract_batch(img[None])
    c = int(counts[0])
    _check_frame(ref[0], kps[0, :c], desc[0, :c], [ex.pyramid_level(l) for l in range(nl)])


# ---- the CUDA path against the reference's own code (GPU box) ----

@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("name,w,h,nf,n,seed,prm", CASES, ids=[c[0] for c in CASES])
def test_cuda_equals_reference_extractor(name, w, h, nf, n, seed, prm):
    from orb_slam2_detailed_comments_b200 import ORBextractor
    sf, nl, ini, mn = prm
    frames = synth.synth_batch(w, h, n, seed0=seed, distinct=n)
    _, ref = ref_binary.ref_extract(frames, nf, sf, nl, ini, mn)
    ex = ORBextractor(nf, sf, nl, ini, mn)
    kps, desc, counts, _ = ex.extract_batch(frames)
    for i in range(n):
        c = int(counts[i])
        _check_frame(ref[i], kps[i, :c], desc[i, :c])
    kps1, desc1, _, _ = ex.extract_batch(frames[-1:])       # mvImagePyramid of the last call
    _check_frame(ref[-1], kps1[0, :int(counts[-1])], desc1[0, :int(counts[-1])], [ex.pyramid_level(l) for l in range(nl)])


@pytest.mark.gpu
@needs_ref_bow
@pytest.mark.parametrize("vkw", [dict(k=10, L=3, seed=11), dict(k=4, L=5, seed=12, prune=0.15), dict(k=10, L=4, seed=21, prune=0.0)],
                         ids=lambda v: f"k{v['k']}L{v['L']}")
def test_cuda_equals_reference_dbow2(tmp_path, vkw):
    import orb_slam2_detailed_comments_b200 as pkg
    sizes = [2000, 1, 0, 777]
    for levelsup in (4, 1):
        voc, feats, ref = _bow_case(tmp_path, vkw, levelsup, sizes)
        gv = pkg.ORBVocabulary(*voc)
        desc = np.zeros((len(sizes), 2003, 32), np.uint8)
        for f, a in enumerate(feats):
            desc[f, :len(a)] = a
        r = gv.transform(desc, np.array(sizes, np.int32), levelsup)
        for f, n in enumerate(sizes):
            gbow, gfv = pkg.ORBVocabulary.frame_maps(r, f)
            _check_bow(gbow, gfv, r["node"][f, :n], ref[f], voc[1] - levelsup)


def test_avx2_timing_build_gives_the_same_output(tmp_path):
    """bench.py's reference arm times oracle/_ref/orb_ref_extract_avx2 (-O3 -march=x86-64-v3, standing in for the reference's
    -march=native): its output file must equal the plain build's byte for byte."""
    import subprocess
    fast = ref_binary.BIN + "_avx2"
    if not os.access(fast, os.X_OK):
        pytest.skip("oracle/_ref/orb_ref_extract_avx2 not built")
    frames = synth.synth_batch(640, 480, 2, seed0=77, distinct=2)
    fin = str(tmp_path / "in.bin")
    ref_binary.write_input(fin, frames, 1000)
    outs = []
    for exe in (ref_binary.BIN, fast):
        fout = str(tmp_path / (os.path.basename(exe) + ".out"))
        r = subprocess.run([exe, fin, fout], capture_output=True)
        if r.returncode != 0 and exe == fast:
            pytest.skip("this host cannot run the AVX2 build")
        assert r.returncode == 0
        outs.append(open(fout, "rb").read())
    assert outs[0] == outs[1]
