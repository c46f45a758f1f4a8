"""GPU parity of the extraction path: CUDA (through the C ABI) vs the CPU oracle, stage by stage.
Bit-exact bar for pyramid bytes, raw FAST keypoints (order included), kept keypoints (order, coordinates,
octave, response, size), blurred bytes, descriptors; angles bit-exact (float replica of cv::fastAtan2)."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

pytestmark = pytest.mark.gpu

CASES = [
    ("tum", 640, 480, 1000, 1234),
    ("tum", 640, 480, 2000, 1235),    # mono-init extractor (2 x nFeatures, src/Tracking.cc:180)
    ("kitti", 1241, 376, 2000, 1236),
    ("euroc", 752, 480, 1200, 1237),
    ("small", 200, 160, 500, 1238),
    ("odd", 333, 257, 700, 1239),
]


def _extractor(nf, **kw):
    from orb_slam2_detailed_comments_b200 import ORBextractor
    return ORBextractor(nf, 1.2, 8, 20, 7, **kw)


def _compare(img, nf, oracle, ini=20, mn=7, nlevels=8, sf=1.2, flags=0):
    from orb_slam2_detailed_comments_b200 import ORBextractor
    ex = ORBextractor(nf, sf, nlevels, ini, mn, flags=flags)
    oflags = (oracle.LEGACY_BLUR if flags & 2 else 0) | (oracle.UPSTREAM_PYRAMID if flags & 1 else 0)
    core = (lambda a: a[19:-19, 19:-19]) if flags & 1 else (lambda a: a)   # the oracle keeps the padded buffer, mvImagePyramid[l] is its interior
    oe = oracle.OracleExtractor(nf, sf, nlevels, ini, mn, flags=oflags)
    okps, odesc, olc = oe(img)
    kps, desc, counts, lcounts = ex.extract_batch(img[None])
    n = int(counts[0])
    for l in range(nlevels):
        assert np.array_equal(ex.pyramid_level(l), core(oe.level(l))), f"pyramid level {l}"
    for l in range(nlevels):
        r_o, r_g = oe.raw_keypoints(l), ex.raw_keypoints(l)
        assert len(r_o) == len(r_g), f"raw count level {l}: {len(r_g)} vs {len(r_o)}"
        for f in ("x", "y", "response"):
            assert np.array_equal(r_o[f], r_g[f]), f"raw {f} level {l}"
    assert np.array_equal(lcounts[0], olc), f"level counts {lcounts[0]} vs {olc}"
    assert n == len(okps)
    for l in range(nlevels):
        ob = oe.level(l, blurred=True)
        if ob is not None:
            assert np.array_equal(ex.pyramid_level(l, blurred=True), ob), f"blurred level {l}"
    k = kps[0, :n]
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(k[f], okps[f]), f"keypoint field {f}"
    assert np.array_equal(k["angle"].view(np.uint32), okps["angle"].view(np.uint32)), \
        f"angle max diff {np.abs(k['angle'] - okps['angle']).max()}"
    assert np.array_equal(desc[0, :n], odesc)
    return n


@pytest.mark.parametrize("name,w,h,nf,seed", CASES)
def test_extract_matches_oracle(oracle, name, w, h, nf, seed):
    img = synth.synth_frame(w, h, seed)
    n = _compare(img, nf, oracle)
    assert n > 0


def test_thresholds_and_levels(oracle):
    img = synth.synth_frame(480, 360, 77)
    _compare(img, 800, oracle, ini=12, mn=7)            # KITTI04-12 stereo uses ini 12
    _compare(img, 300, oracle, nlevels=4, sf=1.5)
    _compare(img, 600, oracle, flags=2)                 # legacy blur taps


@pytest.mark.parametrize("name,w,h,nf,seed", CASES)
def test_upstream_pyramid_switch(oracle, name, w, h, nf, seed):
    """ORB_FLAG_UPSTREAM_PYRAMID: raulmur's un-padded pyramid - the same kernels with a stored border of 0 instead of 19."""
    img = synth.synth_frame(w, h, seed)
    if name == "small":
        # un-padded, level 7 of a 200 x 160 image is 56 x 45: narrower than one 30-px cell plus its borders, where the reference divides by
        # nCols == 0 (src/ORBextractor.cc:1459-1462). The library refuses the shape; five levels fit.
        from orb_slam2_detailed_comments_b200 import ORBextractor
        from orb_slam2_detailed_comments_b200._lib import OrbError
        with pytest.raises(OrbError):
            ORBextractor(nf, 1.2, 8, 20, 7, flags=1).extract_batch(img[None])
        assert _compare(img, nf, oracle, nlevels=5, flags=1) > 0
        return
    n = _compare(img, nf, oracle, flags=1)
    assert n > 0
    if name == "tum":
        _compare(img, nf, oracle, flags=3)             # together with the legacy blur taps
        _compare(img, 300, oracle, nlevels=4, sf=1.5, flags=1)


def test_upstream_pyramid_stereo(oracle):
    """ComputeStereoMatches over the un-padded pyramid: the SAD windows read mvImagePyramid[l] = the interior."""
    from orb_slam2_detailed_comments_b200 import ORBextractor
    left, right = synth.synth_stereo_pair(752, 480, 41, 9)
    ex = ORBextractor(1200, 1.2, 8, 20, 7, flags=1)
    kps, desc, counts, ur, dp, matched = ex.extract_stereo_batch(left[None], right[None], 47.9, 0.11)
    oL = oracle.OracleExtractor(1200, flags=oracle.UPSTREAM_PYRAMID)
    oR = oracle.OracleExtractor(1200, flags=oracle.UPSTREAM_PYRAMID)
    kl, dl, _ = oL(left)
    kr, dr, _ = oR(right)
    our, odp, kept = oracle.stereo_match(oL, oR, kl, dl, kr, dr, 47.9, 0.11)
    nl = int(counts[0])
    assert nl == len(kl) and int(counts[1]) == len(kr)
    assert np.array_equal(ur[0, :nl].view(np.uint32), our.view(np.uint32)), "mvuRight differs"
    assert np.array_equal(dp[0, :nl].view(np.uint32), odp.view(np.uint32)), "mvDepth differs"
    assert int(matched[0]) == kept and kept > 100


def test_flat_and_noise_images(oracle):
    flat = np.full((240, 320), 128, np.uint8)
    from orb_slam2_detailed_comments_b200 import ORBextractor
    ex = ORBextractor(500, 1.2, 8, 20, 7)
    k, d = ex(flat)
    assert len(k) == 0 and d.shape == (0, 32)
    rng = np.random.RandomState(3)
    noise = rng.randint(0, 256, (240, 320), dtype=np.uint8)
    _compare(noise, 500, oracle)


def test_batch_equals_single(oracle):
    frames = synth.synth_batch(640, 480, 5, seed0=500)
    ex = _extractor(1000)
    kps, desc, counts, lc = ex.extract_batch(frames)
    oe = oracle.OracleExtractor(1000)
    for i in range(len(frames)):
        ok, od, olc = oe(frames[i])
        n = int(counts[i])
        assert n == len(ok) and np.array_equal(lc[i], olc)
        assert np.array_equal(kps[i, :n].view(np.uint8), ok.view(np.uint8))
        assert np.array_equal(desc[i, :n], od)


def test_strided_input_and_call_operator(oracle):
    big = synth.synth_frame(700, 500, 9)
    view = big[10:490, 20:660]      # non-contiguous rows (treated as isolated)
    ex = _extractor(1000)
    k, d = ex(view)
    ok, od, _ = oracle.OracleExtractor(1000)(np.ascontiguousarray(view))
    assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)
    k2, d2 = ex(np.zeros((0, 0), np.uint8))
    assert len(k2) == 0


def test_device_resident_path(oracle):
    import torch
    frames = synth.synth_batch(752, 480, 6, seed0=900)
    ex = _extractor(1200)
    t = torch.from_numpy(frames).cuda()
    kps, desc, counts = ex.extract_batch_device(t)
    kps, desc, counts = kps.cpu().numpy(), desc.cpu().numpy(), counts.cpu().numpy()
    oe = oracle.OracleExtractor(1200)
    for i in range(len(frames)):
        ok, od, _ = oe(frames[i])
        n = int(counts[i])
        assert n == len(ok)
        assert np.array_equal(kps[i, :n].view(np.uint8).reshape(-1), ok.view(np.uint8).reshape(-1))
        assert np.array_equal(desc[i, :n], od)


def test_errors():
    from orb_slam2_detailed_comments_b200 import ORBextractor, OrbError
    with pytest.raises(OrbError):
        ORBextractor(0, 1.2, 8, 20, 7)
    ex = ORBextractor(100, 1.2, 8, 20, 7)
    with pytest.raises(OrbError):       # level 7 too small for the 19-px border
        ex(np.zeros((64, 64), np.uint8))
    with pytest.raises(OrbError):       # portrait: nIni == 0 in the reference
        ex(np.zeros((900, 300), np.uint8))
    with pytest.raises(TypeError):
        ex.extract_batch(np.zeros((1, 100, 100), np.float32))


def test_caller_device_preserved():
    """A handle on another GPU must not move the calling thread's current CUDA device (extractor and vocabulary entry points)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import orb_slam2_detailed_comments_b200 as pkg
    torch.cuda.set_device(0)
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, device=1)
    img = synth.synth_frame(320, 240, 5)
    k1, d1 = ex(img)
    assert torch.cuda.current_device() == 0
    t = torch.zeros(4, device="cuda")
    assert t.device.index == 0
    k0, d0 = pkg.ORBextractor(500, 1.2, 8, 20, 7, device=0)(img)
    assert np.array_equal(k0.view(np.uint8), k1.view(np.uint8)) and np.array_equal(d0, d1)
    voc = synth.synth_vocabulary(10, 3, seed=11)
    gv = pkg.ORBVocabulary(*voc, device=1)
    assert torch.cuda.current_device() == 0
    del gv, ex
    assert torch.cuda.current_device() == 0
