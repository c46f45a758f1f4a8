"""The monocular Frame constructor pipeline on device (extract -> undistort -> bag of words, src/Frame.cc:338-420) against the same
composition of the oracle pieces."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

pytestmark = pytest.mark.gpu
TUM1 = (517.306408, 516.469215, 318.643040, 255.313989, [0.262383, -0.953104, -0.005358, 0.002628, 1.163314])


def test_frame_frontend_matches_oracle_composition(oracle):
    import orb_slam2_detailed_comments_b200 as pkg
    voc = synth.synth_vocabulary(10, 4, seed=21)
    fe = pkg.FrameFrontEnd(pkg.ORBextractor(1000, 1.2, 8, 20, 7), *TUM1, vocabulary=pkg.ORBVocabulary(*voc))
    frames = synth.synth_batch(640, 480, 5, seed0=300, distinct=5)
    r = fe.process(frames)
    ov = oracle.OracleVocabulary(*voc)
    oe = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    for f in range(5):
        okps, odesc, _ = oe(frames[f])
        n = len(okps)
        assert r["counts"][f] == n and n > 500
        assert np.array_equal(r["keys"][f, :n].view(np.uint8), okps.view(np.uint8)) and np.array_equal(r["desc"][f, :n], odesc)
        und = oracle.undistort_points(np.stack([okps["x"], okps["y"]], 1), *TUM1)
        got = np.stack([r["keys_un"]["x"][f, :n], r["keys_un"]["y"][f, :n]], 1)
        assert np.array_equal(got.view(np.uint32), und.view(np.uint32))
        assert np.array_equal(r["keys_un"]["octave"][f, :n], okps["octave"]) and np.array_equal(r["keys_un"]["angle"][f, :n], okps["angle"])
        _, _, bow, fv = ov.transform(odesc, 4)
        gbow, gfv = pkg.ORBVocabulary.frame_maps(r, f)
        assert gbow == bow and gfv == fv


def test_frame_frontend_without_vocabulary_or_distortion():
    import orb_slam2_detailed_comments_b200 as pkg
    fe = pkg.FrameFrontEnd(pkg.ORBextractor(500, 1.2, 8, 20, 7), 718.856, 718.856, 607.1928, 185.2157, [0.0, 0.0, 0.0, 0.0])
    r = fe.process(synth.synth_batch(1241, 376, 2, seed0=5, distinct=2))
    assert "bow_word" not in r
    n = int(r["counts"][0])
    assert np.array_equal(r["keys"][0, :n].view(np.uint8), r["keys_un"][0, :n].view(np.uint8))   # mvKeysUn = mvKeys (src/Frame.cc:772-776)
