"""N1: Frame::ComputeStereoMatches on the device (orb_extract_stereo_batch) versus the oracle's restatement of
src/Frame.cc:880-1176 run on the oracle's own keypoints / descriptors / pyramids. mvuRight and mvDepth bit-exact (float32)."""
import numpy as np
import pytest

from orb_slam2_detailed_comments_b200 import synth

pytestmark = pytest.mark.gpu


def _pair(w, h, seed, shift):
    return synth.synth_stereo_pair(w, h, seed, shift)


def _oracle_pair(oracle, nf, left, right, bf, b):
    oL, oR = oracle.OracleExtractor(nf), oracle.OracleExtractor(nf)
    kl, dl, _ = oL(left)
    kr, dr, _ = oR(right)
    ur, dp, kept = oracle.stereo_match(oL, oR, kl, dl, kr, dr, bf, b)
    return kl, dl, kr, dr, ur, dp, kept


@pytest.mark.parametrize("w,h,nf,bf,b", [(1241, 376, 2000, 386.1448, 0.5372), (752, 480, 1200, 47.9, 0.11), (640, 480, 1000, 40.0, 0.08)])
def test_stereo_matches_oracle(oracle, w, h, nf, bf, b):
    from orb_slam2_detailed_comments_b200 import ORBextractor
    P = 5
    pairs = [_pair(w, h, 700 + i, 4 + 5 * i) for i in range(P)]
    left = np.stack([p[0] for p in pairs])
    right = np.stack([p[1] for p in pairs])
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    kps, desc, counts, ur, dp, matched = ex.extract_stereo_batch(left, right, bf, b)
    total = 0
    for i in range(P):
        kl, dl, kr, dr, our, odp, kept = _oracle_pair(oracle, nf, left[i], right[i], bf, b)
        nl, nr = int(counts[2 * i]), int(counts[2 * i + 1])
        assert nl == len(kl) and nr == len(kr)
        assert np.array_equal(kps[2 * i, :nl].view(np.uint8), kl.view(np.uint8)) and np.array_equal(desc[2 * i + 1, :nr], dr)
        assert np.array_equal(ur[i, :nl].view(np.uint32), our.view(np.uint32)), f"pair {i}: mvuRight differs"
        assert np.array_equal(dp[i, :nl].view(np.uint32), odp.view(np.uint32)), f"pair {i}: mvDepth differs"
        assert int(matched[i]) == kept == int((our >= 0).sum())
        total += kept
    assert total > 100 * P, "too few stereo matches for a meaningful test"


def test_stereo_multipass_and_degenerate(oracle, monkeypatch):
    """Pairs never straddle passes (pass size forced to 6 frames = 3 pairs, ramp-up passes of 2); a textureless right image
    yields no match and skips the median filter."""
    from orb_slam2_detailed_comments_b200 import ORBextractor
    monkeypatch.setenv("ORB_B200_CHUNK", "6")
    P = 11
    pairs = [_pair(480, 360, 900 + i, 3 + i) for i in range(P)]
    left = np.stack([p[0] for p in pairs])
    right = np.stack([p[1] for p in pairs])
    right[4] = 128
    ex = ORBextractor(600, 1.2, 8, 20, 7)
    kps, desc, counts, ur, dp, matched = ex.extract_stereo_batch(left, right, 60.0, 0.1)
    assert matched[4] == 0 and counts[9] == 0 and np.all(ur[4, :counts[8]] == -1)
    for i in (0, 3, 5, 10):
        kl, dl, kr, dr, our, odp, kept = _oracle_pair(oracle, 600, left[i], right[i], 60.0, 0.1)
        nl = int(counts[2 * i])
        assert np.array_equal(ur[i, :nl].view(np.uint32), our.view(np.uint32))
        assert np.array_equal(dp[i, :nl].view(np.uint32), odp.view(np.uint32))
        assert int(matched[i]) == kept
