"""Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:770-806, 832-868). The oracle is PINNED to the real
cv2.undistortPoints (tests/golden/undistort_cv2.npz, made by tests/golden/make_golden_undistort.py); the CUDA path is compared with
the oracle bit for bit (float32 outputs of a double-precision iteration)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "undistort_cv2.npz")
NAMES = ["tum1", "euroc", "rational8", "strong"]


def _cam(g, name):
    c = g[name + "_cam"]
    return int(c[0]), int(c[1]), float(c[2]), float(c[3]), float(c[4]), float(c[5]), [float(x) for x in c[6:]]


@pytest.mark.parametrize("name", NAMES)
def test_oracle_equals_cv2_golden(oracle, name):
    g = np.load(GOLD)
    w, h, fx, fy, cx, cy, k = _cam(g, name)
    got = oracle.undistort_points(g[name + "_pts"], fx, fy, cx, cy, k)
    assert np.array_equal(got.view(np.uint32), g[name + "_und"].view(np.uint32))


def test_oracle_no_distortion_is_identity(oracle):
    pts = np.random.default_rng(0).uniform(0, 600, (100, 2)).astype(np.float32)
    assert np.array_equal(oracle.undistort_points(pts, 500., 500., 320., 240., [0.0, 0.3, 0.0, 0.0]), pts)   # only k1 == 0 is tested (:772)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_gpu_equals_golden_and_oracle(oracle, name):
    from orb_slam2_detailed_comments_b200 import frame, KP_DTYPE
    g = np.load(GOLD)
    w, h, fx, fy, cx, cy, k = _cam(g, name)
    rng = np.random.default_rng(3)
    F, cap = 3, 70000
    counts = np.array([6000, 70000, 0], np.int32)
    kps = np.zeros((F, cap), KP_DTYPE)
    kps["x"], kps["y"] = rng.uniform(-30, w + 30, (F, cap)), rng.uniform(-30, h + 30, (F, cap))
    kps["x"][0, :6000], kps["y"][0, :6000] = g[name + "_pts"][:, 0], g[name + "_pts"][:, 1]
    kps["angle"], kps["octave"], kps["response"] = rng.uniform(0, 360, (F, cap)), rng.integers(0, 8, (F, cap)), 40
    un = frame.UndistortKeyPoints(kps, counts, fx, fy, cx, cy, k)
    got0 = np.stack([un["x"][0, :6000], un["y"][0, :6000]], 1)
    assert np.array_equal(got0.view(np.uint32), g[name + "_und"].view(np.uint32))           # == real OpenCV
    want1 = oracle.undistort_points(np.stack([kps["x"][1], kps["y"][1]], 1), fx, fy, cx, cy, k)
    assert np.array_equal(np.stack([un["x"][1], un["y"][1]], 1).view(np.uint32), want1.view(np.uint32))
    for fld in ("size", "angle", "response", "octave", "class_id"):                        # "cv::KeyPoint kp = mvKeys[i]" keeps the rest
        assert np.array_equal(un[fld][:2], kps[fld][:2])
    assert np.array_equal(un[0, 6000:].view(np.uint8), kps[0, 6000:].view(np.uint8))        # rows >= n untouched
    # ComputeImageBounds = min / max over the four undistorted corners (the first four golden points)
    c = g[name + "_und"][:4]
    want_b = (min(c[0, 0], c[2, 0]), min(c[0, 1], c[1, 1]), max(c[1, 0], c[3, 0]), max(c[2, 1], c[3, 1]))
    assert frame.ComputeImageBounds(w, h, fx, fy, cx, cy, k) == tuple(float(x) for x in want_b)


@pytest.mark.gpu
def test_gpu_no_distortion():
    from orb_slam2_detailed_comments_b200 import frame, KP_DTYPE
    kps = np.zeros((1, 50), KP_DTYPE)
    kps["x"], kps["y"] = np.arange(50), np.arange(50) * 2
    un = frame.UndistortKeyPoints(kps, np.array([50], np.int32), 500., 500., 320., 240., [0.0, 0.0, 0.0, 0.0])
    assert np.array_equal(un.view(np.uint8), kps.view(np.uint8))
    assert frame.ComputeImageBounds(1241, 376, 718.856, 718.856, 607.19, 185.2, [0.0, 0.0, 0.0, 0.0]) == (0.0, 0.0, 1241.0, 376.0)
