"""CPU: the C++ oracle of the guided search against the independent Python restatement (tests/guided_ref.py).
No upstream fixture exists for this path (the reference's matcher does not build without OpenCV/DBoW2/g2o), so the
oracle is cross-checked, not pinned: DESIGN.md says "parity unpinned" for SURVEY 8(f) N2."""
import numpy as np
import pytest

from oracle import oracle_py as O
import guided_ref as R
from guided_cases import make_case


@pytest.mark.parametrize("bounds", [(0.0, 0.0, 752.0, 480.0), (-21.3, -14.9, 771.6, 493.2)])
def test_features_in_area_order_and_content(bounds):
    c = make_case(3, n=900, bounds=bounds)
    k = c["kps"]
    k["x"][:5] = [bounds[0], bounds[2], bounds[2] - 0.01, -500, 5000]   # on / outside the grid box: PosInGrid rejects posX == 64
    grid = R.Grid(k, bounds)
    rng = np.random.default_rng(1)
    for _ in range(200):
        x, y = rng.uniform(-50, 800), rng.uniform(-50, 530)
        r = float(rng.choice([0.5, 3, 9.7, 30, 120, 2000]))
        lo, hi = [(-1, -1), (0, 3), (2, -1), (3, 4), (1, 0)][rng.integers(0, 5)]
        got = O.features_in_area(k, bounds, x, y, r, lo, hi)
        assert got.tolist() == grid.in_area(x, y, r, lo, hi)


def test_three_maxima_rules():
    assert R.three_maxima([0] * 30) == (-1, -1, -1)
    assert R.three_maxima([5, 5, 5] + [0] * 27) == (0, 1, 2)            # strict >: the first of equal bins ranks higher
    assert R.three_maxima([100, 9, 8] + [0] * 27) == (0, -1, -1)        # max2 < 0.1 * max1 drops both
    assert R.three_maxima([100, 10, 9] + [0] * 27) == (0, 1, -1)        # 10 is not < 10.0
    assert R.three_maxima([1, 2, 3, 4] + [0] * 26) == (3, 2, 1)


@pytest.mark.parametrize("seed,kw", [
    (0, dict()),
    (1, dict(radius=40.0, dup=0.5)),                   # heavy competition for the same keypoints
    (2, dict(stereo=True)),
    (3, dict(claims=0.3)),                             # non-claiming queries: a keypoint can be re-assigned
    (4, dict(n=40, m=300, radius=200.0)),              # every window holds every keypoint
    (5, dict(bounds=(-21.3, -14.9, 771.6, 493.2))),
    (6, dict(rot=0.0)),
])
def test_oracle_matches_python_restatement(seed, kw):
    c = make_case(seed, **kw)
    for nn_ratio, ori, th in ((0.0, True, 100), (0.8, False, 100), (0.9, True, 50)):
        want = R.search_by_projection(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], th, nn_ratio, ori, c["uright"], c["claimed"])
        got = O.search_by_projection(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], th, nn_ratio, ori, c["uright"], c["claimed"])
        assert got[2] == want[2]
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        assert want[2] > 20   # the scenario exercises the accept path


def test_empty_sets():
    c = make_case(7, n=50, m=20)
    a, q, nm = O.search_by_projection(c["kps"][:0], c["desc"][:0], c["queries"], c["qdesc"], c["bounds"])
    assert nm == 0 and len(a) == 0 and (q == -1).all()
    a, q, nm = O.search_by_projection(c["kps"], c["desc"], c["queries"][:0], c["qdesc"][:0], c["bounds"])
    assert nm == 0 and (a == -1).all()
