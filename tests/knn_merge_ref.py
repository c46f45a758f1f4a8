"""Host-side (numpy) statement of the shard merge rule, test infrastructure only: the product merges on the device (orb_hamming_knn_merge /\norb_hamming_knn_sharded). Moved out of the product package in round 2."""
import numpy as np

from orb_slam2_detailed_comments_b200._lib import KNN_DTYPE


def merge_records_host(parts: list, sentinel: int = 256) -> np.ndarray:
    """Host-side statement of the shard merge rule (the two lexicographically smallest (d, i) with d < sentinel)
    for the gloo tests of the exchange logic; the product path merges on the device (orb_hamming_knn_merge)."""
    Q = len(parts[0])
    d = np.stack([np.stack([p["d1"], p["d2"]], 1) for p in parts], 1).reshape(Q, -1).astype(np.int64)
    i = np.stack([np.stack([p["i1"], p["i2"]], 1) for p in parts], 1).reshape(Q, -1).astype(np.int64)
    key = np.where((i < 0) | (d >= sentinel), np.iinfo(np.int64).max, (d << 40) | np.maximum(i, 0))
    key.sort(axis=1)
    out = np.zeros(Q, KNN_DTYPE)
    big = np.iinfo(np.int64).max
    a, b = key[:, 0], key[:, 1]
    out["d1"] = np.where(a == big, sentinel, a >> 40)
    out["i1"] = np.where(a == big, -1, a & ((1 << 40) - 1))
    ok2 = (a != big) & (b != big)
    out["d2"] = np.where(ok2, b >> 40, sentinel)
    out["i2"] = np.where(ok2, b & ((1 << 40) - 1), -1)
    return out
