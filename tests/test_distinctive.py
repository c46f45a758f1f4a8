"""MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:424-517): oracle vs a numpy statement of the same rule
(CPU), CUDA kernel vs oracle (GPU). No upstream fixture: parity unpinned, cross-checked."""
import numpy as np
import pytest


def _points(seed, sizes):
    rng = np.random.default_rng(seed)
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    desc = np.zeros((int(offsets[-1]), 32), np.uint8)
    for p, n in enumerate(sizes):
        if n == 0:
            continue
        centre = rng.integers(0, 256, 32, dtype=np.uint8)
        rows = np.repeat(centre[None], n, 0)
        flips = rng.random((n, 256)) < rng.uniform(0.02, 0.25)   # observations of one point: noisy copies of one descriptor
        rows ^= np.packbits(flips, axis=1)
        if n > 3 and p % 3 == 0:
            rows[n // 2] = rows[0]   # exact duplicates -> equal medians, the first row must win
        desc[offsets[p]:offsets[p + 1]] = rows
    return desc, offsets


def _numpy_rule(desc, offsets):
    bi, bm = [], []
    for p in range(len(offsets) - 1):
        d = desc[offsets[p]:offsets[p + 1]]
        n = len(d)
        if n == 0:
            bi.append(-1); bm.append(-1); continue
        bits = np.unpackbits(d, axis=1).astype(np.int32)
        D = (bits[:, None, :] != bits[None, :, :]).sum(-1)
        med = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
        bi.append(int(np.argmin(med))); bm.append(int(med.min()))   # argmin = first minimum = strict <
    return np.array(bi, np.int32), np.array(bm, np.int32)


SIZES = [1, 2, 3, 4, 5, 0, 7, 8, 16, 31, 32, 33, 64, 65, 100, 6, 6, 6, 2, 1, 0, 40]


def test_oracle_equals_numpy_rule(oracle):
    desc, offsets = _points(1, SIZES)
    bi, bm = oracle.distinctive_descriptors(desc, offsets, nthreads=3)
    wi, wm = _numpy_rule(desc, offsets)
    assert np.array_equal(bi, wi) and np.array_equal(bm, wm)


@pytest.mark.gpu
def test_gpu_equals_oracle(oracle):
    import orb_slam2_detailed_comments_b200 as pkg
    rng = np.random.default_rng(5)
    sizes = SIZES + [300, 513] + rng.integers(0, 25, 3000).tolist()
    desc, offsets = _points(2, sizes)
    bi, bm = pkg.ORBmatcher.ComputeDistinctiveDescriptors(desc, offsets)
    wi, wm = oracle.distinctive_descriptors(desc, offsets, nthreads=8)
    assert np.array_equal(bi, wi) and np.array_equal(bm, wm)
    assert (bi[np.array(sizes) == 0] == -1).all()


@pytest.mark.gpu
def test_gpu_device_mode_and_empty(oracle):
    import ctypes as C
    import torch
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import lib, check, MEM_DEVICE
    desc, offsets = _points(3, [5, 9, 0, 12, 1])
    td, to = torch.from_numpy(desc).cuda(), torch.from_numpy(offsets).cuda()
    bi = torch.empty(5, dtype=torch.int32, device="cuda")
    check(lib().orb_distinctive_descriptors(MEM_DEVICE, C.c_void_p(td.data_ptr()), C.c_void_p(to.data_ptr()), 5, C.c_void_p(bi.data_ptr()), None,
                                            C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert np.array_equal(bi.cpu().numpy(), oracle.distinctive_descriptors(desc, offsets)[0])
    e = pkg.ORBmatcher.ComputeDistinctiveDescriptors(np.zeros((0, 32), np.uint8), np.zeros(1, np.int64))
    assert len(e[0]) == 0
    z = pkg.ORBmatcher.ComputeDistinctiveDescriptors(np.zeros((0, 32), np.uint8), np.zeros(4, np.int64))
    assert (z[0] == -1).all()
