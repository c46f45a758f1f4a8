"""The C-ABI library loads and exports every symbol include/orb_b200.h declares; without a GPU the compute entry
points fail loudly (ORB_E_CUDA) instead of falling back to a CPU path. No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "orb_b200.h")


def _declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orb_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    from orb_slam2_detailed_comments_b200 import _lib
    L = _lib.lib()
    syms = _declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/orb_b200.h but not exported"
    assert sorted(_lib.SYMBOLS) == syms, "python SYMBOLS list out of sync with the header"
    assert L.orb_abi_version() == 1


def test_struct_layouts():
    from orb_slam2_detailed_comments_b200 import _lib
    assert _lib.KP_DTYPE.itemsize == 28     # cv::KeyPoint
    assert _lib.KNN_DTYPE.itemsize == 24
    assert C.sizeof(_lib.OrbParams) == 24


def test_product_never_imports_oracle():
    """The package, the C++ drop-in layer, the public headers and the tools never import, include or link the oracle
    (only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may)."""
    for top in ("orb_slam2_detailed_comments_b200", "cpp", "include", "tools"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".cc", "Makefile")):
                    txt = open(os.path.join(dirpath, f), errors="replace").read()
                    for needle in ("oracle_py", "liborb_oracle", "orb_oracle.h", "lorb_oracle", "from oracle", "import oracle"):
                        assert needle not in txt, (os.path.join(top, f), needle)


def test_no_cpu_fallback_without_device():
    from orb_slam2_detailed_comments_b200 import _lib
    L = _lib.lib()
    if L.orb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    p = _lib.OrbParams(1000, 1.2, 8, 20, 7, 0)
    rc = L.orb_extractor_create(C.byref(p), 0, C.byref(h))
    assert rc == _lib.ORB_E_CUDA and not h.value
    assert b"no CUDA device" in L.orb_last_error()
    q = np.zeros((4, 32), np.uint8)
    out = np.zeros(4, _lib.KNN_DTYPE)
    rc = L.orb_hamming_knn(q.ctypes.data_as(C.c_void_p), 0, 4, q.ctypes.data_as(C.c_void_p), 0, 4, 0, 256,
                           out.ctypes.data_as(C.c_void_p), 0, None)
    assert rc == _lib.ORB_E_CUDA
    from orb_slam2_detailed_comments_b200 import ORBextractor, OrbError
    with pytest.raises(OrbError):
        ORBextractor(1000, 1.2, 8, 20, 7)


def test_bad_arguments_rejected_before_cuda():
    from orb_slam2_detailed_comments_b200 import _lib
    L = _lib.lib()
    h = C.c_void_p()
    assert L.orb_extractor_create(None, 0, C.byref(h)) == _lib.ORB_E_BADARG
    for bad in [(0, 1.2, 8, 20, 7, 0), (100, 1.0, 8, 20, 7, 0), (100, 1.2, 0, 20, 7, 0), (100, 1.2, 17, 20, 7, 0),
                (100, 1.2, 8, 5, 7, 0), (100, 1.2, 8, 20, 0, 0), (100, 1.2, 8, 20, 7, 4), (100, 1.2, 8, 20, 7, -1)]:
        p = _lib.OrbParams(*bad)
        assert L.orb_extractor_create(C.byref(p), 0, C.byref(h)) == _lib.ORB_E_BADARG, bad
    assert L.orb_hamming_knn(None, 0, 1, None, 0, 0, 0, 256, None, 0, None) == _lib.ORB_E_BADARG
    assert L.orb_extractor_get_levels(None) == 0
    assert L.orb_extractor_max_keypoints(None) == 0


def test_peer_knn_bad_arguments_rejected_before_cuda():
    """The peer-store kNN (include/orb_b200.h orb_knn_peer_*) validates ranks, capacities and pointers before it touches the device; on a box
    without a GPU a valid create reports ORB_E_CUDA (no CPU fallback)."""
    from orb_slam2_detailed_comments_b200 import _lib
    L = _lib.lib()
    h = C.c_void_p()
    for rank, nranks, max_q in [(0, 0, 16), (0, 17, 16), (-1, 2, 16), (2, 2, 16), (0, 2, 0)]:
        assert L.orb_knn_peer_create(rank, nranks, max_q, C.byref(h)) == _lib.ORB_E_BADARG, (rank, nranks, max_q)
        assert not h.value
    assert L.orb_knn_peer_create(0, 2, 16, None) == _lib.ORB_E_BADARG
    buf = (C.c_uint8 * 64)()
    assert L.orb_knn_peer_export(None, buf) == _lib.ORB_E_BADARG
    assert L.orb_knn_peer_connect(None, buf) == _lib.ORB_E_BADARG
    assert L.orb_knn_peer_check(None, None) == _lib.ORB_E_BADARG
    assert L.orb_hamming_knn_peer(None, None, 0, 1, None, 0, 0, 0, 256, None, 0, None) == _lib.ORB_E_BADARG
    L.orb_knn_peer_destroy(None)                       # a null handle is ignored
    if L.orb_device_count() == 0:
        assert L.orb_knn_peer_create(0, 2, 16, C.byref(h)) == _lib.ORB_E_CUDA and not h.value
