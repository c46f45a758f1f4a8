"""Loader for liborb_b200.so (the C ABI declared in include/orb_b200.h).

There is no CPU fallback: if the shared library is missing or fails to load, import of the compute
classes raises. The library is built in-tree by `__graft_entry__.build()` / csrc/Makefile.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "liborb_b200.so")
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
KNN_DTYPE = np.dtype([("d1", "<i4"), ("d2", "<i4"), ("i1", "<i8"), ("i2", "<i8")])
# orb_proj_query / orb_search_params (guided matching)
QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("radius", "<f4"), ("ur", "<f4"), ("angle", "<f4"),
                        ("min_level", "<i2"), ("max_level", "<i2"), ("flags", "<i4"), ("reserved", "<i4")])
QUERY_SKIP, QUERY_CLAIMS = 1, 2
assert KP_DTYPE.itemsize == 28 and KNN_DTYPE.itemsize == 24 and QUERY_DTYPE.itemsize == 32

ORB_OK, ORB_E_BADARG, ORB_E_CAPACITY, ORB_E_CUDA, ORB_E_SHAPE = 0, -1, -2, -3, -4
MEM_HOST, MEM_DEVICE = 0, 1
FLAG_UPSTREAM_PYRAMID, FLAG_LEGACY_BLUR = 1, 2

# every symbol include/orb_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "orb_last_error", "orb_abi_version", "orb_device_count", "orb_extractor_create", "orb_extractor_destroy",
    "orb_extractor_get_levels", "orb_extractor_get_scale_factor", "orb_extractor_get_table",
    "orb_extractor_get_features_per_level", "orb_extractor_max_keypoints", "orb_extract_batch",
    "orb_extract_batch_device", "orb_extract_stereo_batch", "orb_extractor_pyramid_dims", "orb_extractor_pyramid_read",
    "orb_extractor_raw_keypoints", "orb_extractor_stream", "orb_extractor_launch_count",
    "orb_extractor_check", "orb_extractor_pyramid_read_all", "orb_extractor_set_pyramid_download", "orb_host_alloc", "orb_host_free",
    "orb_extractor_set_profiling", "orb_extractor_stage_times", "orb_hamming_knn", "orb_hamming_knn_noalloc",
    "orb_hamming_knn_merge", "orb_hamming_knn_sharded", "orb_knn_peer_create", "orb_knn_peer_export", "orb_knn_peer_connect", "orb_hamming_knn_peer", "orb_knn_peer_check", "orb_knn_peer_destroy", "orb_nccl_unique_id", "orb_nccl_comm_create", "orb_nccl_comm_destroy", "orb_descriptor_distance_batch", "orb_descriptor_distance_host", "orb_matcher_launch_count", "orb_measure_popc_peak",
    "orb_search_by_projection", "orb_distinctive_descriptors",
    "orb_vocabulary_create", "orb_vocabulary_destroy", "orb_vocabulary_words", "orb_bow_transform",
    "orb_undistort_keypoints", "orb_image_bounds", "orb_search_by_bow", "orb_search_for_initialization",
    "orb_search_for_triangulation", "orb_fuse_search",
]


EPI_DTYPE = np.dtype([("f12", "<f4", (9,)), ("ex", "<f4"), ("ey", "<f4"), ("level_sigma2", "<f4", (16,)), ("scale_factor", "<f4", (16,)),
                      ("reserved", "<f4")])   # == orb_epipolar (176 bytes)


class OrbParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("flags", C.c_int32)]


class OrbCamera(C.Structure):
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("dist", C.c_float * 8), ("n_dist", C.c_int32)]

    @classmethod
    def make(cls, fx, fy, cx, cy, dist):
        d = list(dist) + [0.0] * (8 - len(dist))
        return cls(fx, fy, cx, cy, (C.c_float * 8)(*d), len(dist))


class OrbSearchParams(C.Structure):
    _fields_ = [("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("th_dist", C.c_int32), ("nn_ratio", C.c_float), ("check_orientation", C.c_int32), ("reserved", C.c_int32)]


class OrbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"orb_b200 error {code}: {msg}")
        self.code = code


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into liborb_b200.so (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", os.path.join(_PKG, "csrc"), "-j8"],
                          stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    L.orb_last_error.restype = C.c_char_p
    L.orb_extractor_create.argtypes = [C.POINTER(OrbParams), i32, C.POINTER(vp)]
    L.orb_extractor_destroy.argtypes = [vp]
    L.orb_extractor_destroy.restype = None
    L.orb_extractor_get_levels.argtypes = [vp]
    L.orb_extractor_get_scale_factor.argtypes = [vp]
    L.orb_extractor_get_scale_factor.restype = f32
    L.orb_extractor_get_table.argtypes = [vp, i32, vp]
    L.orb_extractor_get_features_per_level.argtypes = [vp, vp]
    L.orb_extractor_max_keypoints.argtypes = [vp]
    L.orb_extract_batch.argtypes = [vp, vp, i32, i32, i32, i32, i32, vp, vp, i32, i32, vp, vp]
    L.orb_extract_stereo_batch.argtypes = [vp, vp, i32, i32, i32, i32, i32, vp, vp, i32, i32, vp, vp, f32, f32, vp, vp, vp]
    L.orb_descriptor_distance_host.argtypes = [vp, vp, i64, vp]
    L.orb_extract_batch_device.argtypes = [vp, vp, i32, i32, i32, i32, vp, vp, i32, vp, i32]
    L.orb_extractor_pyramid_dims.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32)]
    L.orb_extractor_pyramid_read.argtypes = [vp, i32, i32, i32, vp, i32]
    L.orb_extractor_raw_keypoints.argtypes = [vp, i32, i32, vp, i32]
    L.orb_extractor_stream.argtypes = [vp]
    L.orb_extractor_stream.restype = vp
    L.orb_extractor_launch_count.argtypes = [vp]
    L.orb_extractor_launch_count.restype = i64
    L.orb_extractor_set_profiling.argtypes = [vp, i32]
    L.orb_extractor_stage_times.argtypes = [vp, vp, vp]
    L.orb_hamming_knn.argtypes = [vp, i32, i32, vp, i32, i64, i64, i32, vp, i32, vp]
    L.orb_hamming_knn_merge.argtypes = [vp, i32, i32, i32, vp, vp]
    L.orb_hamming_knn_noalloc.argtypes = [vp, i32, vp, i64, i64, i32, vp, vp]
    L.orb_hamming_knn_sharded.argtypes = [vp, i32, i32, vp, i32, i64, i64, i32, vp, i32, vp, vp]
    L.orb_knn_peer_create.argtypes = [i32, i32, i32, C.POINTER(vp)]
    L.orb_knn_peer_export.argtypes = [vp, vp]
    L.orb_knn_peer_connect.argtypes = [vp, vp]
    L.orb_hamming_knn_peer.argtypes = [vp, vp, i32, i32, vp, i32, i64, i64, i32, vp, i32, vp]
    L.orb_knn_peer_check.argtypes = [vp, vp]
    L.orb_knn_peer_destroy.argtypes = [vp]
    L.orb_knn_peer_destroy.restype = None
    L.orb_nccl_unique_id.argtypes = [vp]
    L.orb_nccl_comm_create.argtypes = [vp, i32, i32, C.POINTER(vp)]
    L.orb_nccl_comm_destroy.argtypes = [vp]
    L.orb_nccl_comm_destroy.restype = None
    L.orb_descriptor_distance_batch.argtypes = [vp, vp, i64, vp, vp]
    L.orb_matcher_launch_count.restype = i64
    L.orb_search_for_initialization.argtypes = [i32, vp, vp, vp, i32, vp, vp, vp, i32, i32, C.POINTER(OrbSearchParams), vp, vp, vp, vp]
    L.orb_search_by_bow.argtypes = [i32] + [vp] * 7 + [i32] + [vp] * 7 + [i32, i32, i32, f32, i32, vp, vp, vp, vp]
    L.orb_search_for_triangulation.argtypes = [i32] + [vp] * 8 + [i32] + [vp] * 8 + [i32, i32, vp, i32, i32, vp, vp, vp]
    L.orb_fuse_search.argtypes = [i32, vp, vp, vp, vp, i32, vp, vp, vp, i32, i32, C.POINTER(OrbSearchParams), vp, vp, vp, vp]
    L.orb_undistort_keypoints.argtypes = [i32, C.POINTER(OrbCamera), vp, vp, i32, i32, vp, vp]
    L.orb_image_bounds.argtypes = [C.POINTER(OrbCamera), i32, i32, vp]
    L.orb_distinctive_descriptors.argtypes = [i32, vp, vp, i32, vp, vp, vp]
    L.orb_search_by_projection.argtypes = [i32, vp, vp, vp, vp, vp, i32, vp, vp, vp, i32, i32, C.POINTER(OrbSearchParams), vp, vp, vp, vp]
    L.orb_measure_popc_peak.argtypes = [i32]
    L.orb_measure_popc_peak.restype = C.c_double
    _LIB = L
    return L


def check(rc: int) -> int:
    if rc < 0:
        raise OrbError(rc, lib().orb_last_error().decode("utf-8", "replace"))
    return rc
