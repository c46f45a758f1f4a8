"""Frame-level geometry of the reference's Frame constructors on the GPU (include/Frame.h): UndistortKeyPoints and
ComputeImageBounds (src/Frame.cc:770-806, 832-868) through the C ABI. No CPU path."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, MEM_HOST, OrbCamera, check


def UndistortKeyPoints(kps: np.ndarray, n: np.ndarray, fx, fy, cx, cy, dist) -> np.ndarray:
    """kps [F,cap] KP_DTYPE (mvKeys of F frames), n [F]; dist = mDistCoef (4, 5 or 8 floats). Returns mvKeysUn [F,cap]."""
    kps = np.ascontiguousarray(kps)
    assert kps.dtype == KP_DTYPE and kps.ndim == 2
    n = np.ascontiguousarray(n, np.int32)
    out = np.empty_like(kps)
    cam = OrbCamera.make(fx, fy, cx, cy, dist)
    check(_lib.lib().orb_undistort_keypoints(MEM_HOST, C.byref(cam), kps.ctypes.data_as(C.c_void_p), n.ctypes.data_as(C.c_void_p),
                                             kps.shape[1], kps.shape[0], out.ctypes.data_as(C.c_void_p), None))
    return out


def ComputeImageBounds(cols: int, rows: int, fx, fy, cx, cy, dist):
    """(mnMinX, mnMinY, mnMaxX, mnMaxY) of a cols x rows image."""
    cam = OrbCamera.make(fx, fy, cx, cy, dist)
    b = np.empty(4, np.float32)
    check(_lib.lib().orb_image_bounds(C.byref(cam), cols, rows, b.ctypes.data_as(C.c_void_p)))
    return tuple(float(x) for x in b)
