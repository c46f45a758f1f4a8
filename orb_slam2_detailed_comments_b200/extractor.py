"""Host-side mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:80-250) over the C ABI.

Same constructor arguments, getters and call semantics as the reference class; the work is done by the
sm_100a kernels behind `orb_extract_batch`. No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, MEM_DEVICE, MEM_HOST, OrbParams, check


class ORBextractor:
    HARRIS_SCORE, FAST_SCORE = 0, 1  # include/ORBextractor.h:87-91 (unused by the reference too)

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int, minThFAST: int,
                 device: int = 0, flags: int = 0):
        """include/ORBextractor.h:104 / src/ORBextractor.cc:776-925."""
        self._L = _lib.lib()
        self._h = C.c_void_p()
        p = OrbParams(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, flags)
        check(self._L.orb_extractor_create(C.byref(p), device, C.byref(self._h)))
        self.nfeatures, self.nlevels, self.device = nfeatures, nlevels, device
        self._last_n = 0

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orb_extractor_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    # ---- getters, include/ORBextractor.h:134-182 ----
    def GetLevels(self) -> int:
        return self._L.orb_extractor_get_levels(self._h)

    def GetScaleFactor(self) -> float:
        return self._L.orb_extractor_get_scale_factor(self._h)

    def _table(self, what: int) -> np.ndarray:
        out = np.zeros(self.nlevels, np.float32)
        check(self._L.orb_extractor_get_table(self._h, what, out.ctypes.data_as(C.c_void_p)))
        return out

    def GetScaleFactors(self):
        return self._table(0)

    def GetInverseScaleFactors(self):
        return self._table(1)

    def GetScaleSigmaSquares(self):
        return self._table(2)

    def GetInverseScaleSigmaSquares(self):
        return self._table(3)

    def features_per_level(self) -> np.ndarray:
        out = np.zeros(self.nlevels, np.int32)
        check(self._L.orb_extractor_get_features_per_level(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def max_keypoints(self) -> int:
        return self._L.orb_extractor_max_keypoints(self._h)

    @property
    def launch_count(self) -> int:
        return self._L.orb_extractor_launch_count(self._h)

    STAGES = ("pyramid", "fast_cells", "blur", "quadtree", "describe")

    def set_profiling(self, on: bool):
        check(self._L.orb_extractor_set_profiling(self._h, int(on)))

    def stage_times(self):
        """({stage: summed ms}, internal passes) since set_profiling(True)."""
        ms = (C.c_double * 8)()
        passes = C.c_int64()
        n = check(self._L.orb_extractor_stage_times(self._h, ms, C.byref(passes)))
        return {self.STAGES[i]: ms[i] for i in range(n)}, passes.value

    # ---- operator(), include/ORBextractor.h:122-126 ----
    def __call__(self, image: np.ndarray, mask=None):
        """Returns (keypoints[N] structured cv::KeyPoint records, descriptors[N,32] u8). `mask` is ignored
        like in the reference (include/ORBextractor.h:112). An empty image returns empty outputs (:1966)."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        kps, desc, counts, _ = self.extract_batch(image[None])
        n = int(counts[0])
        return kps[0, :n].copy(), desc[0, :n].copy()

    @staticmethod
    def _frame_pointers(images: np.ndarray) -> np.ndarray:
        """uint64 array of the B frame addresses (the `images` argument of the C ABI), built without a Python loop."""
        base = images.__array_interface__["data"][0]
        return np.uint64(base) + np.arange(images.shape[0], dtype=np.uint64) * np.uint64(images.strides[0])

    def extract_batch(self, images: np.ndarray, out=None):
        """images [B,H,W] u8 host array (rows may be strided). Returns (kps[B,cap], desc[B,cap,32], counts[B],
        level_counts[B,nlevels]) host arrays. `out` may pass preallocated (kps, desc) e.g. pinned buffers."""
        if images.dtype != np.uint8 or images.ndim != 3:
            raise TypeError("images must be uint8 [B,H,W] (CV_8UC1, src/ORBextractor.cc:1972)")
        if images.strides[2] != 1 or images.strides[0] < 0 or images.strides[1] < images.shape[2]:
            images = np.ascontiguousarray(images)
        B, H, W = images.shape
        step = images.strides[1]
        cap = self.max_keypoints()
        if out is None:
            kps = np.empty((B, cap), KP_DTYPE)
            desc = np.empty((B, cap, 32), np.uint8)
        else:
            kps, desc = out
        counts = np.zeros(B, np.int32)
        lcounts = np.zeros((B, self.nlevels), np.int32)
        ptrs = self._frame_pointers(images)
        check(self._L.orb_extract_batch(self._h, ptrs.ctypes.data_as(C.c_void_p), MEM_HOST, step, W, H, B, kps.ctypes.data_as(C.c_void_p),
                                        desc.ctypes.data_as(C.c_void_p), MEM_HOST, cap, counts.ctypes.data_as(C.c_void_p),
                                        lcounts.ctypes.data_as(C.c_void_p)))
        self._last_n = B
        return kps, desc, counts, lcounts

    def extract_stereo_batch(self, left: np.ndarray, right: np.ndarray, bf: float, b: float):
        """Stereo Frame constructor on the device: operator() for left[i] and right[i] (src/Frame.cc:158-168) followed by
        Frame::ComputeStereoMatches (src/Frame.cc:880-1176). left/right [P,H,W] u8. Returns (kps[2P,cap], desc[2P,cap,32],
        counts[2P] — images interleaved L0, R0, L1, ... — uright[P,cap], depth[P,cap], matched[P])."""
        if left.shape != right.shape or left.dtype != np.uint8 or right.dtype != np.uint8 or left.ndim != 3:
            raise TypeError("left/right must be uint8 [P,H,W] of one shape")
        P, H, W = left.shape
        inter = np.empty((2 * P, H, W), np.uint8)
        inter[0::2], inter[1::2] = left, right
        return self.extract_stereo_interleaved(inter, bf, b)

    def extract_stereo_interleaved(self, images: np.ndarray, bf: float, b: float, out=None):
        """Same with the images already interleaved [2P,H,W] (L0, R0, L1, R1, ...). `out` may pass preallocated
        (kps[2P,cap], desc[2P,cap,32], uright[P,cap], depth[P,cap]) e.g. pinned buffers."""
        if images.dtype != np.uint8 or images.ndim != 3 or images.shape[0] % 2 or not images.flags.c_contiguous:
            raise TypeError("images must be a C-contiguous uint8 [2P,H,W] array")
        B, H, W = images.shape
        P = B // 2
        cap = self.max_keypoints()
        if out is None:
            kps, desc = np.empty((B, cap), KP_DTYPE), np.empty((B, cap, 32), np.uint8)
            ur, dp = np.empty((P, cap), np.float32), np.empty((P, cap), np.float32)
        else:
            kps, desc, ur, dp = out
        counts = np.zeros(B, np.int32)
        matched = np.zeros(P, np.int32)
        ptrs = self._frame_pointers(images)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        check(self._L.orb_extract_stereo_batch(self._h, vp(ptrs), MEM_HOST, W, W, H, P, vp(kps), vp(desc), MEM_HOST, cap, vp(counts), None,
                                               bf, b, vp(ur), vp(dp), vp(matched)))
        self._last_n = B
        return kps, desc, counts, ur, dp, matched

    def extract_batch_device(self, images, kps=None, desc=None, counts=None, sync: bool = True):
        """Device-resident path: `images` is a torch.uint8 CUDA tensor [B,H,W] (last dim contiguous, uniform strides).
        Returns torch tensors (kps [B,cap,7] float32 view of cv::KeyPoint records, desc [B,cap,32] u8, counts [B] i32)."""
        import torch
        assert images.is_cuda and images.dtype == torch.uint8 and images.dim() == 3 and images.stride(2) == 1
        B, H, W = images.shape
        assert images.stride(0) == images.stride(1) * H
        cap = self.max_keypoints()
        dev = images.device
        if kps is None:
            kps = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
        if desc is None:
            desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
        if counts is None:
            counts = torch.empty((B,), dtype=torch.int32, device=dev)
        check(self._L.orb_extract_batch_device(self._h, images.data_ptr(), images.stride(1), W, H, B, kps.data_ptr(),
                                               desc.data_ptr(), cap, counts.data_ptr(), int(sync)))
        self._last_n = B
        return kps, desc, counts

    def check(self):
        """orb_extractor_check: waits for queued work; raises OrbError(ORB_E_CAPACITY) if a call since the last check overflowed a capacity
        (what an asynchronous extract_batch_device(sync=False) caller has to poll)."""
        check(self._L.orb_extractor_check(self._h))

    # ---- mvImagePyramid, include/ORBextractor.h:185 ----
    def pyramid_level(self, level: int, index: int = 0, blurred: bool = False) -> np.ndarray:
        c, r = C.c_int(), C.c_int()
        check(self._L.orb_extractor_pyramid_dims(self._h, level, C.byref(c), C.byref(r)))
        out = np.empty((r.value, c.value), np.uint8)
        check(self._L.orb_extractor_pyramid_read(self._h, index, level, int(blurred), out.ctypes.data_as(C.c_void_p), c.value))
        return out

    @property
    def mvImagePyramid(self):
        """Padded CV_8UC1 levels of the last image, fetched lazily from device memory (read by
        Frame::ComputeStereoMatches, src/Frame.cc:910, 1040)."""
        return [self.pyramid_level(l, max(self._last_n - 1, 0)) for l in range(self.nlevels)]

    def raw_keypoints(self, level: int, index: int = 0) -> np.ndarray:
        cap = 1 << 20
        out = np.zeros(cap, KP_DTYPE)
        n = check(self._L.orb_extractor_raw_keypoints(self._h, index, level, out.ctypes.data_as(C.c_void_p), cap))
        return out[:n].copy()
