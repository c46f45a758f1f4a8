"""The monocular Frame constructor's data-parallel part as one device-resident pipeline (src/Frame.cc:338-420):
ExtractORB -> UndistortKeyPoints -> ComputeBoW, every stage consuming the previous stage's device buffers on the extractor's CUDA
stream; the host waits once after extraction (error flag) and once before the results are copied back. (AssignFeaturesToGrid happens inside the guided search, which takes mvKeysUn.)
torch is used for device buffers and the stream only."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, MEM_DEVICE, OrbCamera, check


class FrameFrontEnd:
    def __init__(self, extractor, fx, fy, cx, cy, dist, vocabulary=None, levelsup: int = 4):
        self.ex, self.voc, self.levelsup = extractor, vocabulary, levelsup
        self.cam = OrbCamera.make(fx, fy, cx, cy, dist)
        self._L = _lib.lib()

    def process(self, images: np.ndarray) -> dict:
        """images [B,H,W] uint8 (host). Returns host arrays: keys / keys_un [B,cap] KP_DTYPE, desc [B,cap,32], counts [B], and with a
        vocabulary the orb_bow_transform outputs (bow_word, bow_value, bow_count, fv_node, fv_start, fv_feature, fv_count)."""
        import torch
        B, H, W = images.shape
        cap = self.ex.max_keypoints()
        st_ptr = self._L.orb_extractor_stream(self.ex._h)   # every stage runs on the extractor's stream
        vp = lambda t: C.c_void_p(t.data_ptr())
        # torch only allocates here (its own stream is never used for the kernels); the two device-wide waits below keep its
        # caching allocator and the extractor's stream apart
        d_img = torch.from_numpy(np.ascontiguousarray(images)).cuda()
        torch.cuda.current_stream().synchronize()   # the copy ran on torch's stream; the extractor's stream is non-blocking
        kps, desc, counts = self.ex.extract_batch_device(d_img, sync=True)   # sync: the capacity / error flag is checked here
        kps_un = torch.empty_like(kps)
        check(self._L.orb_undistort_keypoints(MEM_DEVICE, C.byref(self.cam), vp(kps), vp(counts), cap, B, vp(kps_un), C.c_void_p(st_ptr)))
        out = {"keys": kps, "keys_un": kps_un, "desc": desc, "counts": counts}
        if self.voc is not None:
            dev = d_img.device
            i32 = lambda *s: torch.empty(s, dtype=torch.int32, device=dev)
            r = dict(bow_word=i32(B, cap), bow_count=i32(B), fv_node=i32(B, cap), fv_start=i32(B, cap + 1), fv_feature=i32(B, cap),
                     fv_count=i32(B), bow_value=torch.empty((B, cap), dtype=torch.float64, device=dev))
            check(self._L.orb_bow_transform(self.voc._h, MEM_DEVICE, vp(desc), vp(counts), cap, B, self.levelsup, None, None,
                                            vp(r["bow_word"]), vp(r["bow_value"]), vp(r["bow_count"]), vp(r["fv_node"]), vp(r["fv_start"]),
                                            vp(r["fv_feature"]), vp(r["fv_count"]), C.c_void_p(st_ptr)))
            out.update(r)
        torch.cuda.synchronize()
        host = {k: v.cpu() for k, v in out.items()}
        res = {k: v.numpy() for k, v in host.items()}
        for k in ("keys", "keys_un"):
            res[k] = res[k].view(np.uint8).reshape(B, cap, 28).view(KP_DTYPE).reshape(B, cap)
        return res
