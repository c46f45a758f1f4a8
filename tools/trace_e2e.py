"""Timeline of orb_extract_batch's copy / compute pipeline (ORB_B200_TRACE=1) for the KITTI bench workload, with the
input and the output side switched between host (pinned) and device memory to see what each copy direction costs."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import orb_slam2_detailed_comments_b200 as pkg
from orb_slam2_detailed_comments_b200 import synth
from orb_slam2_detailed_comments_b200._lib import lib, check
B, W, H = 512, 1241, 376
frames = synth.synth_batch(W, H, B, seed0=1234, distinct=8)
ex = pkg.ORBextractor(2000, 1.2, 8, 20, 7)
cap = ex.max_keypoints()
L = lib()
pin_in = torch.from_numpy(frames).pin_memory()
dev_in = pin_in.cuda()
pk = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
pd = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
dk, dd = pk.cuda(), pd.cuda()
counts = np.zeros(B, np.int32)
def run(in_dev, out_dev):
    src = dev_in if in_dev else pin_in
    ptrs = (src.data_ptr() + np.arange(B, dtype=np.uint64) * np.uint64(W * H)).astype(np.uint64)
    k, d = (dk, dd) if out_dev else (pk, pd)
    t0 = time.perf_counter()
    check(L.orb_extract_batch(ex._h, ptrs.ctypes.data_as(C.c_void_p), int(in_dev), W, W, H, B, C.c_void_p(k.data_ptr()), C.c_void_p(d.data_ptr()),
                              int(out_dev), cap, counts.ctypes.data_as(C.c_void_p), None))
    return 1e3 * (time.perf_counter() - t0)
for in_dev, out_dev in ((0, 0), (1, 0), (0, 1), (1, 1)):
    os.environ.pop("ORB_B200_TRACE", None)
    for _ in range(3):
        run(in_dev, out_dev)
    ms = min(run(in_dev, out_dev) for _ in range(5))
    print(f"input {'device' if in_dev else 'host'}, output {'device' if out_dev else 'host'}: {ms:.3f} ms per {B} frames", flush=True)
    os.environ["ORB_B200_TRACE"] = "1"
    run(in_dev, out_dev)
