#!/usr/bin/env python
"""Host-to-host throughput of the LocalMapping / LoopClosing matchers built on the guided-search kernels:
ORBmatcher::SearchForTriangulation (orb_search_for_triangulation) and the candidate search of ORBmatcher::Fuse (orb_fuse_search),
next to the CPU oracle (the checker and CPU baseline here: this script lives under tests/ because it runs the oracle) on all host cores (ThreadPool over pairs; the ctypes calls release the GIL). Prints one JSON line.
usage: python tests/matcher_bench.py [--pairs 256] [--features 2000] [--steps 5]"""
import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=256)
    ap.add_argument("--features", type=int, default=2000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--cpu-pairs", type=int, default=64)
    args = ap.parse_args()
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200 import synth
    from orb_slam2_detailed_comments_b200._lib import KP_DTYPE, EPI_DTYPE
    from oracle import oracle_py
    import test_search_triangulation as tst
    import test_fuse_search as tfs
    from test_search_by_bow import _csr

    P, N = args.pairs, args.features
    cores = len(os.sched_getaffinity(0))
    voc = synth.synth_vocabulary(10, 4, seed=21, prune=0.0)
    gv, ov = pkg.ORBVocabulary(*voc), oracle_py.OracleVocabulary(*voc)
    distinct = 8
    cases = [tst._pair(voc, 500 + i, N, N) for i in range(distinct)]
    D1 = np.stack([cases[i % distinct][0] for i in range(P)]); K1 = np.stack([cases[i % distinct][1] for i in range(P)])
    H1 = np.stack([cases[i % distinct][2] for i in range(P)]); U1 = np.stack([cases[i % distinct][3] for i in range(P)])
    D2 = np.stack([cases[i % distinct][4] for i in range(P)]); K2 = np.stack([cases[i % distinct][5] for i in range(P)])
    H2 = np.stack([cases[i % distinct][6] for i in range(P)]); U2 = np.stack([cases[i % distinct][7] for i in range(P)])
    E = np.stack([cases[i % distinct][8] for i in range(P)]).astype(EPI_DTYPE)
    n = np.full(P, N, np.int32)
    bow1, bow2 = gv.transform(D1, n, 2), gv.transform(D2, n, 2)
    m = pkg.ORBmatcher(0.6, True)
    m.SearchForTriangulation(D1, K1, H1, U1, bow1, D2, K2, H2, U2, bow2, E)   # warm
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m12, nm = m.SearchForTriangulation(D1, K1, H1, U1, bow1, D2, K2, H2, U2, bow2, E)
    tri_dt = (time.perf_counter() - t0) / args.steps
    # the same call with every array resident in HBM (what a device-side Frame / KeyFrame pipeline sees), CUDA events on the current stream
    import ctypes as C
    import torch
    from orb_slam2_detailed_comments_b200._lib import lib, check, MEM_DEVICE
    L = lib()
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    ptr = lambda t: C.c_void_p(t.data_ptr())
    free1 = (H1 == 0).astype(np.uint8); blocked2 = (H2 != 0).astype(np.uint8)
    tri_in = [dev(x) for x in (D1, K1, free1, U1.astype(np.float32), bow1["fv_node"], bow1["fv_start"], bow1["fv_feature"], bow1["fv_count"],
                               D2, K2, blocked2, U2.astype(np.float32), bow2["fv_node"], bow2["fv_start"], bow2["fv_feature"], bow2["fv_count"], E)]
    d_m12 = torch.empty(P * N, dtype=torch.int32, device="cuda"); d_nm = torch.empty(P, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def tri_dev():
        a = [ptr(t) for t in tri_in]
        check(L.orb_search_for_triangulation(MEM_DEVICE, *a[:8], N, *a[8:16], N, P, a[16], 50, 1, ptr(d_m12), ptr(d_nm), st))

    def timed(fn):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps * 1e-3
    tri_dev_dt = timed(tri_dev)
    assert np.array_equal(d_nm.cpu().numpy(), nm), "device-memory mode differs from host mode"
    # CPU oracle, same pairs
    fvs = [(_csr(ov.transform(c[0], 2)[3], N), _csr(ov.transform(c[4], 2)[3], N)) for c in cases]

    def cpu_tri(i):
        c, (f1, f2) = cases[i % distinct], fvs[i % distinct]
        e = c[8]
        return oracle_py.search_for_triangulation(c[0], c[1], c[2], c[3], f1, c[4], c[5], c[6], c[7], f2, e["f12"], e["ex"], e["ey"],
                                                  e["level_sigma2"], e["scale_factor"], False, 50, True)[1]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        cpu_nm = list(ex.map(cpu_tri, range(args.cpu_pairs)))
    cpu_tri_dt = time.perf_counter() - t0
    assert [int(x) for x in nm[:distinct]] == cpu_nm[:distinct], "GPU and oracle disagree"

    # Fuse candidate search: N projected points against N keypoints per keyframe, chi-square test on
    fc = [tfs._case(700 + i, N, N, True) for i in range(distinct)]
    K = np.stack([fc[i % distinct]["kps"] for i in range(P)]); D = np.stack([fc[i % distinct]["desc"] for i in range(P)])
    Q = np.stack([fc[i % distinct]["queries"] for i in range(P)]); QD = np.stack([fc[i % distinct]["qdesc"] for i in range(P)])
    U = np.stack([fc[i % distinct]["uright"] for i in range(P)])
    m.FuseSearch(K, D, n, Q, QD, n, fc[0]["bounds"], uright=U, inv_level_sigma2=tfs.INV_SIGMA2)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        bi, bd = m.FuseSearch(K, D, n, Q, QD, n, fc[0]["bounds"], uright=U, inv_level_sigma2=tfs.INV_SIGMA2)
    fuse_dt = (time.perf_counter() - t0) / args.steps
    from orb_slam2_detailed_comments_b200._lib import OrbSearchParams
    fuse_in = [dev(x) for x in (K, D, U.astype(np.float32), n, Q, QD, n)]
    d_bi = torch.empty(P * N, dtype=torch.int32, device="cuda"); d_bd = torch.empty(P * N, dtype=torch.int32, device="cuda")
    b = fc[0]["bounds"]
    prm = OrbSearchParams(float(b[0]), float(b[1]), float(b[2]), float(b[3]), 50, 0.0, 0, 0)
    sig = np.zeros(16, np.float32); sig[:8] = tfs.INV_SIGMA2

    def fuse_dev():
        a = [ptr(t) for t in fuse_in]
        check(L.orb_fuse_search(MEM_DEVICE, a[0], a[1], a[2], a[3], N, a[4], a[5], a[6], N, P, C.byref(prm), sig.ctypes.data_as(C.c_void_p), ptr(d_bi),
                                ptr(d_bd), st))
    fuse_dev_dt = timed(fuse_dev)
    assert np.array_equal(d_bi.cpu().numpy().reshape(P, N), bi), "device-memory mode differs from host mode"

    def cpu_fuse(i):
        c = fc[i % distinct]
        return int((oracle_py.fuse_search(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], 50, c["uright"], tfs.INV_SIGMA2)[0] >= 0).sum())
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        cpu_f = list(ex.map(cpu_fuse, range(args.cpu_pairs)))
    cpu_fuse_dt = time.perf_counter() - t0
    assert [int((bi[i] >= 0).sum()) for i in range(distinct)] == cpu_f[:distinct], "GPU and oracle disagree"

    print(json.dumps({
        "config": f"{P} keyframe pairs x {N} features each, host arrays in and out (H2D/D2H inside), vocabulary k=10 L=4, levelsup 2",
        "search_for_triangulation": {"value": P / tri_dt, "unit": "keyframe pairs/s", "ms_per_call": 1e3 * tri_dt,
                                     "resident": {"value": P / tri_dev_dt, "ms_per_call": 1e3 * tri_dev_dt},
                                     "matches_per_pair": float(np.mean(nm)),
                                     "cpu_baseline": {"value": args.cpu_pairs / cpu_tri_dt, "cores": cores, "kind": "port",
                                                      "sample": f"{args.cpu_pairs} pairs, {cpu_tri_dt:.2f} s wall"}},
        "fuse_search": {"value": P / fuse_dt, "unit": "keyframes/s", "ms_per_call": 1e3 * fuse_dt,
                        "resident": {"value": P / fuse_dev_dt, "ms_per_call": 1e3 * fuse_dev_dt},
                        "fused_per_keyframe": float(np.mean((bi >= 0).sum(axis=1))),
                        "cpu_baseline": {"value": args.cpu_pairs / cpu_fuse_dt, "cores": cores, "kind": "port",
                                         "sample": f"{args.cpu_pairs} keyframes, {cpu_fuse_dt:.2f} s wall"}},
    }))


if __name__ == "__main__":
    main()
