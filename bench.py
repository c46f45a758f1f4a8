#!/usr/bin/env python
"""Benchmark of the ORB front-end hot path on B200 (contract: see the task statement / DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload kitti|tum|euroc|4k]

One "step" = ORBextractor::operator() over one batch of synthetic frames (default: BASELINE.json configs[1],
KITTI-shaped 1241x376 stereo pairs, nFeatures 2000, left+right images batched on the GPU). Metric = images
(frames) per second, whole job. `value` is measured with the frames resident in HBM; `e2e` goes through the
reference-facing host API (orb_extract_batch) with pinned HOST buffers, H2D and D2H inside the timed region.
A secondary object `knn` reports the ORBmatcher path (config C4: 2000 queries x 1M database rows, k=2).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from orb_slam2_detailed_comments_b200 import geometry, synth  # noqa: E402

METRIC = "ORB frames/sec (extract+describe)"
UNIT = "images/s"
WORKLOADS = {
    # name: (config label, width, height, nFeatures, images per step per GPU)
    "kitti": ("KITTI-shaped 1241x376 stereo pairs, nFeatures=2000, 8 levels, scale 1.2, FAST 20/7, L+R batched", 1241, 376, 2000, 512),
    "tum": ("TUM-shaped 640x480 mono, nFeatures=1000, 8 levels, scale 1.2, FAST 20/7", 640, 480, 1000, 1024),
    "euroc": ("EuRoC-shaped 752x480 stereo, nFeatures=1200, 256 frame pairs per GPU", 752, 480, 1200, 512),
    "4k": ("4K 3840x2160, nFeatures=8000, 8 levels", 3840, 2160, 8000, 16),
}
KNN_Q, KNN_D = 2000, 1_000_000


def shared_config(workload: str) -> dict:
    """The `config` object BOTH arms print, key for key and value for value (the driver compares them): what is computed, on which synthetic
    input. Everything that differs between the arms (images per step, timing method, parallelism) goes to the line's `run` object."""
    label, W, H, nf, B = WORKLOADS[workload]
    return {"workload": label, "width": W, "height": H, "nfeatures": nf, "levels": 8, "scale_factor": 1.2, "fast_thresholds": [20, 7],
            "frames": "synth.synth_batch: 64 distinct seeded frames per rank (8 at 4K), tiled with per-copy shifts",
            "l2": "GPU arm: inputs larger than L2 (%.0f MB of frames per step per GPU, + pyramid workspaces), no flush needed; "
                  "reference arm: host cores, not applicable" % (B * W * H / 1e6)}


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


# ---- CPU legs (the only place bench.py executes oracle/: cpu_baseline and --impl reference) ------------------
def _host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_extract_rate(width, height, nf, n_images, threads, seed0=4321):
    """Oracle (CPU restatement of the reference, kind "port") over `n_images` frames, one extractor per thread."""
    from oracle import oracle_py
    frames = synth.synth_batch(width, height, n_images, seed0=seed0, distinct=min(8, n_images))
    oracle_py.extract_many(frames[:min(threads, n_images)], nf, threads)   # warm: page in, spawn
    t0 = time.perf_counter()
    total = oracle_py.extract_many(frames, nf, threads)
    dt = time.perf_counter() - t0
    return n_images / dt, dt, int(total)


def cpu_knn_rate(threads, q_n=256, d_n=200_000):
    from oracle import oracle_py
    q, db = synth.synth_descriptors(q_n, d_n, seed=7)
    t0 = time.perf_counter()
    oracle_py.hamming_knn(q, db, nthreads=threads)
    dt = time.perf_counter() - t0
    return q_n * d_n / dt, dt


REF_BIN = os.path.join(ROOT, "oracle", "_ref", "orb_ref_extract")


def _ref_binary_rate(frames, nf, threads, repeat):
    """oracle/_ref/orb_ref_extract --time: the reference's own src/ORBextractor.cc (compiled unmodified against oracle/cvshim,
    DESIGN.md §2), one ORBextractor per thread over the frames. Returns (images, seconds)."""
    import struct
    import subprocess
    import tempfile
    n, H, W = frames.shape
    with tempfile.TemporaryDirectory() as d:
        fin = os.path.join(d, "in.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<8if", 0x4F524231, n, W, H, nf, 8, 20, 7, 1.2))
            f.write(np.ascontiguousarray(frames, np.uint8).tobytes())
        out = None
        for exe in (REF_BIN + "_avx2", REF_BIN):   # -O3 + AVX2 build (the reference's CMake asks for -march=native); plain -O3 if it cannot run here
            if not os.access(exe, os.X_OK):
                continue
            try:
                out = subprocess.run([exe, "--time", fin, str(threads), str(repeat)], check=True, capture_output=True, text=True, timeout=900)
                break
            except subprocess.CalledProcessError:
                continue
        if out is None:
            raise RuntimeError("oracle/_ref/orb_ref_extract did not run")
    images, sec, _ = out.stdout.split()
    return int(images), float(sec)


def run_reference(args):
    """--impl reference: the reference's CPU path on all host cores, same config/metric/unit. Rank 0 only.
    kind "reference" = oracle/_ref/orb_ref_extract (the reference's own ORBextractor.cc; its OpenCV primitives are the cv2-pinned
    restatements of oracle/cv_primitives.cpp, because the image has no OpenCV C++); kind "port" = the oracle, if the binary is absent."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    label, W, H, nf, _ = WORKLOADS[args.workload]
    cores = _host_cores()
    n = max(2 * cores, 16)   # images per step: a bounded sample of the workload
    frames = synth.synth_batch(W, H, n, seed0=4321, distinct=min(64, n))
    if os.access(REF_BIN, os.X_OK):
        kind = "reference"
        how = "one reference ORBextractor (src/ORBextractor.cc compiled -O3 + AVX2 against oracle/cvshim) per thread"
        if args.warmup:
            _ref_binary_rate(frames, nf, cores, args.warmup)
        images, dt = _ref_binary_rate(frames, nf, cores, args.steps)   # the binary warms one pass itself, then times `steps` passes
        assert images == n * args.steps
    else:
        kind, how = "port", "one oracle extractor per thread"
        from oracle import oracle_py
        for _ in range(args.warmup):
            oracle_py.extract_many(frames, nf, cores)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            oracle_py.extract_many(frames, nf, cores)
        dt = time.perf_counter() - t0
    v = n * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": shared_config(args.workload),
        "run": {"images_per_step": n, "timing": "host wall clock", "ranks": "rank 0's host cores only, whatever --gpus is: at N > 1 the ratio is "
                "N GPUs against ONE host's cores"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{n} images per step x {args.steps} steps, {how}"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- GPU arm ---------------------------------------------------------------------------------------------
_NUMA = None


def _bind_to_gpu_numa_node(local):
    """Pin this rank to the CPUs of its GPU's NUMA node before any pinned buffer is allocated (first touch puts the
    pages on that node), so that N ranks do not pull their frames across the socket interconnect. Best effort: a box
    without the sysfs entries stays unbound."""
    global _NUMA
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            _NUMA = {"node": node, "cpus": len(cpus)}
    except Exception:
        pass


def _dist_setup(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        _bind_to_gpu_numa_node(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def _barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def bench_extract_resident(ex, frames_dev, steps, warmup, world):
    """`value`: frames already in HBM, outputs left in HBM; timed with CUDA events on the extractor's stream."""
    import torch
    st = torch.cuda.ExternalStream(ex._L.orb_extractor_stream(ex._h))
    kps, desc, counts = ex.extract_batch_device(frames_dev, sync=True)
    for _ in range(warmup):
        ex.extract_batch_device(frames_dev, kps, desc, counts, sync=False)
    torch.cuda.synchronize()
    l0 = ex.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    e0.record(st)
    for _ in range(steps):
        ex.extract_batch_device(frames_dev, kps, desc, counts, sync=False)
    e1.record(st)
    e1.synchronize()
    _barrier(world)
    ms = _max_over_ranks(e0.elapsed_time(e1), world)
    launches = ex.launch_count - l0
    # stage breakdown: a second, untimed run of the same steps with per-stage CUDA events; profiling serialises the passes
    # on one stream (the timed run above overlaps consecutive passes on two streams), so the stage sum exceeds ms
    ex.set_profiling(True)
    for _ in range(steps):
        ex.extract_batch_device(frames_dev, kps, desc, counts, sync=False)
    torch.cuda.synchronize()
    stage_ms, passes = ex.stage_times()
    ex.set_profiling(False)
    return ms, launches, stage_ms, passes, int(counts.sum().item())


def bench_extract_e2e(ex, frames_host, steps, warmup, world):
    """`e2e`: orb_extract_batch with pinned HOST frames in, HOST keypoints/descriptors out (copies inside the call)."""
    import torch
    B, H, W = frames_host.shape
    cap = ex.max_keypoints()
    pin_in = torch.from_numpy(frames_host).pin_memory()
    pin_k = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    pin_d = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    from orb_slam2_detailed_comments_b200 import KP_DTYPE
    k_np = pin_k.numpy().view(KP_DTYPE).reshape(B, cap)
    out = (k_np, pin_d.numpy())
    src = pin_in.numpy()
    for _ in range(max(warmup, 1)):
        ex.extract_batch(src, out=out)
    _barrier(world)
    t0 = time.perf_counter()
    nk = 0
    for _ in range(steps):
        _, _, counts, _ = ex.extract_batch(src, out=out)
        nk += int(counts.sum())
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    _barrier(world)
    ms = _max_over_ranks(dt * 1e3, world)
    h2d = B * H * W
    d2h = B * cap * 60 + B * 4 * (1 + ex.nlevels)
    return ms, h2d, d2h, nk


def bench_latency(ex, frames_host, iters=200):
    """Per-call latency the way Frame::ExtractORB calls the extractor (src/Frame.cc:468-481): ONE image per call, pinned host buffers
    in and out, wall clock around the public call. Reported beside the throughput numbers, not part of them."""
    import torch
    _, H, W = frames_host.shape
    cap = ex.max_keypoints()
    n = min(16, len(frames_host))
    src = torch.from_numpy(frames_host[:n].copy()).pin_memory().numpy()
    from orb_slam2_detailed_comments_b200 import KP_DTYPE
    out = (torch.empty((1, cap, 28), dtype=torch.uint8).pin_memory().numpy().view(KP_DTYPE).reshape(1, cap),
           torch.empty((1, cap, 32), dtype=torch.uint8).pin_memory().numpy())
    for i in range(20):
        ex.extract_batch(src[i % n:i % n + 1], out=out)
    ts = []
    for i in range(iters):
        t0 = time.perf_counter()
        ex.extract_batch(src[i % n:i % n + 1], out=out)
        ts.append(time.perf_counter() - t0)
    ts.sort()
    return {"metric": "wall clock of one ORBextractor::operator() call (1 image, host in / host out)", "unit": "ms",
            "median": 1e3 * ts[len(ts) // 2], "p99": 1e3 * ts[int(len(ts) * 0.99)], "calls": iters}


def bench_cpp_dropin_latency(frame: np.ndarray, nf: int):
    """The call Frame::ExtractORB really makes: the C++ drop-in class (cpp/ORBextractor.cc) through tests/cpp/test_dropin --latency, a separate
    process with its own CUDA context; pageable cv::Mat in, std::vector<cv::KeyPoint> + cv::Mat out, mvImagePyramid filled / not filled."""
    exe = os.path.join(ROOT, "tests", "cpp", "test_dropin")
    if not os.access(exe, os.X_OK):
        return {"unavailable": "tests/cpp/test_dropin not built"}
    H, W = frame.shape
    with tempfile.NamedTemporaryFile(suffix=".raw", delete=False) as f:
        f.write(np.ascontiguousarray(frame).tobytes())
        path = f.name
    try:
        r = subprocess.run([exe, "--latency", path, str(W), str(H), str(nf), "300"], capture_output=True, text=True, timeout=300)
        if r.returncode != 0:
            return {"unavailable": "test_dropin --latency failed: " + (r.stderr or r.stdout)[-300:]}
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:   # noqa: BLE001
        return {"unavailable": str(e)[:300]}
    finally:
        os.unlink(path)


def bench_stereo_e2e(ex, frames_host, steps, warmup, world, bf=386.1448, b=0.5372):
    """Secondary: the stereo Frame constructor (two operator() calls + Frame::ComputeStereoMatches, src/Frame.cc:158-168,
    880-1176) through orb_extract_stereo_batch with pinned host buffers; frames (2p, 2p+1) are treated as left / right."""
    import torch
    from orb_slam2_detailed_comments_b200 import KP_DTYPE
    B, H, W = frames_host.shape
    cap = ex.max_keypoints()
    pin_in = torch.from_numpy(frames_host).pin_memory()
    pin_k = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    pin_d = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    pin_u = torch.empty((B // 2, cap), dtype=torch.float32).pin_memory()
    pin_z = torch.empty((B // 2, cap), dtype=torch.float32).pin_memory()
    out = (pin_k.numpy().view(KP_DTYPE).reshape(B, cap), pin_d.numpy(), pin_u.numpy(), pin_z.numpy())
    src = pin_in.numpy()
    for _ in range(max(warmup, 1)):
        ex.extract_stereo_interleaved(src, bf, b, out=out)
    _barrier(world)
    t0 = time.perf_counter()
    nm = 0
    for _ in range(steps):
        r = ex.extract_stereo_interleaved(src, bf, b, out=out)
        nm += int(r[5].sum())
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    _barrier(world)
    return _max_over_ranks(dt * 1e3, world), nm / steps


GUIDED_FRAMES, GUIDED_N, GUIDED_M = 512, 2000, 2000


def _guided_cases(W, H, distinct=8):
    return [synth.synth_guided_case(900 + s, n=GUIDED_N, m=GUIDED_M, w=W, h=H, radius=15.0, stereo=True) for s in range(distinct)]


def bench_guided(W, H, steps, warmup, world):
    """Secondary: ORBmatcher::SearchByProjection(CurrentFrame, LastFrame) candidate search for 512 independent
    frame pairs per step, 2000 projected map points against 2000 keypoints each, arrays resident in HBM. Returns (ms for
    `steps` steps max over ranks, matches per frame, kernel launches)."""
    import ctypes as C
    import torch
    from orb_slam2_detailed_comments_b200._lib import lib, check, OrbSearchParams, MEM_DEVICE
    cases = _guided_cases(W, H)
    F = GUIDED_FRAMES
    pick = [cases[i % len(cases)] for i in range(F)]
    dev = lambda key, dt: torch.from_numpy(np.stack([c[key] for c in pick]).view(np.uint8).reshape(F, -1)).cuda()
    t = {k: dev(k, None) for k in ("kps", "desc", "uright", "claimed", "queries", "qdesc")}
    nk = torch.full((F,), GUIDED_N, dtype=torch.int32, device="cuda")
    nq = torch.full((F,), GUIDED_M, dtype=torch.int32, device="cuda")
    a = torch.empty((F, GUIDED_N), dtype=torch.int32, device="cuda")
    qm = torch.empty((F, GUIDED_M), dtype=torch.int32, device="cuda")
    nm = torch.empty((F,), dtype=torch.int32, device="cuda")
    prm = OrbSearchParams(*[float(b) for b in cases[0]["bounds"]], 100, 0.0, 1, 0)
    st = torch.cuda.current_stream().cuda_stream
    p = lambda x: C.c_void_p(x.data_ptr())
    L = lib()

    def step():
        check(L.orb_search_by_projection(MEM_DEVICE, p(t["kps"]), p(t["desc"]), p(t["uright"]), p(t["claimed"]), p(nk), GUIDED_N,
                                         p(t["queries"]), p(t["qdesc"]), p(nq), GUIDED_M, F, C.byref(prm), p(a), p(qm), p(nm), C.c_void_p(st)))
    for _ in range(max(warmup, 1)):
        step()
    l0 = L.orb_matcher_launch_count()
    _barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    _barrier(world)
    return _max_over_ranks(ms, world), float(nm.float().mean().item()), L.orb_matcher_launch_count() - l0


def cpu_guided_rate(W, H, threads, frames=4096):
    """The oracle's SearchByProjection on `frames` frame pairs of the same workload, one frame per task on `threads` threads."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle_py
    cases = _guided_cases(W, H)
    oracle_py.search_by_projection(cases[0]["kps"], cases[0]["desc"], cases[0]["queries"], cases[0]["qdesc"], cases[0]["bounds"])   # load + argtypes
    def one(i):
        c = cases[i % len(cases)]
        return oracle_py.search_by_projection(c["kps"], c["desc"], c["queries"], c["qdesc"], c["bounds"], 100, 0.0, True, c["uright"], c["claimed"])[2]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(one, range(frames)))
    dt = time.perf_counter() - t0
    return frames / dt, dt


BOW_FRAMES, BOW_N = 512, 2000


def _bow_inputs():
    voc = synth.synth_vocabulary_full(10, 6, seed=5)
    rng = np.random.default_rng(17)
    leaves = np.flatnonzero(voc[3])
    base = voc[4][leaves[rng.integers(0, len(leaves), (8, BOW_N))]]   # descriptors near words, 8 distinct frames
    base ^= np.packbits(rng.random((8, BOW_N, 256)) < 0.05, axis=2)
    return voc, base


def bench_bow(steps, warmup, world, voc, base):
    """Secondary: Frame::ComputeBoW for 512 frames x 2000 descriptors per step against a 10^6-word vocabulary
    (k = 10, L = 6, the ORB vocabulary's shape), descriptors resident in HBM."""
    import ctypes as C
    import torch
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200._lib import lib, check, MEM_DEVICE
    gv = pkg.ORBVocabulary(*voc, device=torch.cuda.current_device())
    F, cap = BOW_FRAMES, BOW_N
    desc = torch.from_numpy(np.stack([base[i % len(base)] for i in range(F)])).cuda()
    n = torch.full((F,), BOW_N, dtype=torch.int32, device="cuda")
    i32 = lambda *shape: torch.empty(shape, dtype=torch.int32, device="cuda")
    word, node, bw, bc, fn, fs, ff, fc = i32(F, cap), i32(F, cap), i32(F, cap), i32(F), i32(F, cap), i32(F, cap + 1), i32(F, cap), i32(F)
    bv = torch.empty((F, cap), dtype=torch.float64, device="cuda")
    p = lambda x: C.c_void_p(x.data_ptr())
    st = torch.cuda.current_stream().cuda_stream
    L = lib()

    def step():
        check(L.orb_bow_transform(gv._h, MEM_DEVICE, p(desc), p(n), cap, F, 4, p(word), p(node), p(bw), p(bv), p(bc), p(fn), p(fs), p(ff), p(fc),
                                  C.c_void_p(st)))
    for _ in range(max(warmup, 1)):
        step()
    _barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    _barrier(world)
    return _max_over_ranks(ms, world), float(bc.float().mean().item())


def cpu_bow_rate(voc, base, threads, frames=4096):
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle_py
    ov = oracle_py.OracleVocabulary(*voc)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda i: len(ov.transform(base[i % len(base)], 4)[2]), range(frames)))
    dt = time.perf_counter() - t0
    return frames / dt, dt


def bench_sustained(ex, frames_dev, B, world, local, rank, seconds=2.0):
    """The headline workload back to back for >= `seconds` (the default timed region is ~0.1 s: too short for the power / clock governor to show):
    resident images/s over the whole loop with its own clock record."""
    import torch
    st = torch.cuda.ExternalStream(ex._L.orb_extractor_stream(ex._h))
    kps, desc, counts = ex.extract_batch_device(frames_dev, sync=True)
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    t0 = time.perf_counter()
    e0.record(st)
    steps = 0
    while True:
        for _ in range(8):
            ex.extract_batch_device(frames_dev, kps, desc, counts, sync=False)
        steps += 8
        e1.record(st)
        e1.synchronize()
        done = time.perf_counter() - t0 >= seconds
        if world > 1:   # every rank runs the same number of steps
            import torch.distributed as dist
            t = torch.tensor([0 if done else 1], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            done = int(t.item()) == 0
        if done:
            break
    _barrier(world)
    ms = _max_over_ranks(e0.elapsed_time(e1), world)
    clocks = sampler.stop() if sampler else {}
    return {"value": B * world * steps / (ms * 1e-3), "unit": UNIT, "steps": steps, "seconds": ms * 1e-3, "clocks": clocks,
            "what": "the headline workload back to back, frames resident in HBM, CUDA events around the whole loop"}


def bench_other_config(name, rank, world, local, steps=4, warmup=3):
    """One of the other BASELINE configs (TUM, EuRoC, 4K) in the same line: resident value, e2e, pipeline roofline fraction, clocks."""
    import torch
    import orb_slam2_detailed_comments_b200 as pkg
    label, W, H, nf, B = WORKLOADS[name]
    frames = synth.synth_batch(W, H, B, seed0=777 + 100000 * rank, distinct=min(32, B) if name != "4k" else min(4, B))
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, device=local)
    frames_dev = torch.from_numpy(frames).cuda()
    sampler = ClockSampler(local) if rank == 0 else None
    ms, launches, stage_ms, passes, nkp = bench_extract_resident(ex, frames_dev, steps, warmup, world)
    clocks = sampler.stop() if sampler else {}
    value = B * world * steps / (ms * 1e-3)
    e_ms, h2d, d2h, _ = bench_extract_e2e(ex, frames, 2, 1, world)
    ex.close()
    del frames_dev
    torch.cuda.empty_cache()
    peak, _ = _peaks()
    A = geometry.algorithmic_bytes(W, H, nf)
    return {"config": shared_config(name), "images_per_step_per_gpu": B, "steps": steps, "value": value, "unit": UNIT, "ms_per_step": ms / steps,
            "e2e": {"value": B * world * 2 / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": 2},
            "pipeline": {"algorithmic_bytes_per_image": A, "achieved": value / world * A / 1e9, "frac": value / world * A / 1e9 / peak},
            "stages_ms_per_step": {k: v / steps for k, v in stage_ms.items()}, "keypoints_per_image": nkp / B,
            "gpu_launches": int(launches), "clocks": clocks}


def knn_parity(rank, world, out_merged, q, nsample=64):
    """N > 1: the merged result of the sharded scan (all_gather + device merge) == ONE scan of rank 0 over the concatenated shards, for the
    first `nsample` queries. Rank 0 regenerates every rank's shard from its generator seed (same device type, same seed -> same bytes)."""
    import torch
    from orb_slam2_detailed_comments_b200 import matcher
    if world == 1 or rank != 0:
        return None
    per = KNN_D // world
    shards = []
    for r in range(world):
        g = torch.Generator(device="cuda").manual_seed(100 + r)
        shards.append(torch.randint(0, 256, (per, 32), dtype=torch.uint8, device="cuda", generator=g))
    full = torch.cat(shards)
    one = matcher.hamming_knn_device(q[:nsample].contiguous(), full)
    torch.cuda.synchronize()
    return bool(torch.equal(one, out_merged[:nsample]))


def bench_knn(rank, world, steps, warmup):
    """Config C4: 2000 queries x 1M database rows, k=2; database sharded contiguously over ranks. pairs/s = Q*D / time (max over ranks).
    world > 1: both exchange paths are timed, each behind ONE C-ABI call per rank - orb_hamming_knn_peer (the reduction kernel stores the records
    into the peers' mailboxes over NVLink and merges: no collective launch) and orb_hamming_knn_sharded (shard scan + ncclAllGather + merge
    kernel). The faster one with parity true is the reported value; both are listed under `variants`."""
    import torch
    from orb_slam2_detailed_comments_b200 import _lib, matcher
    rng = np.random.RandomState(7)
    per = KNN_D // world
    base = rank * per
    q = torch.from_numpy(rng.randint(0, 256, (KNN_Q, 32), dtype=np.uint8)).cuda()
    gdb = torch.Generator(device="cuda").manual_seed(100 + rank)
    db = torch.randint(0, 256, (per, 32), dtype=torch.uint8, device="cuda", generator=gdb)
    out = torch.empty((KNN_Q, 3), dtype=torch.int64, device="cuda")

    def timed(step):
        for _ in range(max(warmup, 3)):
            step()
        l0 = _lib.lib().orb_matcher_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        _barrier(world)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        e1.synchronize()
        _barrier(world)
        ms = _max_over_ranks(e0.elapsed_time(e1), world)
        launches = _lib.lib().orb_matcher_launch_count() - l0
        return KNN_Q * per * world * steps / (ms * 1e-3), ms / steps, launches, knn_parity(rank, world, out, q)

    if world == 1:
        r = timed(lambda: matcher.hamming_knn_device(q, db, out=out, index_base=base))
        return r + (None, None)
    variants = {}
    comm = matcher.nccl_comm_ptr()   # torch.distributed's ncclComm_t, handed to the C ABI
    r_nccl = timed(lambda: matcher.hamming_knn_sharded(q, db, base, out=out, comm=comm))
    variants["nccl_allgather"] = {"api": "orb_hamming_knn_sharded", "value": r_nccl[0], "ms_per_step": r_nccl[1], "gpu_launches": int(r_nccl[2]),
                                  "parity": r_nccl[3]}
    best, name = r_nccl, "orb_hamming_knn_sharded: shard scan + ncclAllGather + device top-2 merge in one C-ABI call"
    try:
        grp = matcher.KnnPeerGroup(max_q=KNN_Q)
        out.zero_()
        r_peer = timed(lambda: grp.knn(q, db, base, out=out))
        bad = 0.0
        try:
            grp.check()
        except Exception:   # noqa: BLE001 - a wait that timed out on this rank: every rank has to learn it (the ranks stay in step)
            bad = 1.0
        if _max_over_ranks(bad, world) > 0.0:
            raise RuntimeError("a peer flag did not arrive on some rank")
        variants["peer_stores"] = {"api": "orb_hamming_knn_peer", "value": r_peer[0], "ms_per_step": r_peer[1], "gpu_launches": int(r_peer[2]),
                                   "parity": r_peer[3]}
        ok = _max_over_ranks(0.0 if (rank != 0 or r_peer[3]) else 1.0, world) == 0.0   # rank 0 holds the parity verdict
        if ok and r_peer[0] > r_nccl[0]:
            best, name = r_peer, ("orb_hamming_knn_peer: shard scan, then ONE kernel that reduces, stores the records into the peers' mailboxes over "
                                  "NVLink (CUDA IPC), waits for their flags and merges - no collective launch")
        _barrier(world)
        grp.close()
    except Exception as e:   # noqa: BLE001 - no peer access between the GPUs of this box: the NCCL path stands
        variants["peer_stores"] = {"api": "orb_hamming_knn_peer", "unavailable": str(e)[:200]}
    return best + (name, variants)


def run_b200(args):
    import torch
    import orb_slam2_detailed_comments_b200 as pkg
    if not torch.cuda.is_available() or pkg.lib().orb_device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU arm)")
    rank, world, local = _dist_setup(args)
    label, W, H, nf, B = WORKLOADS[args.workload]
    if args.batch:
        B = args.batch
    frames = synth.synth_batch(W, H, B, seed0=1234 + 100000 * rank, distinct=min(64, B) if args.workload != "4k" else min(8, B))
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, device=local)
    frames_dev = torch.from_numpy(frames).cuda()

    sampler = ClockSampler(local) if rank == 0 else None
    ms, launches, stage_ms, passes, nkp = bench_extract_resident(ex, frames_dev, args.steps, args.warmup, world)
    clocks = sampler.stop() if sampler else {}
    value = B * world * args.steps / (ms * 1e-3)

    e_steps = max(1, min(args.steps, 5))
    e_ms, h2d, d2h, _ = bench_extract_e2e(ex, frames, e_steps, args.warmup, world)
    e2e_value = B * world * e_steps / (e_ms * 1e-3)

    st_ms, st_matches = bench_stereo_e2e(ex, frames, e_steps, args.warmup, world)
    knn_rate, knn_ms, knn_launches, knn_par, knn_path, knn_variants = bench_knn(rank, world, max(args.steps, 5), args.warmup)
    g_steps = max(args.steps, 5)
    g_ms, g_matches, g_launches = bench_guided(W, H, g_steps, args.warmup, world)
    bow_voc, bow_base = _bow_inputs()
    b_ms, b_words = bench_bow(g_steps, args.warmup, world, bow_voc, bow_base)
    popc_peak = pkg.lib().orb_measure_popc_peak(local) if rank == 0 else 0.0
    latency = None
    if world == 1:
        ex1 = pkg.ORBextractor(nf, 1.2, 8, 20, 7, device=local)   # its own handle: single-image workspaces
        latency = bench_latency(ex1, frames)
        ex1.close()
        latency["cpp_dropin"] = bench_cpp_dropin_latency(frames[0], nf)

    sustained = bench_sustained(ex, frames_dev, B, world, local, rank)
    others = {}
    if args.workload == "kitti" and not args.batch and not args.only_headline:
        del frames_dev
        torch.cuda.empty_cache()
        for name in ("tum", "euroc", "4k"):
            others[name] = bench_other_config(name, rank, world, local)

    if rank != 0:
        return
    peak, peak_src = _peaks()
    A = geometry.algorithmic_bytes(W, H, nf)
    sb = geometry.stage_bytes(W, H, nf)
    per_rank_rate = value / world
    total_stage = sum(stage_ms.values()) or 1.0
    dom = max(stage_ms, key=stage_ms.get)
    n_launch_dom = passes * (8 if dom == "pyramid" else 1)
    frames_per_launch = B * args.steps / max(passes, 1)
    dom_ms = stage_ms[dom] / max(passes, 1)                       # per internal pass (pyramid: its 8 launches together)
    dom_bytes = sb[dom] * frames_per_launch
    dom_achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    traffic = None   # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed ncu --set full capture
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tj = json.load(f)
        if args.workload == "kitti":
            t = tj["dram_bytes_per_launch"][dom]
            traffic = (t["read"] + t["write"]) / tj["frames_per_launch"] * frames_per_launch
    except Exception:
        traffic = None
    issue = None   # instruction-issue view of the same step (the extraction kernels are issue bound)
    try:
        if args.workload == "kitti" and clocks.get("sm_mhz"):
            wi = sum(v for v in tj["warp_instructions_per_launch"].values() if isinstance(v, int)) / tj["frames_per_launch"]
            peak_ips = 148 * 4 * clocks["sm_mhz"] * 1e6   # one warp instruction per cycle per SM sub-partition
            issue = {"warp_instructions_per_image": wi, "achieved": per_rank_rate * wi, "peak": peak_ips, "unit": "warp instr/s",
                     "frac": per_rank_rate * wi / peak_ips, "source": "profiles/ncu_traffic.json (smsp__inst_executed.sum of the committed ncu capture)"}
    except Exception:
        issue = None
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": dom_achieved, "peak": peak, "unit": "GB/s",
        "frac": dom_achieved / peak, "traffic": traffic, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": dom_bytes, "avg_launch_ms": dom_ms, "launches": n_launch_dom,
        "share_of_step": stage_ms[dom] / total_stage,
        "pipeline": {"algorithmic_bytes_per_image": A, "achieved": per_rank_rate * A / 1e9, "frac": per_rank_rate * A / 1e9 / peak},
        "stages_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
        "issue": issue,
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": shared_config(args.workload),
        "run": {"images_per_step_per_gpu": B, "keypoints_per_image": nkp / B, "parallelism": f"frames sharded over {world} GPU(s), no collective",
                "timing": "CUDA events on the extractor's stream, max over ranks"},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e_steps, "ms_per_step": e_ms / e_steps, "api": "orb_extract_batch (pinned host in/out)"},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "stereo": {"metric": "stereo frame pairs/sec (extract L+R + Frame::ComputeStereoMatches on device, host to host)",
                   "value": (B // 2) * world * e_steps / (st_ms * 1e-3), "unit": "pairs/s", "ms_per_step": st_ms / e_steps,
                   "stereo_matches_per_pair": st_matches / (B // 2), "api": "orb_extract_stereo_batch (pinned host in/out)"},
        "guided": {"metric": "guided searches/sec (ORBmatcher::SearchByProjection frame-to-frame form: grid + window scan + claims + rotation filter)",
                   "value": GUIDED_FRAMES * world * g_steps / (g_ms * 1e-3), "unit": "frame pairs/s", "ms_per_step": g_ms / g_steps,
                   "config": f"{GUIDED_FRAMES} frame pairs per step per GPU, {GUIDED_M} projected map points x {GUIDED_N} keypoints each, resident in HBM",
                   "matches_per_pair": g_matches, "gpu_launches": int(g_launches), "api": "orb_search_by_projection (device memory)"},
        "bow": {"metric": "ComputeBoW frames/sec (DBoW2 transform: tree walk + BowVector + FeatureVector, levelsup 4)",
                "value": BOW_FRAMES * world * g_steps / (b_ms * 1e-3), "unit": "frames/s", "ms_per_step": b_ms / g_steps,
                "config": f"{BOW_FRAMES} frames x {BOW_N} descriptors per step per GPU, vocabulary k=10 L=6 (10^6 words), resident in HBM",
                "words_per_frame": b_words, "gpu_launches": 2 * g_steps, "api": "orb_bow_transform (device memory)"},
        "knn": {"metric": "Hamming matches/sec (k=2 brute force)", "value": knn_rate, "unit": "pairs/s",
                "config": f"{KNN_Q} queries x {KNN_D} db rows sharded over {world} GPU(s)" + (", " + knn_path if knn_path else ""),
                "variants": knn_variants,
                "ms_per_step": knn_ms, "gpu_launches": int(knn_launches),
                "parity": knn_par if world > 1 else "n/a at 1 GPU (no merge; tests/test_gpu_knn.py checks the scan)",
                "roofline": {"bound": "popc", "achieved": knn_rate * 8 / world, "peak": popc_peak, "unit": "popc.b32/s",
                             "frac": (knn_rate * 8 / world / popc_peak) if popc_peak else None}},
    }
    line["sustained"] = sustained
    if others:
        line["configs"] = others
    if _NUMA:
        line["e2e"]["numa"] = _NUMA   # this rank's binding (rank 0); every rank binds to its own GPU's node
    if world == 1:
        line["latency"] = latency
    if world == 1 and not args.no_cpu:
        cores = _host_cores()
        n = 512   # ~20-30 core-seconds of CPU work
        v, dt, _ = cpu_extract_rate(W, H, nf, n, cores)
        kv, kdt = cpu_knn_rate(cores)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{n} images of the same workload, one oracle extractor per thread, {dt:.1f} s wall",
                                "knn_pairs_per_s": kv}
        if os.access(REF_BIN, os.X_OK):   # the reference's own ORBextractor.cc (oracle/_ref), same frames, all host cores
            rframes = synth.synth_batch(W, H, max(2 * cores, 16), seed0=4321, distinct=8)
            rimg, rdt = _ref_binary_rate(rframes, nf, cores, 2)
            line["cpu_baseline"]["reference_binary"] = {
                "value": rimg / rdt, "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"{rimg} images, one reference ORBextractor (src/ORBextractor.cc over oracle/cvshim) per thread, {rdt:.1f} s wall"}
        bwv, bwdt = cpu_bow_rate(bow_voc, bow_base, cores)
        line["bow"]["cpu_baseline"] = {"value": bwv, "unit": "frames/s", "cores": cores, "kind": "port",
                                       "sample": f"4096 frames of the same workload, {bwdt:.1f} s wall"}
        gv, gdt = cpu_guided_rate(W, H, cores)
        line["guided"]["cpu_baseline"] = {"value": gv, "unit": "frame pairs/s", "cores": cores, "kind": "port",
                                          "sample": f"4096 frame pairs of the same workload, {gdt:.1f} s wall"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="kitti", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="images per step per GPU (default: per workload)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--only-headline", action="store_true", help="skip the `configs` entries (TUM / EuRoC / 4K)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
