"""Pinned host<->device copy bandwidth of the box: the ceiling of bench.py's e2e number.

  python tools/pcie_bw.py                 one process, GPU 0: H2D and D2H alone
  python tools/pcie_bw.py --ranks N       N concurrent processes, process r on GPU r (bound to that GPU's NUMA node like bench.py's ranks), each
                                          doing ONLY the copies of one bench step - 239 MB H2D + 65 MB D2H, both directions in flight at once on two
                                          streams - for --seconds; prints per-rank and aggregate GB/s as one JSON line.
The aggregate at N ranks, divided by the bytes one image needs (466,616 in + 127,716 out at KITTI), is the most images/s ANY host-to-host path
through these PCIe links / this host memory system can reach; bench.py's e2e at N GPUs is to be read against it (profiles/r2_pcie_ranks.md)."""
import argparse
import json
import os
import subprocess
import sys
import time

H2D_BYTES = 512 * 1241 * 376            # frames of one KITTI step per GPU
D2H_BYTES = 512 * 2128 * 60 + 512 * 36  # keypoints + descriptors (cap rows) + counts


def _bind(local):
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return node
    except Exception:   # noqa: BLE001
        return None


def worker(rank, seconds, start_at, bind):
    import torch
    torch.cuda.set_device(rank)
    node = _bind(rank) if bind else None
    h_in = torch.empty(H2D_BYTES, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(D2H_BYTES, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(H2D_BYTES, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(D2H_BYTES, dtype=torch.uint8, device="cuda")
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(3):
        with torch.cuda.stream(s_in):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s_out):
            h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    while time.time() < start_at:   # all ranks start together
        time.sleep(0.001)
    t0 = time.perf_counter()
    steps = 0
    while time.perf_counter() - t0 < seconds:
        for _ in range(4):
            with torch.cuda.stream(s_in):
                d_in.copy_(h_in, non_blocking=True)
            with torch.cuda.stream(s_out):
                h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        steps += 4
    dt = time.perf_counter() - t0
    print(json.dumps({"rank": rank, "numa": node, "steps": steps, "h2d_gbs": steps * H2D_BYTES / dt / 1e9, "d2h_gbs": steps * D2H_BYTES / dt / 1e9,
                      "steps_per_s": steps / dt}), flush=True)


def single():
    import torch
    n = H2D_BYTES
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    out = {}
    for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        e1.synchronize()
        out[name + "_alone_gbs"] = 10 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=0)
    ap.add_argument("--seconds", type=float, default=3.0)
    ap.add_argument("--no-bind", action="store_true")
    ap.add_argument("--worker", type=int, default=-1)
    ap.add_argument("--start-at", type=float, default=0.0)
    a = ap.parse_args()
    if a.worker >= 0:
        worker(a.worker, a.seconds, a.start_at, not a.no_bind)
        return
    if a.ranks <= 0:
        single()
        return
    start_at = time.time() + 12.0   # CUDA context + pinned allocation of every rank fit in here
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", str(r), "--seconds", str(a.seconds), "--start-at", str(start_at)]
                              + (["--no-bind"] if a.no_bind else []), stdout=subprocess.PIPE, text=True) for r in range(a.ranks)]
    rows = []
    for p in procs:
        out, _ = p.communicate(timeout=300)
        rows += [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    per_image = 1241 * 376 + 2128 * 60
    agg_h2d, agg_d2h = sum(r["h2d_gbs"] for r in rows), sum(r["d2h_gbs"] for r in rows)
    print(json.dumps({"ranks": a.ranks, "bound_to_numa": not a.no_bind, "aggregate_h2d_gbs": agg_h2d, "aggregate_d2h_gbs": agg_d2h,
                      "aggregate_steps_per_s": sum(r["steps_per_s"] for r in rows),
                      "copy_ceiling_images_per_s": sum(r["steps_per_s"] for r in rows) * 512,
                      "bytes_per_image": per_image, "per_rank": rows}), flush=True)


if __name__ == "__main__":
    main()
