"""Runs only the guided-search leg of bench.py (512 frame pairs, 2000 x 2000) — for ncu captures of the three kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
ms, matches, launches = bench.bench_guided(1241, 376, int(sys.argv[1]) if len(sys.argv) > 1 else 5, 2, 1)
print("ms per step", ms / (int(sys.argv[1]) if len(sys.argv) > 1 else 5), "matches per pair", matches, "launches", launches)
