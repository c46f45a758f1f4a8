#!/bin/bash
# round 2, step q: ncu --set full of all extraction kernels of one 64-frame KITTI pass (after the same command exited 0 without ncu)
set -x
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > gpurun_out/r2q_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_(fast|blur|describe|quadtree|pyr)" -s 13 -c 13 -o gpurun_out/r2q_full -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > gpurun_out/r2q_ncu.log 2>&1; echo "ncu exit $?"
