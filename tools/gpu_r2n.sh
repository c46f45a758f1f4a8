#!/bin/bash
# round 2, step n: ncu --set full of the FAST kernel with source (after the same command exited 0 without ncu)
set -x
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > gpurun_out/r2n_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_fast" -s 1 -c 1 -o gpurun_out/r2n_fast -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > gpurun_out/r2n_ncu.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out/
