#!/bin/bash
# round 2, step l: upstream-pyramid switch - full GPU suite + headline bench (regression check)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2l_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2l_tests.log
tail -15 gpurun_out/r2l_tests.log
timeout 600 python bench.py --only-headline > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; tail -c 1500 gpurun_out/r2l_bench.json
