#!/bin/bash
O=gpurun_out; mkdir -p $O
( for c in "256 38 0 1 32 0" "256 38 0 1 37 0" "256 38 0 3 32 16" "256 38 0 3 32 0" "32 31 0 1 32 0" "32 31 0 1 37 0" "32 31 0 3 32 16" "48 37 0 1 48 0" "48 37 0 1 36 0" "48 37 0 1 33 0" "32 31 0 4 16 4"; do tests/cuda/tma_probe $c; done ) > $O/r2c_probe.log 2>&1
cat $O/r2c_probe.log
