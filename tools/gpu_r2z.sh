#!/bin/bash
# 2 GPUs: the peer-store kNN test (+ the NCCL one)
set -x
mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -6
timeout 900 python -m pytest tests/test_gpu_knn_nccl.py -m gpu -x -q > gpurun_out/r2z_tests.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/r2z_tests.log
