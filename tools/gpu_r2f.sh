#!/bin/bash
# round 2, call F (2 GPUs): NCCL-sharded kNN through the C ABI (2-rank test), 2-GPU bench line with knn.parity
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_knn_nccl.py -q -x > $O/r2f_pytest_nccl.log 2>&1; echo "nccl test exit $?"; tail -5 $O/r2f_pytest_nccl.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --only-headline > $O/r2f_bench_n2.json 2> $O/r2f_bench_n2.err; echo "bench n2 exit $?"
python - <<'PY'
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2f_bench_n2.json").read().splitlines() if l.startswith("{")][-1])
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "sustained", round(d["sustained"]["value"]))
    print("knn", d["knn"]["value"], d["knn"]["roofline"]["frac"], "parity", d["knn"]["parity"], d["knn"]["config"])
except Exception as e: print("parse failed", e)
PY
tail -5 $O/r2f_bench_n2.err
