#!/bin/bash
# round 2, final multi-GPU run: bench at N = 2, 4, 8 on one 8-GPU box (+ the 2-rank NCCL kNN test and the new reference pins on GPU)
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_reference_matcher.py tests/test_gpu_knn_nccl.py -m gpu -q > $O/r2y_tests.log 2>&1; tail -3 $O/r2y_tests.log
for n in 2 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 10 --warmup 3 --only-headline > $O/r2y_bench_n$n.json 2> $O/r2y_bench_n$n.err; echo "bench n$n exit $?"
python - <<PY
import json
try:
    d=json.loads([l for l in open("$O/r2y_bench_n$n.json").read().splitlines() if l.startswith("{")][-1])
    print("N=$n value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "knn", round(d["knn"]["value"]/1e9), d["knn"]["roofline"]["frac"], d["knn"]["parity"])
except Exception as e: print("parse failed", e)
PY
done
