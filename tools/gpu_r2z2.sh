#!/bin/bash
# N GPUs: bench with both kNN exchange paths
set -u
O=gpurun_out; mkdir -p $O
n=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2971$n bench.py --gpus $n --steps 10 --warmup 3 --only-headline > $O/r2z_bench_n$n.json 2> $O/r2z_bench_n$n.err; echo "bench n$n exit $?"
tail -3 $O/r2z_bench_n$n.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("$O/r2z_bench_n$n.json").read().splitlines() if l.startswith("{")][-1])
    print("N=$n value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "knn", round(d["knn"]["value"]/1e9), d["knn"]["roofline"]["frac"], d["knn"]["parity"])
    print(d["knn"]["config"])
    print(json.dumps(d["knn"]["variants"], indent=1))
except Exception as e: print("parse failed", e)
PY
