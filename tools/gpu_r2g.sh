#!/bin/bash
# round 2, call G (8 GPUs): copy ceiling with N concurrent ranks vs bench e2e at N = 4, 8 on the same box
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L | wc -l; nproc; numactl -H 2>/dev/null | head -4
python tools/pcie_bw.py > $O/r2g_pcie_1alone.json 2>&1; cat $O/r2g_pcie_1alone.json
for n in 1 2 4 8; do timeout 200 python tools/pcie_bw.py --ranks $n --seconds 3 2>/dev/null | tail -1 > $O/r2g_pcie_$n.json; python -c "
import json; d=json.load(open('$O/r2g_pcie_$n.json')); print($n, 'h2d', round(d['aggregate_h2d_gbs'],1), 'd2h', round(d['aggregate_d2h_gbs'],1), 'ceiling img/s', round(d['copy_ceiling_images_per_s']), [r['numa'] for r in d['per_rank']])"; done
timeout 200 python tools/pcie_bw.py --ranks 8 --seconds 3 --no-bind 2>/dev/null | tail -1 > $O/r2g_pcie_8_nobind.json; python -c "
import json; d=json.load(open('$O/r2g_pcie_8_nobind.json')); print('8 nobind h2d', round(d['aggregate_h2d_gbs'],1), 'ceiling', round(d['copy_ceiling_images_per_s']))"
for n in 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 10 --warmup 3 --only-headline > $O/r2g_bench_n$n.json 2> $O/r2g_bench_n$n.err; echo "bench n$n exit $?"
python - <<PY
import json
try:
    d=json.loads([l for l in open("$O/r2g_bench_n$n.json").read().splitlines() if l.startswith("{")][-1])
    print("N=$n value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "knn", round(d["knn"]["value"]/1e9), d["knn"]["roofline"]["frac"], d["knn"]["parity"])
except Exception as e: print("parse failed", e)
PY
done
