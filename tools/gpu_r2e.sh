#!/bin/bash
# round 2, call E: full GPU suite after the oracle pins / WsOrder / drop-in rework; default bench line (configs, sustained, cpp latency); ncu full capture
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/r2e_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $O/r2e_pytest.log
t0=$(date +%s)
timeout 900 python bench.py > $O/r2e_bench.json 2> $O/r2e_bench.err; echo "bench exit $? in $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r2e_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["pipeline"]["frac"],4))
    print("sustained", round(d["sustained"]["value"]), d["sustained"]["steps"], d["sustained"]["clocks"])
    for k,v in d.get("configs",{}).items(): print(k, round(v["value"]), round(v["e2e"]["value"]), round(v["pipeline"]["frac"],4), v["stages_ms_per_step"])
    print("latency", d.get("latency"))
    print("knn", d["knn"]["value"], d["knn"]["roofline"]["frac"], d["knn"]["parity"])
except Exception as e: print("parse failed", e)
PY
tail -3 $O/r2e_bench.err
python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > $O/r2e_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_(fast|blur|describe|quadtree|pyr)" -s 13 -c 13 -o $O/r2e_full -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > $O/r2e_ncu_f.log 2>&1; echo "ncu full exit $?"
