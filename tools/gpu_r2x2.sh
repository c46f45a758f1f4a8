#!/bin/bash
# round 2, final state: full GPU suite + default bench + ncu --set full of one 32-frame pass (after the same command exited 0 without ncu)
set -x
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/r2x2_tests.log 2>&1; echo "tests rc=$?" >> $O/r2x2_tests.log; tail -3 $O/r2x2_tests.log
timeout 900 python bench.py > $O/r2x2_bench.json 2> $O/r2x2_bench.err; echo "bench rc=$?"
python bench.py --steps 1 --warmup 1 --batch 32 --no-cpu --only-headline > $O/r2x2_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_(fast|blur|describe|quadtree|pyr)" -s 13 -c 13 -o $O/r2x2_full -f python bench.py --steps 1 --warmup 1 --batch 32 --no-cpu --only-headline > $O/r2x2_ncu_f.log 2>&1; echo "ncu full exit $?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2x2_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["pipeline"]["frac"],4), "fast frac", round(d["roofline"]["frac"],4))
for k,v in d.get("configs",{}).items(): print(k, round(v["value"]), round(v["e2e"]["value"]))
PY
