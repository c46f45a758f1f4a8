#!/bin/bash
# round 2, final check W: smoke(), full GPU suite, default bench line, reference arm
set -x
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2w_smoke.log 2>&1; tail -2 gpurun_out/r2w_smoke.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2w_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2w_tests.log
tail -4 gpurun_out/r2w_tests.log
timeout 900 python bench.py > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r2w_ref.json 2> gpurun_out/r2w_ref.err; echo "ref rc=$?"; tail -c 600 gpurun_out/r2w_ref.json
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r2w_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["pipeline"]["frac"],4), "fast frac", round(d["roofline"]["frac"],4), "issue", round(d["roofline"]["issue"]["frac"],3), "launches", d.get("gpu_launches"))
    for k,v in d.get("configs",{}).items(): print(k, round(v["value"]), round(v["e2e"]["value"]), round(v["pipeline"]["frac"],4))
    print("knn", d["knn"]["value"], d["knn"]["roofline"]["frac"])
except Exception as e: print("parse failed", e)
PY
