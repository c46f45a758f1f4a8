#!/bin/bash
# round 2, step v: full GPU suite + the default bench line (all configs) with the new defaults
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2v_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2v_tests.log
tail -5 gpurun_out/r2v_tests.log
timeout 900 python bench.py > gpurun_out/r2v_bench.json 2> gpurun_out/r2v_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r2v_bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["pipeline"]["frac"],4), "fast frac", round(d["roofline"]["frac"],4), "launches", d.get("gpu_launches"))
    print("sustained", round(d["sustained"]["value"]), d["sustained"]["clocks"])
    for k,v in d.get("configs",{}).items(): print(k, round(v["value"]), round(v["e2e"]["value"]), round(v["pipeline"]["frac"],4), v["stages_ms_per_step"])
    print("latency", d.get("latency",{}).get("cpp_dropin"))
    print("knn", d["knn"]["value"], d["knn"]["roofline"]["frac"], d["knn"]["parity"])
    print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"].get("reference_binary",{}).get("value"))
except Exception as e: print("parse failed", e)
PY
