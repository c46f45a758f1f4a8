#!/bin/bash
set -x
mkdir -p gpurun_out
for h in 0 34 30; do
  if [ $h = 0 ]; then unset ORB_B200_FAST_HACK; else export ORB_B200_FAST_HACK=$h; fi
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2o_$h.json 2> gpurun_out/r2o_$h.err
done
python - <<'PY'
import json
for n in ("0","34","30"):
    try:
        d=json.loads(open(f"gpurun_out/r2o_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["value"]), round(d["e2e"]["value"]), d["roofline"].get("achieved"))
    except Exception as e: print(n, "failed", e)
PY
