#!/bin/bash
set -x
mkdir -p gpurun_out
ORB_B200_DESC_WIDE=1 timeout 600 python -m pytest tests/test_gpu_extract.py -m gpu -x -q > gpurun_out/r2p_tests.log 2>&1; tail -3 gpurun_out/r2p_tests.log
for rep in 1 2; do
  ORB_B200_DESC_WIDE=1 timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2p_wide_$rep.json 2> gpurun_out/r2p_wide_$rep.err
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2p_def_$rep.json 2> gpurun_out/r2p_def_$rep.err
done
python - <<'PY'
import json
for n in ("wide_1","def_1","wide_2","def_2"):
    try:
        d=json.loads(open(f"gpurun_out/r2p_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(n, "failed", e)
PY
