#!/bin/bash
# A/B of one env switch: $1 = variable that selects the OLD path
set -x
mkdir -p gpurun_out
V=$1
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_extract_scale.py tests/test_reference_binary.py tests/test_gpu_stereo.py -m gpu -x -q > gpurun_out/r2r_tests.log 2>&1; tail -3 gpurun_out/r2r_tests.log
for rep in 1 2; do
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2r_new_$rep.json 2> gpurun_out/r2r_new_$rep.err
  env $V=1 timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2r_old_$rep.json 2> gpurun_out/r2r_old_$rep.err
done
python - <<'PY'
import json
for n in ("new_1","old_1","new_2","old_2"):
    try:
        d=json.loads(open(f"gpurun_out/r2r_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(n, "failed", e)
PY
