#!/bin/bash
set -x
mkdir -p gpurun_out
ORB_B200_PYR_SIDE=1 timeout 600 python -m pytest tests/test_gpu_extract.py tests/test_gpu_threads.py -m gpu -x -q > gpurun_out/r2t_tests.log 2>&1; tail -3 gpurun_out/r2t_tests.log
for rep in 1 2; do
  ORB_B200_PYR_SIDE=1 timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2t_new_$rep.json 2> gpurun_out/r2t_new_$rep.err
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2t_old_$rep.json 2> gpurun_out/r2t_old_$rep.err
done
ORB_B200_PYR_SIDE=1 ORB_B200_STREAMS=3 timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2t_new_3.json 2> gpurun_out/r2t_new_3.err
python - <<'PY'
import json
for n in ("new_1","old_1","new_2","old_2","new_3"):
    try:
        d=json.loads(open(f"gpurun_out/r2t_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(n, "failed", e)
PY
