#!/bin/bash
set -x
mkdir -p gpurun_out
ORB_B200_BLUR_TH=64 timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_extract_scale.py -m gpu -x -q > gpurun_out/r2b2_tests.log 2>&1; tail -3 gpurun_out/r2b2_tests.log
for rep in 1 2; do
  ORB_B200_BLUR_TH=64 timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2b2_64_$rep.json 2> gpurun_out/r2b2_64_$rep.err
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2b2_32_$rep.json 2> gpurun_out/r2b2_32_$rep.err
done
for wl in tum 4k; do
  ORB_B200_BLUR_TH=64 timeout 300 python bench.py --workload $wl --only-headline --no-cpu > gpurun_out/r2b2_64_$wl.json 2>/dev/null
  timeout 300 python bench.py --workload $wl --only-headline --no-cpu > gpurun_out/r2b2_32_$wl.json 2>/dev/null
done
python - <<'PY'
import json
for n in ("64_1","32_1","64_2","32_2","64_tum","32_tum","64_4k","32_4k"):
    try:
        d=json.loads(open(f"gpurun_out/r2b2_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(n, "failed", e)
PY
