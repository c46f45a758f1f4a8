#!/bin/bash
# (1) the seeded configuration sweep against the reference's own code on the GPU; (2) 4K pass-size / blur-tile sweep on one box
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_reference_binary.py -m gpu -x -q > gpurun_out/r2u_tests.log 2>&1; tail -3 gpurun_out/r2u_tests.log
run() { # name, env...
  local name=$1; shift
  env "$@" timeout 300 python bench.py --workload 4k --only-headline --no-cpu > gpurun_out/r2u_$name.json 2> gpurun_out/r2u_$name.err
}
for rep in 1 2; do
  run def_$rep X=1
  run c3_$rep ORB_B200_CHUNK=3
  run c8_$rep ORB_B200_CHUNK=8
  run c16_$rep ORB_B200_CHUNK=16
  run b128_$rep ORB_B200_BLUR_TH=128
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2u_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(f, "failed", e)
PY
