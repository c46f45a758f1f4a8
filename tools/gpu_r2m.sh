#!/bin/bash
# round 2, step m: FAST kernel A/B on one box: previous build (liborb_b200_base.so) vs this build; quick parity first
set -x
mkdir -p gpurun_out
P=orb_slam2_detailed_comments_b200
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_extract_scale.py tests/test_reference_binary.py -m gpu -x -q > gpurun_out/r2m_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2m_tests.log
tail -5 gpurun_out/r2m_tests.log
for rep in 1 2; do
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2m_new_$rep.json 2> gpurun_out/r2m_new_$rep.err
  cp $P/liborb_b200.so /tmp/new.so; cp $P/liborb_b200_base.so $P/liborb_b200.so
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2m_base_$rep.json 2> gpurun_out/r2m_base_$rep.err
  cp /tmp/new.so $P/liborb_b200.so
done
python - <<'PY'
import json
for n in ("new_1","base_1","new_2","base_2"):
    try:
        d=json.loads(open(f"gpurun_out/r2m_{n}.json").read().strip().splitlines()[-1])
        print(n, round(d["value"]), round(d["e2e"]["value"]), d.get("stages_ms_per_step"))
    except Exception as e: print(n, "failed", e)
PY
