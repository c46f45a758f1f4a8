#!/bin/bash
# per-call pass size (ensure_shape): parity of the touched paths + the mid-size host-to-host points
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_extract_scale.py tests/test_gpu_stereo.py tests/test_gpu_dropin.py -m gpu -x -q > gpurun_out/r2u3_tests.log 2>&1; tail -3 gpurun_out/r2u3_tests.log
run() { local name=$1 wl=$2 b=$3; timeout 300 python bench.py --workload $wl --batch $b --only-headline --no-cpu > gpurun_out/r2u3_$name.json 2> gpurun_out/r2u3_$name.err; }
run 4k_b16 4k 16
run kitti_b64 kitti 64
run kitti_b128 kitti 128
run tum_b96 tum 96
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2u3_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(f, "failed", e)
PY
