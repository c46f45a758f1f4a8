#!/bin/bash
# host-to-host throughput of mid-size calls against the pass size (how many passes a call needs to overlap its copies)
set -x
mkdir -p gpurun_out
run() { # name, workload, batch, env
  local name=$1 wl=$2 b=$3; shift 3
  env "$@" timeout 300 python bench.py --workload $wl --batch $b --only-headline --no-cpu > gpurun_out/r2v_$name.json 2> gpurun_out/r2v_$name.err
}
for c in 2 3 4; do run 4k_b16_c$c 4k 16 ORB_B200_CHUNK=$c; done
for c in 3 6; do run 4k_b64_c$c 4k 64 ORB_B200_CHUNK=$c; done
for c in 8 16 32; do run kitti_b64_c$c kitti 64 ORB_B200_CHUNK=$c; done
for c in 8 16 32; do run kitti_b128_c$c kitti 128 ORB_B200_CHUNK=$c; done
for c in 16 32; do run kitti_b256_c$c kitti 256 ORB_B200_CHUNK=$c; done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2v_*_c*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"]), round(d["e2e"]["value"]))
    except Exception as e: print(f, "failed", e)
PY
