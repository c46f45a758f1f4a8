#!/bin/bash
# streams x chunk on the other shapes (pyramid on the high-priority stream)
set -x
mkdir -p gpurun_out
: > gpurun_out/r2u_sweep.txt
for wl in tum euroc 4k; do
for cfg in "2 0" "4 0" "4 32" "4 48" "3 64" "4 96" "4 16" "4 8"; do
  set -- $cfg
  if [ $wl = 4k ] && [ $2 -gt 16 ]; then continue; fi
  export ORB_B200_STREAMS=$1
  if [ $2 = 0 ]; then unset ORB_B200_CHUNK; else export ORB_B200_CHUNK=$2; fi
  timeout 300 python bench.py --workload $wl --only-headline --no-cpu > gpurun_out/r2u_tmp.json 2> gpurun_out/r2u_tmp.err
  python - "$wl $cfg" <<'PY' >> gpurun_out/r2u_sweep.txt
import json,sys
try:
    d=json.loads(open("gpurun_out/r2u_tmp.json").read().strip().splitlines()[-1])
    print("workload streams chunk =", sys.argv[1], "->", round(d["value"]), round(d["e2e"]["value"]))
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
done
cat gpurun_out/r2u_sweep.txt
