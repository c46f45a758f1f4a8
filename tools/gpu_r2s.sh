#!/bin/bash
# sweep of the scheduling switches with the round-2 kernels: streams x side x chunk
set -x
mkdir -p gpurun_out
: > gpurun_out/r2s_sweep.txt
for cfg in "4 2 32" "4 2 24" "4 2 16" "4 2 40" "3 2 32" "3 2 40"; do
  set -- $cfg
  export ORB_B200_STREAMS=$1 ORB_B200_SIDE=$2
  if [ $3 = 0 ]; then unset ORB_B200_CHUNK; else export ORB_B200_CHUNK=$3; fi
  for rep in 1 2; do
  timeout 300 python bench.py --only-headline --no-cpu > gpurun_out/r2s_tmp.json 2> gpurun_out/r2s_tmp.err
  python - "$cfg" <<'PY' >> gpurun_out/r2s_sweep.txt
import json,sys
try:
    d=json.loads(open("gpurun_out/r2s_tmp.json").read().strip().splitlines()[-1])
    print("streams side chunk =", sys.argv[1], "->", round(d["value"]), round(d["e2e"]["value"]))
except Exception as e: print(sys.argv[1], "failed", e)
PY
  done
done
cat gpurun_out/r2s_sweep.txt
