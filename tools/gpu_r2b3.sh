#!/bin/bash
mkdir -p gpurun_out
ORB_B200_BLUR_TH=128 timeout 900 python -m pytest tests/test_gpu_extract.py -m gpu -x -q 2>&1 | tail -2
for th in 32 64 128 64 128; do
  for wl in kitti tum; do
  ORB_B200_BLUR_TH=$th timeout 300 python bench.py --workload $wl --only-headline --no-cpu > gpurun_out/r2b3.json 2>/dev/null
  python -c "
import json; d=json.loads(open('gpurun_out/r2b3.json').read().strip().splitlines()[-1]); print('$th $wl', round(d['value']), round(d['e2e']['value']))"
  done
done
