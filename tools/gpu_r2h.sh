#!/bin/bash
# round 2, call H: quadtree with pipelined key passes + aggregated atomics: parity, then 4K / KITTI A/B against the old CTA sizes
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "extract or reference or dropin or frontend or stereo or zz" > $O/r2h_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/r2h_pytest.log
for wl in 4k kitti tum; do
  for qt in 0 512 1024 192; do
    if [ $qt = 0 ]; then unset ORB_B200_QT_THREADS; else export ORB_B200_QT_THREADS=$qt; fi
    python bench.py --workload $wl --no-cpu --only-headline --steps 8 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl qt=$qt', round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline']['stages_ms_per_step'].items()}, flush=True)"
  done
done
