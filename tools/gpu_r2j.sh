#!/bin/bash
# round 2, call J: tiled separable resize (persistent CTAs, 2-stage TMA): parity, A/B against the 4-pixel window kernel, ncu
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "extract or reference or frontend or stereo or zz" > $O/r2j_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/r2j_pytest.log
for v in tiled untiled; do
  case $v in tiled) E="A=1";; untiled) E="ORB_B200_PYR_UNTILED=1";; esac
  for wl in kitti tum 4k; do
  env $E python bench.py --workload $wl --no-cpu --only-headline --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v $wl', round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline']['stages_ms_per_step'].items()}, flush=True)"
  done
done
python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > $O/r2j_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_pyr" -s 9 -c 9 -o $O/r2j_pyr -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > $O/r2j_ncu.log 2>&1; echo "ncu exit $?"
