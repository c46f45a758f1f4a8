#!/bin/bash
# round 2, call I: blur with TMA tile staging (parity, A/B), capacity-overflow reporting test, 4K pass-size check
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "extract or reference or dropin or frontend or stereo or zz or dense" > $O/r2i_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/r2i_pytest.log
for v in new blur_notma; do
  case $v in new) E="A=1";; blur_notma) E="ORB_B200_BLUR_NO_TMA=1";; esac
  for wl in kitti tum; do
  env $E python bench.py --workload $wl --no-cpu --only-headline --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v $wl', round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline']['stages_ms_per_step'].items()}, flush=True)"
  done
done
for c in 0 8 4; do
  if [ $c = 0 ]; then unset ORB_B200_CHUNK; else export ORB_B200_CHUNK=$c; fi
  python bench.py --workload 4k --no-cpu --only-headline --steps 8 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('4k chunk=$c', round(d['value']), round(d['e2e']['value']), flush=True)"
done
unset ORB_B200_CHUNK
python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > $O/r2i_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_blur" -s 1 -c 1 -o $O/r2i_blur -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu --only-headline > $O/r2i_ncu.log 2>&1; echo "ncu exit $?"
