#!/bin/bash
# round 2, call K: full GPU suite; the measured alternatives (tiled resize, blur / descriptor TMA, round-1 FAST) must stay bit-exact
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/r2k_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/r2k_pytest.log
for E in ORB_B200_PYR_TILED=1 ORB_B200_BLUR_TMA=1 ORB_B200_DESC_TMA=1 ORB_B200_FAST_V1=1 ORB_B200_NO_TMA=1; do
  env $E timeout 900 python -m pytest tests -m gpu -q -x -k "extract or reference_binary or frontend or stereo" > $O/r2k_pytest_alt.log 2>&1; echo "$E exit $?"; tail -1 $O/r2k_pytest_alt.log
done
ORB_B200_PYR_TILED=1 python bench.py --no-cpu --only-headline --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('tiled(all levels) kitti', round(d['value']), {k: round(v,3) for k,v in d['roofline']['stages_ms_per_step'].items()}, flush=True)"
python -c "import __graft_entry__ as g; g.smoke()"
