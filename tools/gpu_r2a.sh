#!/bin/bash
# round 2, call A: parity of the rewritten FAST kernel (TMA window) + A/B against the round-1 kernel + ncu.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/r2a_smi.txt 2>&1
timeout 300 python tests/fast_debug.py > $O/r2a_fast_debug.log 2>&1; echo "fast_debug exit $?"
ORB_B200_NO_TMA=1 timeout 300 python tests/fast_debug.py small tum > $O/r2a_fast_debug_notma.log 2>&1; echo "fast_debug(no tma) exit $?"
timeout 600 python -m pytest tests -m gpu -q -x > $O/r2a_pytest.log 2>&1; echo "pytest exit $?"
tail -3 $O/r2a_pytest.log
for v in new v1 notma; do
  case $v in new) E="";; v1) E="ORB_B200_FAST_V1=1";; notma) E="ORB_B200_NO_TMA=1";; esac
  env $E timeout 300 python bench.py --no-cpu --steps 10 --warmup 3 > $O/r2a_bench_$v.json 2> $O/r2a_bench_$v.err; echo "bench $v exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/r2a_bench_$v.json").read().strip().splitlines()[-1])
    print("$v", round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["stages_ms_per_step"])
except Exception as e: print("$v parse failed", e)
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -s 13 -c 40 --csv --log-file $O/r2a_launches.csv python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu > $O/r2a_ncu_l.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_(fast|blur|describe)" -s 3 -c 3 -o $O/r2a_full -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu > $O/r2a_ncu_f.log 2>&1; echo "ncu full exit $?"
