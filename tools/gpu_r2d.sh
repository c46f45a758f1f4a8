#!/bin/bash
# round 2, call D: no-alloc kNN entry point (alone + after the suite), then the A/B lines and ncu captures for profiles/ (re-taken after the machine swap)
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q > $O/r2d_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -2 $O/r2d_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -q -x > $O/r2d_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/r2d_pytest.log
for v in new fastv1 notma desctma; do
  case $v in new) E="A=1";; fastv1) E="ORB_B200_FAST_V1=1";; notma) E="ORB_B200_NO_TMA=1";; desctma) E="ORB_B200_DESC_TMA=1";; esac
  env $E timeout 300 python bench.py --no-cpu --steps 10 --warmup 3 > $O/r2d_bench_$v.json 2> $O/r2d_bench_$v.err; echo "bench $v exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/r2d_bench_$v.json").read().strip().splitlines()[-1])
    print("$v", round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["stages_ms_per_step"])
except Exception as e: print("$v parse failed", e)
PY
done
python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu > $O/r2d_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -s 13 -c 26 --csv --log-file $O/r2d_launches.csv python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu > $O/r2d_ncu_l.log 2>&1; echo "ncu list exit $?"
