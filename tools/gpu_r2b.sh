#!/bin/bash
# round 2, call B: TMA descriptor kernel parity + A/B against the cp.async kernel + ncu.
set -u
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x > $O/r2b_pytest.log 2>&1; echo "pytest exit $?"
tail -5 $O/r2b_pytest.log
for v in new desctma; do
  case $v in new) E="A=1";; desctma) E="ORB_B200_DESC_TMA=1";; esac
  env $E timeout 300 python bench.py --no-cpu --steps 10 --warmup 3 > $O/r2b_bench_$v.json 2> $O/r2b_bench_$v.err; echo "bench $v exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("$O/r2b_bench_$v.json").read().strip().splitlines()[-1])
    print("$v", round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["stages_ms_per_step"])
except Exception as e: print("$v parse failed", e)
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_(describe)" -s 1 -c 1 -o $O/r2b_full -f python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu > $O/r2b_ncu_f.log 2>&1; echo "ncu full exit $?"
