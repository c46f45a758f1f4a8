#!/bin/bash
# round 2, final capture ZZ (last tree of the round): ncu --set full of one 32-frame KITTI pass (13 launches) + launch list of one bench step (each after the same
# command exited 0 without ncu)
set -x
O=gpurun_out; mkdir -p $O
python bench.py --steps 1 --warmup 1 --batch 32 --no-cpu --only-headline > $O/r2zz_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"orb_(fast|blur|describe|quadtree|pyr)" -s 13 -c 13 -o $O/r2zz_full -f python bench.py --steps 1 --warmup 1 --batch 32 --no-cpu --only-headline > $O/r2zz_ncu_f.log 2>&1; echo "ncu full exit $?"
python bench.py --steps 1 --warmup 1 --no-cpu --only-headline > $O/r2zz_plain2.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"orb_" -c 1200 --csv --log-file $O/r2zz_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu --only-headline > $O/r2zz_ncu_l.log 2>&1; echo "ncu list exit $?"
wc -l $O/r2zz_launches.csv
