#!/bin/bash
# 8 GPUs: kNN tests (1 GPU + 2-rank) and the 8-GPU bench after the chunk-balance change
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_knn.py tests/test_gpu_knn_nccl.py tests/test_zz_no_alloc.py tests/test_reference_matcher.py -m gpu -q > $O/r2z3_tests.log 2>&1; tail -2 $O/r2z3_tests.log
bash tools/gpu_r2z2.sh 8
timeout 300 python bench.py --only-headline --no-cpu > $O/r2z3_n1.json 2>/dev/null; python -c "
import json; d=json.loads(open('$O/r2z3_n1.json').read().strip().splitlines()[-1]); print('N=1 knn', d['knn']['value']/1e9, d['knn']['roofline']['frac'])"
