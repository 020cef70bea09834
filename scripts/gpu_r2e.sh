#!/bin/bash
# Round 2, GPU call E: logistic parity, the wide tests with K3M3 / K3M2x at 50,000 haplotypes, C5 bench line again.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_logit.py tests/test_gpu_wide.py tests/test_gpu_quant.py -x -q 2>&1 | tail -25 > $OUT/pytest_e.log; cat $OUT/pytest_e.log
timeout 1500 python bench.py --config c5 --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_c5.json 2> $OUT/bench_c5.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c5.json'))
q=d['traits']['quantitative']
print('binary ms', d['ms_per_step'], 'quant ms', q['ms_per_step'], q['roofline'].get('snarls_off_fast_path'), q['roofline'].get('other_kernels_ms'), q['parity'])
PY
tail -3 $OUT/bench_c5.err
