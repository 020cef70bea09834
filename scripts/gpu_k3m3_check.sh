#!/bin/bash
# K3M3 (three-path snarls on the integer MMA): quantitative parity tests, then the quantitative bench with and without it
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quant.py tests/test_gpu_shim.py -m gpu -x -q 2>&1 | tail -12
BQ="python bench.py --trait quantitative --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 quant step %.3f ms  value %.4g off-fast %s frac %.3f' % (d['ms_per_step'], d['value'], r.get('snarls_off_fast_path'), r['frac']))"; }
STOAT_GPU_K3M3=0 timeout 300 $BQ 2>&1 | tail -3 | show "K3M3=off"
timeout 300 $BQ 2>&1 | tail -3 | show "K3M3=on"
K='regex:dosage_moments|ols_solve|quant_general|student_tail|lean_rec'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 12 --csv --log-file gpurun_out/launches_quant3.csv $BQ > gpurun_out/ncu_q3.log 2>&1
grep -v "^==" gpurun_out/launches_quant3.csv | python -c "
import csv,sys
r=list(csv.reader(sys.stdin))
for x in r[1:]: print(x[4][:40], x[7], x[8], x[-1])"
