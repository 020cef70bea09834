#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
BQ="python bench.py --trait quantitative --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
$BQ 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('quant step %.3f ms  off-fast %s' % (d['ms_per_step'], r['snarls_off_fast_path']))"
K='regex:dosage_moments|ols_solve|quant_general'
$BQ > gpurun_out/plain_q.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 12 --csv --log-file gpurun_out/launches_quant.csv $BQ > gpurun_out/ncu_q.log 2>&1
grep -v "^==" gpurun_out/launches_quant.csv | python -c "
import csv,sys
r=list(csv.reader(sys.stdin))
for x in r[1:]: print(x[4][:34], x[7], x[8], x[-1])"
