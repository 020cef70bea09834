#!/bin/bash
mkdir -p gpurun_out
BQ="python bench.py --trait quantitative --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $BQ 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('quant step %.3f ms' % d['ms_per_step'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:dosage_moments -c 12 --csv --log-file gpurun_out/k3m_time.csv $BQ > /dev/null 2>&1
grep -v "^==" gpurun_out/k3m_time.csv | python -c "
import csv,sys
r=list(csv.reader(sys.stdin))
for x in r[1:]: print(x[4][:40], x[-1])"
