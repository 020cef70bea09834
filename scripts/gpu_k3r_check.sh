#!/bin/bash
# K3R bring-up: parity tests, quantitative bench under a few configurations, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
BQ="python bench.py --trait quantitative --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 quant step %.3f ms  off-fast %s frac %.3f' % (d['ms_per_step'], r.get('snarls_off_fast_path'), r['frac']))"; }
STOAT_GPU_K3R=0 timeout 300 $BQ 2>&1 | show "K3R=off"
timeout 300 $BQ 2>&1 | show "wpt2"
STOAT_GPU_K3R_WPT=1 timeout 300 $BQ 2>&1 | show "wpt1"
STOAT_GPU_K3R_WPT=4 timeout 300 $BQ 2>&1 | show "wpt4"
STOAT_GPU_K3R_RING=4 timeout 300 $BQ 2>&1 | show "wpt2 ring4"
STOAT_GPU_K3R_CTAS=2 timeout 300 $BQ 2>&1 | show "wpt2 ctas2"
K='regex:dosage_moments|ols_solve|quant_general'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 16 --csv --log-file gpurun_out/launches_quant.csv $BQ > gpurun_out/ncu_q.log 2>&1
grep -v "^==" gpurun_out/launches_quant.csv | python -c "
import csv,sys
r=list(csv.reader(sys.stdin))
for x in r[1:]: print(x[4][:40], x[7], x[8], x[-1])"
