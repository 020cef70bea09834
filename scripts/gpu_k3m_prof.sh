#!/bin/bash
mkdir -p gpurun_out
BQ="python bench.py --trait quantitative --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 quant step %.3f ms  off-fast %s frac %.3f' % (d['ms_per_step'], r.get('snarls_off_fast_path'), r['frac']))"; }
timeout 600 python -m pytest tests/test_gpu_quant.py tests/test_gpu_vs_reference.py -m gpu -x -q 2>&1 | tail -3
timeout 300 $BQ 2>&1 | tail -3 | show "K3M"
BQ1="python bench.py --trait quantitative --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dosage_moments_mma -s 3 -c 1 -f -o gpurun_out/prof_k3m $BQ1 > gpurun_out/ncu_k3m.log 2>&1
tail -2 gpurun_out/ncu_k3m.log
