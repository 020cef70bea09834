#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_quant.py -m gpu -x -q 2>&1 | tail -2
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 step %.3f ms k3m %.3f frac %.3f' % (d['ms_per_step'], r['kernel_ms'], r['frac']))"; }
BQ="python bench.py --trait quantitative --steps 10 --warmup 3 --no-cpu-baseline --no-e2e"
STOAT_GPU_K3M_PREFETCH=0 timeout 300 $BQ 2>&1 | tail -1 | show "prefetch=0"
timeout 300 $BQ 2>&1 | tail -1 | show "prefetch=1"
STOAT_GPU_K3M_CTAS=1 timeout 300 $BQ 2>&1 | tail -1 | show "prefetch=1 ctas=1"
