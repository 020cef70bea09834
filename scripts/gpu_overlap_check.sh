#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 step %.3f ms value %.3e k1 %.3f k2 %s frac %.3f' % (d['ms_per_step'], d['value'], r['kernel_ms'], list(r['other_kernels_ms'].values())[0], r['frac']))"; }
for V in 1 2 4 8; do
  STOAT_GPU_OVERLAP=$V timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show "chunks=$V"
done
