#!/bin/bash
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 step %.3f ms  k2 %s' % (d['ms_per_step'], list(r['other_kernels_ms'].values())[0]))"; }
for V in b1r0 b2r0 b4r0 b1r1 b2r1 b4r1; do
  STOAT_GPU_FISHER=$V timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show $V
done
