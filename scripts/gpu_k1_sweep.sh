#!/bin/bash
# K1 pipeline-shape sweep on a B200: stages x warps per CTA, binary bench, resident numbers only.
python -m pytest tests/test_gpu_binary.py -m gpu -x -q 2>&1 | tail -3
for cfg in "3 8" "2 13" "2 10" "4 6" "3 6" "2 8"; do
  set -- $cfg
  echo "stages=$1 warps=$2"
  STOAT_GPU_K1_STAGES=$1 STOAT_GPU_K1_WARPS=$2 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | \
    python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('  k1 %.3f ms frac %.3f  other %s  step %.3f ms' % (r['kernel_ms'], r['frac'], r['other_kernels_ms'], d['ms_per_step']))"
done
