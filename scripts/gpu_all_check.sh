#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 step %.3f ms value %.3e e2e %s frac %.3f off-fast %s' % (d['ms_per_step'], d['value'], (d.get('e2e') or {}).get('value'), r['frac'], r.get('snarls_off_fast_path')))"; }
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show binary
timeout 300 python bench.py --trait quantitative --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show quant
timeout 600 python bench.py --trait eqtl --steps 2 --warmup 1 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show eqtl
