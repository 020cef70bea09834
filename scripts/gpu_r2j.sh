#!/bin/bash
# Round 2, GPU call J: the tcgen05 int8 eQTL contraction: probe, parity tests, C4 bench against the FP64 DMMA form.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 120 python scripts/umma_probe.py 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_eqtl.py -x -q 2>&1 | tail -15 > $OUT/pytest_eqtl.log; cat $OUT/pytest_eqtl.log
timeout 900 python bench.py --trait eqtl --steps 3 --warmup 3 --no-cpu-baseline 2>$OUT/bench_eqtl_umma.err | grep '^{' > $OUT/bench_eqtl_umma.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_eqtl_umma.json')); print('umma', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['other_kernels_ms'], d['parity'], d['e2e'] and d['e2e']['ms_per_step'])
PY
tail -3 $OUT/bench_eqtl_umma.err
STOAT_GPU_EQTL=dmma timeout 900 python bench.py --trait eqtl --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | grep '^{' > $OUT/bench_eqtl_dmma.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_eqtl_dmma.json')); print('dmma', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['parity'])
PY
