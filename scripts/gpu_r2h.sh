#!/bin/bash
# Round 2, GPU call H: shim (device-formatted columns, logistic analyzer, multi-context chunk), then the shim bench at C3 scale.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_shim.py -x -q 2>&1 | tail -25 > $OUT/pytest_shim.log; cat $OUT/pytest_shim.log
timeout 900 host/shim_selftest bench 1000000 5000 0 $OUT/shim_bench_c3.json 2>&1 | tail -3
timeout 900 host/shim_selftest bench 1000000 5000 0,0 $OUT/shim_bench_c3_2ctx.json 2>&1 | tail -3
