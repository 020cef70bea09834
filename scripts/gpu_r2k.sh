#!/bin/bash
# Round 2, GPU call K: whole GPU suite, the default bench line (what the driver runs), the reference arm, ncu of the tcgen05 kernel.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
( time timeout 1200 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err ) 2>&1 | grep real; cut -c1-300 $OUT/bench_default.json; tail -3 $OUT/bench_default.err
( time timeout 900 python bench.py --impl reference > $OUT/bench_reference.json 2> $OUT/bench_reference.err ) 2>&1 | grep real; cut -c1-400 $OUT/bench_reference.json
B="python bench.py --trait eqtl --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eqtl_dense_umma -s 1 -c 1 -f -o $OUT/prof_eqtl $B > $OUT/ncu_eqtl.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:eqtl' -c 60 --csv --log-file $OUT/launches_eqtl.csv $B > $OUT/ncu_launches_eqtl.log 2>&1
ls -la $OUT | grep -E "prof_eqtl|launches_eqtl"
