#!/bin/bash
# B200: full GPU test suite, the three bench lines, ncu launch list + full capture of the dense eQTL kernel.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench_binary.json 2> $OUT/bench_binary.err; cut -c1-400 $OUT/bench_binary.json; tail -3 $OUT/bench_binary.err
timeout 900 python bench.py --trait eqtl --steps 5 --warmup 3 > $OUT/bench_eqtl.json 2> $OUT/bench_eqtl.err; cat $OUT/bench_eqtl.json; tail -3 $OUT/bench_eqtl.err
timeout 600 python bench.py --trait eqtl --impl reference --steps 2 --warmup 1 > $OUT/bench_eqtl_ref.json 2> $OUT/bench_eqtl_ref.err; cut -c1-300 $OUT/bench_eqtl_ref.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_binary_ref.json 2> $OUT/bench_binary_ref.err; cut -c1-300 $OUT/bench_binary_ref.json
BE="python bench.py --trait eqtl --snarls 37888 --genes 2432 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
K='regex:eqtl|quant_general|binary_prep|repack_rows'
timeout 600 $BE > $OUT/plain_eqtl.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file $OUT/launches_eqtl.csv $BE > $OUT/ncu_launches_eqtl.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eqtl_dense_dmma -s 3 -c 1 -o $OUT/prof_eqtl $BE > $OUT/ncu_eqtl.log 2>&1
ls -la $OUT | head -40
