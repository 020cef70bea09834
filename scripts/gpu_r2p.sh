#!/bin/bash
# Round 2, GPU call P: whole GPU suite, the default bench line (what the driver runs) with all traits, ncu of the tcgen05
# kernel (8 epilogue warps) and of the K3G team kernel, launch list of the default command.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $OUT/r2p_pytest_gpu.log; cat $OUT/r2p_pytest_gpu.log
( time timeout 1200 python bench.py > $OUT/r2p_bench_default.json 2> $OUT/r2p_bench_default.err ) 2>&1 | grep real; cut -c1-300 $OUT/r2p_bench_default.json; tail -3 $OUT/r2p_bench_default.err
B="python bench.py --trait eqtl --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eqtl_dense_umma -s 1 -c 1 -f -o $OUT/r2p_prof_eqtl $B > $OUT/r2p_ncu_eqtl.log 2>&1
tail -2 $OUT/r2p_ncu_eqtl.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2p_launches_default.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --traits headline > $OUT/r2p_ncu_launches.log 2>&1
tail -2 $OUT/r2p_ncu_launches.log
ls -la $OUT | grep -E "r2p_"
