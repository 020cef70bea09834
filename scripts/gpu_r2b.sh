#!/bin/bash
# Round 2, GPU call B: bench lines (C3 default with nested traits, C5 per-GPU shard), ncu launch lists and full captures of
# the new Fisher loop and the column-chunked K1 at the C5 width.
set -u
OUT=gpurun_out
mkdir -p $OUT
K='regex:path_and_popc|binary_prep|binary_stats|fisher_exact|repack_rows|build_case_mask|dosage_moments|lean_rec|ols_solve|student_tail|quant_general|centre_permute|eqtl|slow_list'
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench_default.json 2> $OUT/bench_default.err; cut -c1-400 $OUT/bench_default.json; tail -3 $OUT/bench_default.err
/usr/bin/time -v timeout 1200 python bench.py --config c5 --steps 10 --warmup 3 > $OUT/bench_c5.json 2> $OUT/bench_c5.err; cut -c1-3000 $OUT/bench_c5.json; tail -25 $OUT/bench_c5.err | grep -E "Elapsed|Maximum resident|Error|error" 
B3="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --traits headline"
B5="python bench.py --snarls 200000 --samples 25000 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --traits headline"
$B3 > $OUT/plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file $OUT/launches_binary.csv $B3 > $OUT/ncu_launches_binary.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fisher_exact -s 3 -c 1 -f -o $OUT/prof_fisher $B3 > $OUT/ncu_fisher.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:path_and_popc_tma -s 3 -c 1 -f -o $OUT/prof_k1 $B3 > $OUT/ncu_k1.log 2>&1
$B5 > $OUT/plain_c5w.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:path_and_popc_tma -s 3 -c 1 -f -o $OUT/prof_k1_c5w $B5 > $OUT/ncu_k1_c5w.log 2>&1
ls -la $OUT | head -30
