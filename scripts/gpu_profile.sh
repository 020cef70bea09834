#!/bin/bash
# Run on a B200 box (under gpurun): GPU tests, bench lines, ncu launch lists and full captures of the hot kernels.
# Outputs land in gpurun_out/; summaries are copied into profiles/ by scripts/summarise_profiles.py afterwards.
set -u
OUT=gpurun_out
mkdir -p $OUT
K='regex:path_and_popc|binary_prep|binary_stats|fisher_exact|repack_rows|build_case_mask|dosage_moments|lean_rec|ols_solve|student_tail|quant_general|centre_permute|eqtl'
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > $OUT/bench_binary.json 2> $OUT/bench_binary.err; cat $OUT/bench_binary.json; tail -3 $OUT/bench_binary.err
python bench.py --trait quantitative --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_quant.json 2> $OUT/bench_quant.err; cat $OUT/bench_quant.json; tail -3 $OUT/bench_quant.err
$BENCH > $OUT/plain_binary.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file $OUT/launches_binary.csv $BENCH > $OUT/ncu_launches_binary.log 2>&1
$BENCH > $OUT/plain_binary.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:path_and_popc_tma -s 3 -c 1 -f -o $OUT/prof_k1 $BENCH > $OUT/ncu_k1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fisher_exact -s 3 -c 1 -f -o $OUT/prof_fisher $BENCH > $OUT/ncu_fisher.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:binary_stats -s 3 -c 1 -f -o $OUT/prof_stats $BENCH > $OUT/ncu_stats.log 2>&1
BQ="python bench.py --trait quantitative --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
$BQ > $OUT/plain_quant.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 100 --csv --log-file $OUT/launches_quant.csv $BQ > $OUT/ncu_launches_quant.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dosage_moments_mma_kernel -s 3 -c 1 -f -o $OUT/prof_k3m $BQ > $OUT/ncu_k3m.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dosage_moments_mma3_kernel -s 3 -c 1 -f -o $OUT/prof_k3m3 $BQ > $OUT/ncu_k3m3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:student_tail -s 3 -c 1 -f -o $OUT/prof_k4b $BQ > $OUT/ncu_k4b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ols_solve -s 3 -c 1 -f -o $OUT/prof_k4 $BQ > $OUT/ncu_k4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dosage_moments_tma -s 3 -c 1 -f -o $OUT/prof_k3 $BQ > $OUT/ncu_k3.log 2>&1
ls -la $OUT
timeout 900 python bench.py --trait eqtl --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_eqtl.json 2> $OUT/bench_eqtl.err; cut -c1-300 $OUT/bench_eqtl.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_binary_ref.json 2> $OUT/bench_binary_ref.err; cut -c1-300 $OUT/bench_binary_ref.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
