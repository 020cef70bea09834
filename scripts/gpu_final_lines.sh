#!/bin/bash
# final bench lines + launch lists of the round (one GPU)
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps 10 --warmup 3 > $OUT/bench_binary.json 2> $OUT/bench_binary.err; cut -c1-250 $OUT/bench_binary.json
python bench.py --trait quantitative --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_quant.json 2> $OUT/bench_quant.err; cut -c1-250 $OUT/bench_quant.json
K='regex:path_and_popc|binary_prep|binary_stats|fisher_exact|repack_rows|build_case_mask|dosage_moments|lean_rec|ols_solve|student_tail|quant_general|centre_permute'
BQ="python bench.py --trait quantitative --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 100 --csv --log-file $OUT/launches_quant.csv $BQ > $OUT/ncu_launches_quant.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dosage_moments_mma2x -s 3 -c 1 -f -o $OUT/prof_k3m2x $BQ > $OUT/ncu_k3m2x.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ols_solve -s 3 -c 1 -f -o $OUT/prof_k4 $BQ > $OUT/ncu_k4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:student_tail -s 3 -c 1 -f -o $OUT/prof_k4b $BQ > $OUT/ncu_k4b.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
