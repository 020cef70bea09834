mkdir -p gpurun_out
K='regex:^(binary_|build_case|centre_|covar_|dosage_|eqtl_|fisher_|lean_rec|logit_|ols_solve|path_and|quant_general|repack_rows|student_tail)'
timeout 250 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 600 --csv --log-file gpurun_out/launches_binary.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_l1.log 2>&1
wc -l gpurun_out/launches_binary.csv
