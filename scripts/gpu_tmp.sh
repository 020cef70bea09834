mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:covar_moments_kernel -s 9 -c 1 -f -o gpurun_out/r2q_k3c python bench.py --trait quantitative --ncov 3 --traits headline --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r2q_ncu.log 2>&1
tail -3 gpurun_out/r2q_ncu.log
