#!/bin/bash
mkdir -p gpurun_out
BQ1="python bench.py --trait quantitative --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:student_tail -s 3 -c 1 -f -o gpurun_out/prof_k4b $BQ1 > gpurun_out/ncu_k4b.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dosage_moments_mma_kernel -s 3 -c 1 -f -o gpurun_out/prof_k3m $BQ1 > gpurun_out/ncu_k3m.log 2>&1
ls -la gpurun_out/*.ncu-rep
