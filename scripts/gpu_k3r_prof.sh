#!/bin/bash
mkdir -p gpurun_out
BQ="python bench.py --trait quantitative --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $BQ > gpurun_out/plain_quant.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dosage_moments_reg -s 3 -c 1 -f -o gpurun_out/prof_k3r $BQ > gpurun_out/ncu_k3r.log 2>&1
tail -3 gpurun_out/ncu_k3r.log
