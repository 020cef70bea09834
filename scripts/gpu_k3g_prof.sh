#!/bin/bash
mkdir -p gpurun_out
BQ1="python bench.py --seed 43 --trait quantitative --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:quant_general -s 3 -c 1 -f -o gpurun_out/prof_general $BQ1 > gpurun_out/ncu_general.log 2>&1
tail -2 gpurun_out/ncu_general.log
