#!/bin/bash
mkdir -p gpurun_out
BQ="python bench.py --trait quantitative --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
for n in 1 2 3 4 6 8; do
  STOAT_GPU_K4B_BLOCKS=$n timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:student_tail -c 3 --csv --log-file gpurun_out/k4b_$n.csv $BQ > /dev/null 2>&1
  echo "blocks/SM $n: $(grep -v '^==' gpurun_out/k4b_$n.csv | tail -2 | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' ')"
done
