#!/bin/bash
# Round 2, GPU call G: the block form of the Fisher list (default, 32 steps per block): the whole GPU suite, the step at C3 and
# at C5 against the serial form, an ncu capture of the new kernel.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
timeout 600 python scripts/sweep_binary.py --snarls 1000000 --samples 5000 --out $OUT/sweep_fisher_block.json \
  "STOAT_GPU_FISHER=32" "STOAT_GPU_FISHER=serial" "STOAT_GPU_FISHER=8" "STOAT_GPU_FISHER=16" "STOAT_GPU_FISHER=exact" \
  "STOAT_GPU_FISHER=32 STOAT_GPU_FISHER_BLOCKS=6" "STOAT_GPU_FISHER=32 STOAT_GPU_FISHER_BLOCKS=10" 2>&1 | tail -8
timeout 600 python scripts/sweep_binary.py --snarls 1250000 --samples 25000 --out $OUT/sweep_fisher_block_c5.json \
  "STOAT_GPU_FISHER=32" "STOAT_GPU_FISHER=serial" "STOAT_GPU_FISHER=16" 2>&1 | tail -4
B3="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --traits headline"
ncu --set full --clock-control none --import-source on -k regex:fisher_block -s 3 -c 1 -f -o $OUT/prof_fisher $B3 > $OUT/ncu_fisher.log 2>&1
K='regex:path_and_popc|binary_prep|binary_stats|fisher|repack_rows|build_case_mask'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file $OUT/launches_binary.csv $B3 > $OUT/ncu_launches_binary.log 2>&1
ls -la $OUT | grep -E "prof_fisher|launches"
