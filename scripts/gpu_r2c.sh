#!/bin/bash
# Round 2, GPU call C: the C5 per-GPU shard (1.25M snarls x 50,000 haplotypes, binary + nested quantitative) on one GPU.
set -u
OUT=gpurun_out
mkdir -p $OUT
free -g | head -2; nproc
timeout 1500 python bench.py --config c5 --steps 10 --warmup 3 > $OUT/bench_c5.json 2> $OUT/bench_c5.err; cut -c1-600 $OUT/bench_c5.json; tail -5 $OUT/bench_c5.err
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
