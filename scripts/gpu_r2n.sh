#!/bin/bash
# Round 2, GPU call N (N = 2): whole GPU suite after the K3G / K3M changes, sample-sharded quantitative line, default line.
set -u
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541"
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
timeout 1200 $TR bench.py --gpus 2 --config c5 --sharding sample --trait quantitative --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/bench_c5_sample_quant_2gpu.err | grep '^{' > $OUT/bench_c5_sample_quant_2gpu.json; cut -c1-200 $OUT/bench_c5_sample_quant_2gpu.json; tail -4 $OUT/bench_c5_sample_quant_2gpu.err
timeout 1200 $TR bench.py --gpus 2 --config c3 --sharding sample --trait quantitative --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/bench_c3_sample_quant_2gpu.err | grep '^{' > $OUT/bench_c3_sample_quant_2gpu.json; cut -c1-200 $OUT/bench_c3_sample_quant_2gpu.json; tail -4 $OUT/bench_c3_sample_quant_2gpu.err
