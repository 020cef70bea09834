#!/bin/bash
# Round 2, multi-GPU call F (N = 2): NCCL tests, C3 snarl-sharded, C5 snarl- and sample-sharded bench lines.
set -u
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -8 > $OUT/pytest_multi.log; cat $OUT/pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus 2 --steps 10 --warmup 3 --traits headline > $OUT/bench_c3_2gpu.json 2> $OUT/bench_c3_2gpu.err; cut -c1-300 $OUT/bench_c3_2gpu.json
timeout 1200 $TR bench.py --gpus 2 --config c5 --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_c5_snarl_2gpu.json 2> $OUT/bench_c5_snarl_2gpu.err; cut -c1-300 $OUT/bench_c5_snarl_2gpu.json; tail -3 $OUT/bench_c5_snarl_2gpu.err
timeout 1200 $TR bench.py --gpus 2 --config c5 --sharding sample --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_c5_sample_2gpu.json 2> $OUT/bench_c5_sample_2gpu.err; cut -c1-1200 $OUT/bench_c5_sample_2gpu.json; tail -3 $OUT/bench_c5_sample_2gpu.err
