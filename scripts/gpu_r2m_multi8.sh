#!/bin/bash
# Round 2, multi-GPU call L (N = 8): NCCL tests, the default line with nested traits, C5 both shardings, eQTL gene-sharded.
set -u
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531"
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -3
timeout 1200 $TR bench.py --gpus 8 --steps 10 --warmup 3 2>$OUT/bench_default_8gpu.err | grep '^{' > $OUT/bench_default_8gpu.json; cut -c1-250 $OUT/bench_default_8gpu.json; tail -2 $OUT/bench_default_8gpu.err
timeout 900 $TR bench.py --gpus 8 --impl reference --steps 3 --warmup 1 2>/dev/null | grep '^{' > $OUT/bench_reference_8gpu.json; cut -c1-250 $OUT/bench_reference_8gpu.json
timeout 1200 $TR bench.py --gpus 8 --config c5 --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/bench_c5_snarl_8gpu.err | grep '^{' > $OUT/bench_c5_snarl_8gpu.json; cut -c1-250 $OUT/bench_c5_snarl_8gpu.json
timeout 1200 $TR bench.py --gpus 8 --config c5 --sharding sample --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/bench_c5_sample_8gpu.err | grep '^{' > $OUT/bench_c5_sample_8gpu.json; cut -c1-250 $OUT/bench_c5_sample_8gpu.json
timeout 1200 $TR bench.py --gpus 8 --trait eqtl --sharding gene --steps 3 --warmup 3 --no-cpu-baseline 2>$OUT/bench_eqtl_gene_8gpu.err | grep '^{' > $OUT/bench_eqtl_gene_8gpu.json; cut -c1-250 $OUT/bench_eqtl_gene_8gpu.json
timeout 600 host/shim_selftest bench 1000000 5000 0,1,2,3,4,5,6,7 $OUT/shim_bench_c3_8gpu.json 2>&1 | tail -1
