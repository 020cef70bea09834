#!/bin/bash
# Round 2, multi-GPU call I (N = 2): shim over two devices, sample-sharded C5 with the fixed parity gather, eQTL gene-sharded.
set -u
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
STOAT_SHIM_DEVICES=0,1 timeout 600 python -m pytest tests/test_gpu_shim.py -x -q 2>&1 | tail -4
timeout 900 host/shim_selftest bench 1000000 5000 0,1 $OUT/shim_bench_c3_2gpu.json 2>&1 | tail -2
timeout 1200 $TR bench.py --gpus 2 --config c5 --sharding sample --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/bench_c5_sample_2gpu.err | grep '^{' > $OUT/bench_c5_sample_2gpu.json; cut -c1-200 $OUT/bench_c5_sample_2gpu.json; tail -2 $OUT/bench_c5_sample_2gpu.err
timeout 1500 $TR bench.py --gpus 2 --trait eqtl --sharding gene --steps 3 --warmup 3 --no-cpu-baseline 2>$OUT/bench_eqtl_gene_2gpu.err | grep '^{' > $OUT/bench_eqtl_gene_2gpu.json; cut -c1-300 $OUT/bench_eqtl_gene_2gpu.json; tail -2 $OUT/bench_eqtl_gene_2gpu.err
timeout 1500 $TR bench.py --gpus 2 --trait eqtl --steps 3 --warmup 3 --no-cpu-baseline 2>$OUT/bench_eqtl_snarl_2gpu.err | grep '^{' > $OUT/bench_eqtl_snarl_2gpu.json; cut -c1-300 $OUT/bench_eqtl_snarl_2gpu.json; tail -2 $OUT/bench_eqtl_snarl_2gpu.err
