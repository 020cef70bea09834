#!/bin/bash
# Round 2, GPU call R (N = 2): whole GPU suite on two GPUs after the CTA-team general kernel and the covariate kernels,
# sample-sharded quantitative line (deferred snarls replayed through the general kernel), default line at N = 2.
set -u
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541"
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > $OUT/r2r_pytest_gpu_2gpu.log; cat $OUT/r2r_pytest_gpu_2gpu.log
timeout 900 $TR bench.py --gpus 2 --config c5 --sharding sample --trait quantitative --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/r2r_q.err | grep '^{' > $OUT/r2r_bench_c5_sample_quant_2gpu.json; python -c "import json;d=json.load(open('$OUT/r2r_bench_c5_sample_quant_2gpu.json'));print(d['value'], d['ms_per_step'], d['parity'])"; tail -3 $OUT/r2r_q.err
timeout 900 $TR bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline 2>$OUT/r2r_d.err | grep '^{' > $OUT/r2r_bench_default_2gpu.json; python -c "import json;d=json.load(open('$OUT/r2r_bench_default_2gpu.json'));print(d['value'], d['ms_per_step'], d['parity']['mismatches'], d['e2e']['value'])"; tail -3 $OUT/r2r_d.err
