#!/bin/bash
# 2 GPUs: multi-GPU parity + the bench at N=2 (binary and eQTL), launched as the driver does
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -5
for T in binary quantitative eqtl; do
  EXTRA=""; [ $T = eqtl ] && EXTRA="--steps 3"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --trait $T --warmup 3 $EXTRA 2> gpurun_out/bench2_$T.err | tail -1 > gpurun_out/bench2_$T.json
  python -c "
import json,sys
d=json.loads(open('gpurun_out/bench2_$T.json').read().strip().splitlines()[-1])
print('$T n_gpus', d['n_gpus'], 'value %.4e' % d['value'], 'e2e %.4e' % d['e2e']['value'], 'ms/step %.2f' % d['ms_per_step'])" || tail -5 gpurun_out/bench2_$T.err
done
