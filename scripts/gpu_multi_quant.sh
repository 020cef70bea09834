#!/bin/bash
mkdir -p gpurun_out
for N in 1 2; do
  if [ $N = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"; fi
  timeout 900 $L bench.py --gpus $N --trait quantitative --warmup 3 --steps 10 --no-e2e --no-cpu-baseline 2> gpurun_out/bq$N.err | tail -1 > gpurun_out/bq$N.json
  python -c "
import json
d=json.loads(open('gpurun_out/bq$N.json').read().strip().splitlines()[-1])
print('quant n_gpus', d['n_gpus'], 'value %.4e' % d['value'], 'ms/step %.2f' % d['ms_per_step'], d['roofline']['kernel_ms'], list(d['roofline']['other_kernels_ms'].values()))" || tail -5 gpurun_out/bq$N.err
done
CUDA_VISIBLE_DEVICES=1 timeout 300 python bench.py --trait quantitative --warmup 3 --steps 10 --no-e2e --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('gpu1 alone ms/step %.2f' % d['ms_per_step'])"
