#!/bin/bash
# 4 GPUs: binary and quantitative bench lines, launched as the driver does
mkdir -p gpurun_out
for T in binary quantitative; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --trait $T --warmup 3 --steps 10 2> gpurun_out/b4_$T.err | tail -1 > gpurun_out/bench4_$T.json
  python -c "
import json
d=json.loads(open('gpurun_out/bench4_$T.json').read().strip().splitlines()[-1])
print('$T n_gpus', d['n_gpus'], 'value %.4e' % d['value'], 'e2e %.4e' % d['e2e']['value'], 'ms/step %.2f' % d['ms_per_step'])" || tail -5 gpurun_out/b4_$T.err
done
