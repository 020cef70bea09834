#!/bin/bash
# N = 4 and 8: the default (binary) bench launched as the driver does
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | wc -l
for N in 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --warmup 3 --steps 10 2> gpurun_out/bench${N}_binary.err | tail -1 > gpurun_out/bench${N}_binary.json
  python -c "
import json
d=json.loads(open('gpurun_out/bench${N}_binary.json').read().strip().splitlines()[-1])
print('binary n_gpus', d['n_gpus'], 'value %.4e' % d['value'], 'e2e %.4e' % d['e2e']['value'], 'ms/step %.2f' % d['ms_per_step'])" || tail -5 gpurun_out/bench${N}_binary.err
done
