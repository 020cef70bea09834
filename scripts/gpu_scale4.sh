#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -14
for B in 0 1; do
  if [ $B = 1 ]; then export STOAT_BENCH_NO_BIND=1; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --warmup 3 --steps 10 2> gpurun_out/b4.err | tail -1 > gpurun_out/b4.json
  python -c "
import json
d=json.loads(open('gpurun_out/b4.json').read().strip().splitlines()[-1])
print('nobind=$B binary n_gpus', d['n_gpus'], 'value %.4e' % d['value'], 'e2e %.4e' % d['e2e']['value'], 'ms/step %.2f' % d['ms_per_step'])" || tail -5 gpurun_out/b4.err
  grep -i affinity gpurun_out/b4.err | head -2
done
