#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eqtl.py tests/test_gpu_vs_reference.py tests/test_gpu_binary.py -m gpu -x -q 2>&1 | tail -8
timeout 900 python bench.py --trait eqtl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_eqtl.json 2> gpurun_out/bench_eqtl.err; cat gpurun_out/bench_eqtl.json | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print('value %.4e e2e %.4e step %.1f ms  E1 %.1f ms %.2f TF/s frac %.3f peak %.1f dense %d general %d rest %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['kernel_ms'], r['achieved'], r['frac'], r['peak'], r['snarls_on_tensor_path'], r['snarls_on_general_path'], r['other_kernels_ms']))"; tail -3 gpurun_out/bench_eqtl.err
