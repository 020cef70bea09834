#!/bin/bash
# K3G with a CTA team per snarl on short lists: cliff numbers with and without, the GPU suite, the overlap sweep of the binary run
mkdir -p gpurun_out
{
for n in 5000 25000; do
  STOAT_GPU_K3G_TEAM=0 timeout 300 python scripts/k3g_cliff.py $n
  timeout 300 python scripts/k3g_cliff.py $n
done
} > gpurun_out/r2o_cliff.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2o_pytest.txt 2>&1
tail -5 gpurun_out/r2o_pytest.txt
cat gpurun_out/r2o_cliff.txt
for ov in 1 2 4; do
  STOAT_GPU_OVERLAP=$ov timeout 300 python bench.py --no-cpu-baseline --no-e2e --traits headline --steps 20 --warmup 3 > gpurun_out/r2o_overlap_$ov.json 2>gpurun_out/r2o_overlap_$ov.err
  python -c "import json;d=json.load(open('gpurun_out/r2o_overlap_$ov.json'));print('overlap $ov', d['ms_per_step'], d['parity']['mismatches'])"
done
timeout 300 python bench.py --no-cpu-baseline --no-e2e --trait quantitative --traits headline --steps 20 --warmup 3 > gpurun_out/r2o_quant.json 2>gpurun_out/r2o_quant.err
python -c "import json;d=json.load(open('gpurun_out/r2o_quant.json'));print('quant', d['ms_per_step'], d['parity']['mismatches'], d['roofline'])"
