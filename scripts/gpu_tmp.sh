mkdir -p gpurun_out
for cfg in c3 c5; do
for ov in 1 2 3 4 6 8; do
  STOAT_GPU_OVERLAP=$ov timeout 300 python bench.py --config $cfg --no-cpu-baseline --no-e2e --traits headline --steps 30 --warmup 3 > gpurun_out/r2q_overlap_${cfg}_$ov.json 2>gpurun_out/r2q_overlap.err
  python -c "import json;d=json.load(open('gpurun_out/r2q_overlap_${cfg}_$ov.json'));r=d['roofline'];print('$cfg overlap $ov', round(d['ms_per_step'],4), d['parity']['mismatches'], round(r['kernel_ms'],3), r['other_kernels_ms'])"
done
done
