mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quant.py tests/test_gpu_wide.py tests/test_gpu_vs_reference.py -m gpu -x -q > gpurun_out/r2p_pytest.txt 2>&1
tail -3 gpurun_out/r2p_pytest.txt
for cfg in 0 8x2 10x2 11x2 6x3 7x3; do
  STOAT_GPU_K3M_RING=$cfg timeout 300 python bench.py --no-cpu-baseline --no-e2e --trait quantitative --traits headline --steps 20 --warmup 3 > gpurun_out/r2p_ring_$cfg.json 2>gpurun_out/r2p_ring_$cfg.err
  python -c "import json;d=json.load(open('gpurun_out/r2p_ring_$cfg.json'));r=d['roofline'];print('ring $cfg', d['ms_per_step'], r['kernel_ms'], r['frac'], d['parity']['mismatches'])"
done
