mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_logit.py tests/test_gpu_wide.py tests/test_gpu_shim.py -m gpu -x -q 2>&1 | tail -8
timeout 300 python scripts/fuzz_parity.py 60 50000 > gpurun_out/r2s_fuzz3.txt 2>&1; tail -3 gpurun_out/r2s_fuzz3.txt | cut -c1-700
