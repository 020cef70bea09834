mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_eqtl.py tests/test_gpu_shim.py tests/test_gpu_vs_reference.py -m gpu -x -q 2>&1 | tail -6
timeout 600 python scripts/bench_eqtl_windows.py 100000 5000 2000 5 2>&1 | tail -4 | tee gpurun_out/r2s_eqtl_windows.txt
