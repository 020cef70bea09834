#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eqtl.py tests/test_gpu_vs_reference.py -m gpu -x -q 2>&1 | tail -15
timeout 600 python scripts/eqtl_probe.py 37888 5000 2432 2>&1 | tail -8
timeout 900 python scripts/eqtl_probe.py 200000 5000 20000 2>&1 | tail -8
