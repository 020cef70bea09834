#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_quant.py -m gpu -x -q 2>&1 | tail -30
timeout 300 python bench.py --trait quantitative --steps 2 --warmup 1 --no-cpu-baseline --no-e2e 2>&1 | tail -15
