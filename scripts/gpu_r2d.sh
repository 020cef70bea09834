#!/bin/bash
# Round 2, GPU call D: parity of the logistic path (class counts + IRLS, general kernel).
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_logit.py -x -q 2>&1 | tail -25 > $OUT/pytest_logit.log; cat $OUT/pytest_logit.log
