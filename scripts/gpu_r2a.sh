#!/bin/bash
# Round 2, GPU call A: parity of the re-tiled K1 / the new Fisher loop / the chunked upload, then sweeps of the overlapped
# binary step at C3 and of K1's column chunk at the C5 width.  Outputs land in gpurun_out/.
set -u
OUT=gpurun_out
mkdir -p $OUT
free -g | head -2; nproc; nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
timeout 600 python scripts/sweep_binary.py --snarls 1000000 --samples 5000 --out $OUT/sweep_c3.json \
  "STOAT_GPU_OVERLAP=1" "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_LEAN=1" "STOAT_GPU_OVERLAP=2" "STOAT_GPU_OVERLAP=4" \
  "STOAT_GPU_OVERLAP=8" "STOAT_GPU_OVERLAP=4 STOAT_GPU_FISHER_BLOCKS_BESIDE=3" \
  "STOAT_GPU_OVERLAP=4 STOAT_GPU_FISHER_BLOCKS_BESIDE=5" "STOAT_GPU_OVERLAP=4 STOAT_GPU_FISHER_BLOCKS_BESIDE=8" \
  "STOAT_GPU_OVERLAP=4 STOAT_GPU_K1_LEAN=0" "STOAT_GPU_OVERLAP=1 STOAT_GPU_FISHER_BLOCKS=4" \
  "STOAT_GPU_OVERLAP=1 STOAT_GPU_FISHER_BLOCKS=6" "STOAT_GPU_OVERLAP=1 STOAT_GPU_FISHER_BLOCKS=12" \
  "STOAT_GPU_OVERLAP=1 STOAT_GPU_FISHER=exact" "STOAT_GPU_OVERLAP=4 STOAT_GPU_FISHER=exact" 2>&1 | tail -16
timeout 600 python scripts/sweep_binary.py --snarls 200000 --samples 25000 --out $OUT/sweep_c5w.json \
  "STOAT_GPU_OVERLAP=1" "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_CHUNK=1024" "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_CHUNK=1280" \
  "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_CHUNK=2048" "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_CHUNK=3136" \
  "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_CHUNK=6272" "STOAT_GPU_OVERLAP=1 STOAT_GPU_K1_CHUNK=1536 STOAT_GPU_K1_STAGES=3" \
  "STOAT_GPU_OVERLAP=4" "STOAT_GPU_OVERLAP=4 STOAT_GPU_K1_CHUNK=1024" 2>&1 | tail -10
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench_default.json 2> $OUT/bench_default.err; cut -c1-1500 $OUT/bench_default.json; tail -5 $OUT/bench_default.err
