#!/bin/bash
# Fisher kernel: blocks per SM (every warp ends with a drain in which its lanes run dry one by one)
mkdir -p gpurun_out
BB="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
for n in 3 4 5 6 8 10 12 16; do
  STOAT_GPU_FISHER_BLOCKS=$n timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:fisher_exact -c 3 --csv --log-file gpurun_out/fb_$n.csv $BB > /dev/null 2>&1
  echo "blocks/SM $n: $(grep -v '^==' gpurun_out/fb_$n.csv | tail -2 | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' ') step $(STOAT_GPU_FISHER_BLOCKS=$n timeout 300 $BB 2>/dev/null | tail -1 | python -c "import sys,json; print('%.3f' % json.loads(sys.stdin.read())['ms_per_step'])")"
done
