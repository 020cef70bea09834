ncu --set full --import-source on --clock-control none -k regex:quant_general -s 30 -c 1 -f -o gpurun_out/prof_k3g python scripts/k3g_cliff.py 5000 > /dev/null 2>&1
ls -la gpurun_out/prof_k3g.ncu-rep
