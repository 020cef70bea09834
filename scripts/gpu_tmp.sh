mkdir -p gpurun_out
{
for n in 5000 25000; do
  STOAT_GPU_K3G_TEAM=0 timeout 300 python scripts/k3g_cliff.py $n
  timeout 300 python scripts/k3g_cliff.py $n
done
} > gpurun_out/r2o_cliff.txt 2>&1
cat gpurun_out/r2o_cliff.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2o_pytest.txt 2>&1
tail -5 gpurun_out/r2o_pytest.txt
