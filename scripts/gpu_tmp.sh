mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_eqtl.py -m gpu -x -q 2>&1 | tail -4
for v in "0 1" "2 1"; do
set -- $v
STOAT_GPU_E3C=$1 STOAT_GPU_E3C_CACHE=$2 timeout 200 python bench.py --trait eqtl --ncov 3 --snarls 2000 --genes 2000 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2s_eqtl_ncov3.json 2> gpurun_out/r2s_eqtl_ncov3.err
python -c "import json;d=json.load(open('gpurun_out/r2s_eqtl_ncov3.json'));print('eqtl ncov3 variant $1 cache $2', d['value'], d['ms_per_step'], d['parity']['mismatches'])"; tail -3 gpurun_out/r2s_eqtl_ncov3.err
done
