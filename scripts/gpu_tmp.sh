mkdir -p gpurun_out
for v in 0 1 2 4 8; do
STOAT_GPU_E3C_PREFETCH=$v timeout 200 python bench.py --trait eqtl --ncov 3 --snarls 2000 --genes 2000 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2s_eqtl_ncov3.json 2> gpurun_out/r2s_eqtl_ncov3.err
python -c "import json;d=json.load(open('gpurun_out/r2s_eqtl_ncov3.json'));print('eqtl ncov3 prefetch $v', d['value'], d['ms_per_step'], d['parity']['mismatches'])"; tail -3 gpurun_out/r2s_eqtl_ncov3.err
done
