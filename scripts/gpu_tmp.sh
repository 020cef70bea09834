mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eqtl.py tests/test_gpu_quant.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python scripts/fuzz_parity.py 100 21850 > gpurun_out/r2r_fuzz2.txt 2>&1; tail -3 gpurun_out/r2r_fuzz2.txt | cut -c1-700
timeout 600 python bench.py --trait eqtl --ncov 3 --snarls 2000 --genes 2000 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2q_eqtl_ncov3.json 2> gpurun_out/r2q_eqtl_ncov3.err
python -c "import json;d=json.load(open('gpurun_out/r2q_eqtl_ncov3.json'));print('eqtl ncov3', d['value'], d['ms_per_step'], d['parity']['mismatches'])"; tail -3 gpurun_out/r2q_eqtl_ncov3.err
timeout 600 python bench.py --trait quantitative --ncov 3 --traits headline --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2q_quant_ncov3.json 2> gpurun_out/r2q_quant_ncov.err
python -c "import json;d=json.load(open('gpurun_out/r2q_quant_ncov3.json'));print('quant ncov3', d['value'], d['ms_per_step'], d['parity']['mismatches'], d['roofline'].get('snarls_off_fast_path'))"
