mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quant.py -m gpu -x -q 2>&1 | tail -5
for q in 1 3 5; do
timeout 600 python bench.py --trait quantitative --ncov $q --traits headline --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2q_quant_ncov$q.json 2> gpurun_out/r2q_quant_ncov.err
python -c "import json;d=json.load(open('gpurun_out/r2q_quant_ncov$q.json'));print('quant ncov$q', d['value'], d['ms_per_step'], d['parity']['mismatches'], d['roofline']['kernel_ms'], list(d['roofline']['other_kernels_ms'].values())[0], d['roofline'].get('snarls_off_fast_path'))"; tail -3 gpurun_out/r2q_quant_ncov.err
done
