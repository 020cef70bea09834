timeout 120 python scripts/umma_probe.py 2>&1 | tail -1
timeout 600 python -m pytest tests/test_gpu_eqtl.py -x -q 2>&1 | tail -3
timeout 600 python bench.py --trait eqtl --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],1), 'ms', d['parity']['mismatches'], d['parity'].get('first_mismatch'), d['roofline']['frac'])"
