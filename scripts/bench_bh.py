#!/usr/bin/env python3
"""stoat_gpu_bh_adjust (host buffers in and out, copies included) against the numpy restatement of the reference's
adjust_pvalues_with_BH on the host, n = 1M and 10M p-values.  python scripts/bench_bh.py > gpurun_out/bh.json"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402  (the checker; timed here as the CPU baseline of this row)
import stoat_b200  # noqa: E402

res = []
with stoat_b200.Context(0) as ctx:
    for n in (1_000_000, 10_000_000):
        p = np.random.default_rng(n).random(n) ** 4
        ctx.bh_adjust(p[:1000])
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter()
            got = ctx.bh_adjust(p)
            best = min(best, time.perf_counter() - t0)
        t0 = time.perf_counter()
        ref = oracle.bh_adjust(p)
        t_cpu = time.perf_counter() - t0
        res.append({"n": n, "gpu_ms": best * 1e3, "cpu_numpy_ms": t_cpu * 1e3, "bit_identical": bool(np.array_equal(got, ref))})
print(json.dumps({"row": "SURVEY 8f-3 BH adjustment", "results": res}))
