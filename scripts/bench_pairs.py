#!/usr/bin/env python3
"""set_matrix (dense reference-layout bytes + K0 repack) against set_matrix_pairs (set bits scattered on the device) for a
sparse cohort: 1M rows x 10,000 haplotypes at a few densities, pinned host buffers, wall time of the call (H2D included).
Checks that both uploads give the same device rows on a sample.  python scripts/bench_pairs.py > gpurun_out/pairs.json"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stoat_b200  # noqa: E402

N_ROWS, N_COLS = 1_000_000, 10_000
out = []
dev = torch.device("cuda", 0)
with stoat_b200.Context(0) as ctx:
    for density in (0.002, 0.01, 0.05):
        g = torch.Generator(device=dev).manual_seed(5)
        n_pairs = int(N_ROWS * N_COLS * density)
        r = torch.randint(0, N_ROWS, (n_pairs,), device=dev, generator=g, dtype=torch.int64)
        c = torch.randint(0, N_COLS, (n_pairs,), device=dev, generator=g, dtype=torch.int64)
        # dense reference layout: bit row * n_cols + col, LSB first (src/matrix.cpp:67-84)
        bit = r * N_COLS + c
        dense = torch.zeros((N_ROWS * N_COLS + 7) // 8, dtype=torch.uint8, device=dev)
        one = torch.ones((), dtype=torch.uint8, device=dev)
        for b in range(8):  # OR in the bits of each position within a byte (scatter of equal values: order free)
            sel = bit[(bit & 7) == b]
            dense[sel >> 3] |= one << b
        h_dense = torch.empty(dense.numel(), dtype=torch.uint8, pin_memory=True).copy_(dense)
        h_r = torch.empty(n_pairs, dtype=torch.int32, pin_memory=True).copy_(r.to(torch.int32))
        h_c = torch.empty(n_pairs, dtype=torch.int32, pin_memory=True).copy_(c.to(torch.int32))
        torch.cuda.synchronize()
        lib, h = ctx._lib, ctx._ctx
        import ctypes as C

        def timed(fn, n=3):
            best = 1e9
            for _ in range(n):
                t0 = time.perf_counter()
                fn()
                best = min(best, time.perf_counter() - t0)
            return best * 1e3
        t_dense = timed(lambda: ctx._check(lib.stoat_gpu_set_matrix(h, C.c_void_p(h_dense.data_ptr()), N_ROWS, N_COLS)))
        ctx.n_rows, ctx.n_cols = N_ROWS, N_COLS
        rows_a = [ctx.read_row(k).copy() for k in (0, 1, 777, N_ROWS - 1)]
        t_pairs = timed(lambda: ctx._check(lib.stoat_gpu_set_matrix_pairs(h, C.c_void_p(h_r.data_ptr()), C.c_void_p(h_c.data_ptr()),
                                                                         n_pairs, N_ROWS, N_COLS)))
        rows_b = [ctx.read_row(k).copy() for k in (0, 1, 777, N_ROWS - 1)]
        same = all(np.array_equal(a, b) for a, b in zip(rows_a, rows_b))
        out.append({"density": density, "n_pairs": n_pairs, "dense_upload_ms": t_dense, "dense_bytes": int(dense.numel()),
                    "pairs_upload_ms": t_pairs, "pairs_bytes": 8 * n_pairs, "rows_equal": bool(same)})
        del dense, bit, r, c, h_dense, h_r, h_c
print(json.dumps({"workload": f"{N_ROWS} rows x {N_COLS} haplotypes", "results": out}))
