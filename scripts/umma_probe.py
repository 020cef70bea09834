#!/usr/bin/env python3
"""Probe of the tcgen05 int8 kernel's mechanics: C = A @ B^T against numpy, growing shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stoat_b200

rng = np.random.default_rng(1)
with stoat_b200.Context(0) as ctx:
    for k, rows, kind in ((128, 256, "ones"), (128, 256, "rand"), (512, 256, "rand"), (1024, 768, "rand"), (5120, 1024, "rand")):
        if kind == "ones":
            a = np.ones((128, k), np.int8)
            b = np.ones((rows, k), np.int8)
        else:
            a = rng.integers(0, 3, size=(128, k), dtype=np.int8)
            b = rng.integers(-64, 65, size=(rows, k), dtype=np.int8)
        c = ctx.debug_umma_probe(a, b)
        want = a.astype(np.int32) @ b.astype(np.int32).T
        ok = np.array_equal(c, want)
        print(f"k={k} rows={rows} {kind}: {'OK' if ok else 'MISMATCH'}", flush=True)
        if not ok:
            bad = np.argwhere(c != want)
            print("  mismatches", bad.shape[0], "of", c.size, "first", bad[:5].tolist())
            print("  got", c[:2, :10].tolist(), "\n  want", want[:2, :10].tolist())
            print("  rows ok", int((c == want).all(axis=1).sum()), "cols ok", int((c == want).all(axis=0).sum()))
            sys.exit(1)
print("probe passed")
