#!/usr/bin/env python3
"""What ONE snarl on the general quantitative kernel (K3G) costs: the same 20,000-snarl batch with and without a five-path
snarl appended (its paths reuse rows of other snarls: more than four paths always take the general kernel)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import stoat_b200
from stoat_b200 import SnarlPaths, synth

n_samples = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
d = synth.numpy_dataset(20000, n_samples, seed=5, frac_three=0.0, frac_node0=0.0)
p = d["paths"]
spo, peo, er = p.snarl_path_off.copy(), p.path_edge_off.copy(), p.edge_row.copy()
extra_rows = np.array([0, 5, 9, 14, 22], np.uint32)
spo2 = np.concatenate([spo, [spo[-1] + 5]]).astype(np.uint64)
peo2 = np.concatenate([peo, peo[-1] + 1 + np.arange(5)]).astype(np.uint64)
er2 = np.concatenate([er, extra_rows]).astype(np.uint32)
with stoat_b200.Context(0) as ctx:
    ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    ctx.set_quantitative_phenotype(d["y"])
    for name, paths in (("plain", SnarlPaths(spo, peo, er)), ("plus one general snarl", SnarlPaths(spo2, peo2, er2))):
        ctx.stage_paths(paths)
        for _ in range(3):
            ctx.quantitative_run_staged()
        ctx.timing()
        t0 = time.perf_counter()
        for _ in range(20):
            ctx.quantitative_run_staged()
        t = ctx.timing()
        dt = (time.perf_counter() - t0) / 20 * 1e3
        print(f"{n_samples} samples, {name}: {dt:.3f} ms per step (table {t['table_ms']:.3f} + stats {t['stats_ms']:.3f}), n_general {t['n_general']}")
