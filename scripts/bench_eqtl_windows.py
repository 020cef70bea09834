#!/usr/bin/env python3
"""eQTL with GENE WINDOWS (found_gene_snarl, src/snarl_analyzer.cpp:471-491: a snarl is regressed against the genes around
it, a handful each), the shape `stoat vcf --eqtl` runs: stoat_gpu_eqtl_batch with a CSR of genes per snarl, with the sweep
kernels of csrc/quant_covar.cuh and with the general kernel (STOAT_GPU_COVAR_FAST=0), without and with covariates.

    python scripts/bench_eqtl_windows.py [snarls] [samples] [genes] [genes per snarl]
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def run(S, n, G, per):
    import stoat_b200
    from stoat_b200 import synth

    d = synth.numpy_dataset(S, n, seed=42, frac_three=0.05, frac_node0=0.02)
    rng = np.random.default_rng(7)
    expr = rng.uniform(0, 8, size=(G, 1)) * (1.0 + rng.uniform(-1, 1, size=(G, n)))
    off = (np.arange(S + 1) * per).astype(np.uint64)
    idx = ((np.arange(S)[:, None] * 3 + np.arange(per)[None, :]) % G).astype(np.uint32).reshape(-1)
    out = {}
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
        for ncov in (0, 3):
            ctx.set_covariates(rng.normal(size=(n, ncov)) if ncov else None)
            for _ in range(2):
                t0 = time.perf_counter()
                res = ctx.eqtl_batch(d["paths"], expr, off, idx)
                wall = time.perf_counter() - t0
            t = ctx.timing()
            out[ncov] = (t["table_ms"], wall * 1e3, int((res["code"] == 0).sum()), float(np.nansum(res["p_value"])))
    return out


if __name__ == "__main__":
    if os.environ.get("_WINDOWS_CHILD"):
        S, n, G, per = (int(v) for v in sys.argv[1:5])
        for ncov, (k, w, kept, chk) in run(S, n, G, per).items():
            print(f"RESULT {ncov} {k:.3f} {w:.1f} {kept} {chk!r}")
        sys.exit(0)
    S, n, G, per = (int(v) for v in (sys.argv[1:5] + ["100000", "5000", "2000", "5"][len(sys.argv) - 1:]))
    rows = {}
    for fast in ("1", "0"):
        env = dict(os.environ, _WINDOWS_CHILD="1", STOAT_GPU_COVAR_FAST=fast)
        txt = subprocess.run([sys.executable, __file__, str(S), str(n), str(G), str(per)], env=env, capture_output=True, text=True)
        for line in txt.stdout.splitlines():
            if line.startswith("RESULT"):
                _, ncov, k, w, kept, chk = line.split()
                rows[(fast, int(ncov))] = (float(k), float(w), int(kept), chk)
        if txt.returncode:
            print(txt.stderr[-2000:])
    pairs = S * per
    for ncov in (0, 3):
        a, b = rows.get(("1", ncov)), rows.get(("0", ncov))
        if a and b:
            print(f"{S} snarls x {per} genes each ({pairs} pairs), {n} samples, {ncov} covariates: sweep kernels {a[0]:.1f} ms "
                  f"({pairs / a[0] * 1e3:.3g} pair tests/s), general kernel {b[0]:.1f} ms ({pairs / b[0] * 1e3:.3g}); "
                  f"fits {a[2]} / {b[2]}, sum p {a[3]} / {b[3]}")
