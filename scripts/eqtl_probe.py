#!/usr/bin/env python3
"""Quick timing of the dense eQTL scan on a B200 (not the bench): pairs/s, DMMA flops vs a cuBLAS DGEMM."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import stoat_b200
from stoat_b200 import synth

S, n, G = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (20000, 5000, 2048)))
dev = torch.device("cuda", 0)
d = synth.torch_dataset(S, n, dev, seed=42, frac_three=0.0, frac_node0=0.0)
gen = torch.Generator(device=dev)
gen.manual_seed(7)
base = torch.rand((G, 1), device=dev, generator=gen, dtype=torch.float64) * 8
expr = base * (1 + (torch.rand((G, n), device=dev, generator=gen, dtype=torch.float64) * 2 - 1))
a = torch.randn((4096, 4096), device=dev, dtype=torch.float64)
b = torch.randn((4096, 4096), device=dev, dtype=torch.float64)
for _ in range(2):
    a @ b
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    a @ b
torch.cuda.synchronize()
dgemm = 5 * 2 * 4096**3 / (time.perf_counter() - t0) / 1e12
print(f"cuBLAS DGEMM 4096^3: {dgemm:.1f} TFLOP/s")
with stoat_b200.Context(0) as ctx:
    ctx.set_matrix_dev(d["bits_dev"].data_ptr(), d["n_rows"], d["n_cols"])
    ctx.stage_paths(d["paths"])
    ctx.set_expression_dev(expr.data_ptr(), G)
    for i in range(3):
        ctx.eqtl_dense_run_staged(p_threshold=1e-5)
        t = ctx.timing()
        n_pad = (ctx.row_pitch // 16) * 64
        fl = 2.0 * n_pad * S * G
        print(f"run {i}: E1 {t["table_ms"]:.2f} ms, rest {t["stats_ms"]:.2f} ms ({t['n_general']} snarls), "
              f"{S * G / (t["table_ms"] + t["stats_ms"]) * 1e3:.3e} pairs/s, {fl / t['table_ms'] / 1e9:.2f} TFLOP/s in E1")
    r = ctx.eqtl_summary_fetch()
    print("min p:", np.nanmin(r["min_p"]), "n_below total:", int(r["n_below"].sum()), "filtered:", int(np.isnan(r["min_p"]).sum()))
