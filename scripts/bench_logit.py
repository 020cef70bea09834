#!/usr/bin/env python3
"""The binary trait with covariates (BinaryCovarSnarlAnalyzer: logistic regression per snarl) at the C3
shape on one GPU: resident timing of stoat_gpu_binary_covar_run_staged, parity of the first snarls against the oracle.

    python scripts/bench_logit.py [snarls] [samples] [steps]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np


def main():
    import torch

    import stoat_b200
    from stoat_b200 import synth

    S = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    dev = torch.device("cuda:0")
    d = synth.torch_dataset(S, n, dev, seed=42)
    out = {"workload": f"{S} snarls x {2 * n} haplotypes, binary trait with covariates (logistic regression)"}
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix_dev(d["bits_dev"].data_ptr(), d["n_rows"], d["n_cols"])
        ctx.set_binary_phenotype(d["is_case"])
        ctx.stage_paths(d["paths"])
        for mode in ("default", "general"):
            if mode == "general":
                if S > 100_000:
                    continue
                os.environ["STOAT_GPU_LOGIT"] = "general"
            for _ in range(2):
                ctx.binary_covar_run_staged()
                ctx.binary_covar_fetch()
            t0 = time.perf_counter()
            for _ in range(steps):
                ctx.binary_covar_run_staged()
            res = ctx.binary_covar_fetch()
            dt = (time.perf_counter() - t0) / steps * 1e3
            t = ctx.timing()
            out[mode] = {"ms_per_step": dt, "path_tests_per_s": d["paths"].n_paths / dt * 1e3, "timing": t,
                         "fits": int(((res["filtered"] == 0) & (res["code"] == 0)).sum())}
            os.environ.pop("STOAT_GPU_LOGIT", None)
        # parity of the first snarls against the oracle
        import oracle
        import test_gpu_logit

        k = min(S, 5000)
        st = d["structure"]
        r1 = int(st.row0[k - 1] + st.n_rows_of[k - 1])
        row_bytes = d["n_cols"] // 8
        bits = d["bits_dev"][: r1 * row_bytes].cpu().numpy()
        p = d["paths"]
        P1 = int(p.snarl_path_off[k])
        E1 = int(p.path_edge_off[P1])
        sub = oracle.Paths(p.snarl_path_off[: k + 1].copy(), p.path_edge_off[: P1 + 1].copy(), p.edge_row[:E1].copy())
        ref = oracle.binary_covar_batch(bits, d["n_cols"], d["is_case"], sub)
        got = {kk: (v[:P1] if kk == "allele_paths" else v[:k]) for kk, v in res.items()}
        try:
            out["parity"] = {"checked_snarls": k, "fits": test_gpu_logit.assert_logit_equal(got, ref), "mismatches": 0}
        except AssertionError as e:
            out["parity"] = {"checked_snarls": k, "mismatches": 1, "first": str(e)[:300]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
