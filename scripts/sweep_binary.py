#!/usr/bin/env python3
"""Sweep launch parameters of the binary step on ONE resident data set (the environment variables are read per launch).

  python scripts/sweep_binary.py --snarls 1000000 --samples 5000 --out gpurun_out/sweep_c3.json \
      "STOAT_GPU_OVERLAP=1" "STOAT_GPU_OVERLAP=4" "STOAT_GPU_OVERLAP=4 STOAT_GPU_FISHER_BLOCKS_BESIDE=3" ...

Each positional argument is one configuration: space-separated NAME=VALUE pairs applied on top of the caller's environment."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--snarls", type=int, default=1_000_000)
    ap.add_argument("--samples", type=int, default=5_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--trait", default="binary")
    ap.add_argument("--out", default="")
    ap.add_argument("configs", nargs="*")
    a = ap.parse_args()
    import torch

    import stoat_b200
    from stoat_b200 import synth

    dev = torch.device("cuda", 0)
    d = synth.torch_dataset(a.snarls, a.samples, dev, seed=42)
    ctx = stoat_b200.Context(0)
    ctx.set_matrix_dev(d["bits_dev"].data_ptr(), d["n_rows"], d["n_cols"])
    del d["bits_dev"]
    torch.cuda.empty_cache()
    if a.trait == "binary":
        ctx.set_binary_phenotype(d["is_case"])
        run = ctx.binary_run_staged
    else:
        ctx.set_quantitative_phenotype(d["y"])
        run = ctx.quantitative_run_staged
    ctx.stage_paths(d["paths"])
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    base_env = dict(os.environ)
    rows = []
    ref = None
    for cfg in a.configs or [""]:
        os.environ.clear()
        os.environ.update(base_env)
        for kv in cfg.split():
            k, v = kv.split("=", 1)
            os.environ[k] = v
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(a.steps):
            run()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        tm = ctx.mean_timing(a.steps)
        res = ctx.binary_fetch() if a.trait == "binary" else ctx.quantitative_fetch()
        key = "p_fisher" if a.trait == "binary" else "p_value"
        same = True
        if ref is None:
            ref = res
        else:
            for k in res:
                x, y = res[k], ref[k]
                if x.dtype.kind == "f":
                    same &= bool(np.all((x == y) | (np.isnan(x) & np.isnan(y)) | (np.abs(x - y) <= 1e-12 * np.abs(y))))
                else:
                    same &= bool(np.array_equal(x, y))
        row = {"config": cfg, "ms_per_step": ms, "table_ms": tm["table_ms"], "stats_ms": tm["stats_ms"],
               "same_as_first": same, "n_fisher": ctx.timing()["n_fisher"], "n_general": ctx.timing()["n_general"]}
        rows.append(row)
        print(json.dumps(row), flush=True)
        del key
    if a.out:
        with open(a.out, "w") as f:
            json.dump({"snarls": a.snarls, "samples": a.samples, "trait": a.trait, "rows": rows}, f, indent=1)
    ctx.close()


if __name__ == "__main__":
    main()
