#!/usr/bin/env python3
"""Randomised parity campaign on the GPU: seeded random problems of every shape the generators of tests/helpers.py and
stoat_b200/synth.py can draw, through all four analyzers of the C-ABI, against the oracle.  The parity tests of tests/ pin
chosen cases; this walks the space between them for a time budget and prints every configuration that fails (seed included,
so that it can be turned into a test).

    python scripts/fuzz_parity.py [seconds] [first_seed]
"""
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

import helpers as H
import oracle
import stoat_b200
from stoat_b200 import SnarlPaths, synth
from test_gpu_logit import assert_logit_equal


def assert_ols_equal(got, ref, rtol=1e-9):
    """As helpers.assert_quant_equal, but rows whose fit the reference itself computes from rounding noise are left out:
    a saturated or near-perfect fit (1 - r2 < 1e-8: rss is what the cancellation left) has no digits to compare."""
    for k in ("allele_paths", "filtered", "n_used"):
        np.testing.assert_array_equal(got[k], ref[k], err_msg=k)
    if ref["code"].shape == ref["filtered"].shape:
        keep = ref["filtered"] == 0
        np.testing.assert_array_equal(got["code"][keep], ref["code"][keep], err_msg="code")
        m = keep & (ref["code"] == 0)
    else:  # eQTL: one row per (snarl, gene)
        np.testing.assert_array_equal(got["code"], ref["code"], err_msg="code")
        m = ref["code"] == 0
    with np.errstate(invalid="ignore"):
        m = m & (ref["r2"] < 1.0 - 1e-8) & np.isfinite(ref["se"]) & (ref["se"] > 0)
    atols = dict(p_value=1e-300, r2=1e-12, beta=1e-12, se=1e-300)
    worst = []
    for k in ("p_value", "r2", "beta", "se"):
        a, b = got[k][m], ref[k][m]
        err = np.abs(a - b)
        ok = (np.isnan(a) & np.isnan(b)) | (err <= rtol * np.abs(b) + atols[k])
        bad = np.flatnonzero(~ok)
        if bad.size:
            rel = err[bad] / np.maximum(np.abs(b[bad]), 1e-300)
            j = bad[int(np.argmax(rel))]
            worst.append(f"{k}: {bad.size} mismatches, worst rel {rel.max():.2e} got {a[j]!r} want {b[j]!r} "
                         f"row {np.flatnonzero(m)[j]} r2 {ref['r2'][m][j]!r} n_used {ref['n_used'].ravel()[0]}")
    assert not worst, "; ".join(worst)


def draw_problem(rng):
    kind = rng.choice(["partition", "random", "synth", "wide_partition"], p=[0.35, 0.35, 0.2, 0.1])
    if kind == "partition":
        n = int(rng.choice([1, 2, 3, 7, 31, 32, 33, 64, 100, 257, 512, 700, 1025, 2100]))
        dense, snarls = H.partition_problem(rng, int(rng.integers(5, 200)), n, max_paths=int(rng.integers(2, 9)),
                                            p_missing_hap=float(rng.choice([0.0, 0.01, 0.05, 0.4])))
    elif kind == "wide_partition":
        n = int(rng.choice([6200, 7000, 9000, 13000]))
        dense, snarls = H.partition_problem(rng, int(rng.integers(5, 40)), n, max_paths=int(rng.integers(2, 6)),
                                            p_missing_hap=float(rng.choice([0.0, 0.01, 0.3])))
    elif kind == "random":
        n = int(rng.choice([1, 5, 30, 64, 150, 333, 700, 1500]))
        dense, snarls = H.random_snarl_problem(rng, int(rng.integers(5, 250)), n, int(rng.integers(4, 90)),
                                               max_paths=int(rng.integers(1, 9)), max_edges=int(rng.integers(1, 6)),
                                               p_missing_edge=float(rng.choice([0.0, 0.05, 0.3])),
                                               p_empty_path=float(rng.choice([0.0, 0.03, 0.2])),
                                               correlated=bool(rng.integers(0, 2)))
    else:
        n = int(rng.choice([60, 200, 500, 1000, 2500]))
        d = synth.numpy_dataset(int(rng.integers(50, 1500)), n, seed=int(rng.integers(0, 1 << 30)),
                                frac_missing_hap=float(rng.choice([0.0, 0.01, 0.1])),
                                frac_three=float(rng.choice([0.0, 0.05, 0.5])), frac_node0=float(rng.choice([0.0, 0.02, 0.3])),
                                frac_missing_edge=float(rng.choice([0.0, 0.02])))
        P = d["paths"]
        return kind, n, d["bits"], d["n_rows"], d["n_cols"], (P.snarl_path_off, P.path_edge_off, P.edge_row)
    return kind, n, oracle.pack_matrix(dense), dense.shape[0], dense.shape[1], H.csr_from_lists(snarls)


def run_trait(ctx, trait, prob, rng):
    """One analyzer on one drawn problem; raises AssertionError on a parity failure.  -> False when the case is skipped."""
    n, bits, n_cols, P, OP, thr, covar, ncov = (prob[k] for k in ("n", "bits", "n_cols", "P", "OP", "thr", "covar", "ncov"))
    if trait == "binary":
        ctx.set_binary_phenotype(prob["is_case"])
        got = ctx.binary_batch(P, **thr)
        H.assert_binary_equal(got, oracle.binary_batch(bits, n_cols, prob["is_case"], OP, **thr))
    elif trait == "quantitative":
        ctx.set_quantitative_phenotype(prob["y"])
        ctx.set_covariates(covar)
        got = ctx.quantitative_batch(P, **thr)
        assert_ols_equal(got, oracle.quantitative_batch(bits, n_cols, prob["y"], OP, covar=covar, **thr))
    elif trait == "logistic":
        if n > 2600:
            return False
        use_cov = bool(ncov and rng.integers(0, 2))
        ctx.set_binary_phenotype(prob["is_case"])
        ctx.set_covariates(covar)
        got = ctx.binary_covar_batch(P, use_covariates=use_cov, **thr)
        ref = oracle.binary_covar_batch(bits, n_cols, prob["is_case"], OP, covar=covar, use_covariates=use_cov, **thr)
        assert_logit_equal(got, ref)
    else:
        expr = prob["expr"]
        if P.n_snarls * expr.shape[0] * n > 3e7:
            return False
        ctx.set_covariates(covar)
        got = ctx.eqtl_batch(P, expr, None, None, **thr)
        assert_ols_equal(got, oracle.eqtl_batch(bits, n_cols, expr, OP, None, None, covar=covar, **thr))
    return True


def draw_case(seed):
    rng = np.random.default_rng(seed)
    kind, n, bits, n_rows, n_cols, (spo, peo, er) = draw_problem(rng)
    thr = dict(min_ind=int(rng.choice([0, 1, 2, 3, 10])), min_hap=int(rng.choice([0, 2, 5, 20])),
               maf=float(rng.choice([0.0, 0.01, 0.05, 0.2])))
    ncov = int(rng.choice([0, 0, 1, 3, 5]))
    covar = rng.normal(size=(n, ncov)) if ncov else None
    is_case = (rng.random(n) < rng.uniform(0.1, 0.9)).astype(np.uint8)
    # phenotype scales the reference itself can resolve: with |mean| / sd ~ 1e7 its raw-y sums are rounding noise
    y_sd, y_mean = float(rng.choice([1.0, 3.0, 50.0])), float(rng.choice([0.0, -40.0, 10.0]))
    y = rng.normal(size=n) * y_sd + y_mean
    n_genes = int(rng.integers(1, 12))
    expr = rng.uniform(0, 8, size=(n_genes, 1)) * (1.0 + rng.uniform(-1, 1, size=(n_genes, n)))
    prob = dict(kind=kind, n=n, bits=bits, n_rows=n_rows, n_cols=n_cols, P=SnarlPaths(spo, peo, er),
                OP=oracle.Paths(spo, peo, er), thr=thr, covar=covar, ncov=ncov, is_case=is_case, y=y, expr=expr,
                tag=f"seed {seed} {kind} n={n} snarls={len(spo) - 1} ncov={ncov} y={y_mean}+{y_sd}z thr={thr}")
    return prob, rng


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    t_end = time.time() + budget
    n_ok, fails, counts = 0, [], {}
    with stoat_b200.Context(0) as ctx:
        seed = seed0
        while time.time() < t_end:
            prob, rng = draw_case(seed)
            ctx.set_matrix(prob["bits"], prob["n_rows"], prob["n_cols"])
            for trait in ("binary", "quantitative", "logistic", "eqtl"):
                tag = f"{prob['tag']} {trait}"
                try:
                    if run_trait(ctx, trait, prob, rng):
                        n_ok += 1
                        counts[trait] = counts.get(trait, 0) + 1
                except AssertionError as e:
                    fails.append((tag, str(e)[:500]))
                    print("FAIL", tag, "::", str(e)[:500].replace("\n", " "), flush=True)
                except Exception as e:  # an ABI error is a finding too
                    fails.append((tag, repr(e)[:500]))
                    print("ERROR", tag, "::", repr(e)[:500], flush=True)
                    traceback.print_exc()
                finally:
                    ctx.set_covariates(None)
            seed += 1
    print(f"fuzz: {n_ok} runs passed {counts}, {len(fails)} failed, seeds {seed0}..{seed - 1}")
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
