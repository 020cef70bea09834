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


NOISE = {"ols_rows_nearer_to_exact_than_the_oracle": 0, "worst_oracle_error_rel": 0.0, "worst_gpu_error_rel_on_those": 0.0,
         "ols_saturated_rows": 0, "ols_beta_rows_beyond_1e-9_within_1e-10_se": 0, "logit_runs_beyond_1e-9_within_1e-8": 0}


def assert_ols_equal(got, ref, rtol=1e-9, arbiter=None, min_rows=0):
    """As helpers.assert_quant_equal, except where the reference's own numbers are rounding noise (counted in NOISE):
    * a saturated or near-perfect fit (1 - r2 < 1e-8, or no finite r2 at all: one sample, more features than rows) — rss
      is what a cancellation left, so se, t and p have no digits to compare;
    * a coefficient that is small against its standard error comes out of the cancellation inv(X'X) X'y of terms of the
      size of sum(y): it is compared with 1e-9 |beta| + 1e-10 se (its uncertainty is se, its rounding noise ~1e-12 se)."""
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
        sound = np.isfinite(ref["r2"]) & (ref["r2"] < 1.0 - 1e-8) & np.isfinite(ref["se"]) & (ref["se"] > 0)
    if min_rows:  # as many features as rows: the fit is saturated whatever r2 the pseudo-inverse branch prints
        nu = ref["n_used"] if ref["code"].shape == ref["filtered"].shape else np.repeat(ref["n_used"], ref["code"].size // max(ref["n_used"].size, 1))
        sound = sound & (nu > min_rows)
    NOISE["ols_saturated_rows"] += int((m & ~sound).sum())
    m = m & sound
    atols = dict(p_value=1e-300, r2=1e-12, beta=1e-12, se=1e-300)
    strict = np.abs(got["beta"][m] - ref["beta"][m]) <= rtol * np.abs(ref["beta"][m]) + 1e-12
    atols["beta"] = 1e-10 * ref["se"][m]
    NOISE["ols_beta_rows_beyond_1e-9_within_1e-10_se"] += int(
        (~strict & (np.abs(got["beta"][m] - ref["beta"][m]) <= rtol * np.abs(ref["beta"][m]) + atols["beta"])).sum())
    worst = []
    for k in ("p_value", "r2", "beta", "se"):
        a, b = got[k][m], ref[k][m]
        err = np.abs(a - b)
        ok = (np.isnan(a) & np.isnan(b)) | (err <= rtol * np.abs(b) + atols[k])
        bad = np.flatnonzero(~ok)
        if bad.size and arbiter is not None and k != "r2":
            # rows beyond 1e-9: accepted when the CUDA value is no farther from the extended-precision value than the
            # oracle's (the reference's own order of operations) is, i.e. the difference is the reference's rounding error
            still = []
            for j in bad:
                ex = arbiter(int(np.flatnonzero(m)[j]))[k]
                e_gpu, e_ref = abs(a[j] - ex), abs(b[j] - ex)
                if e_gpu <= e_ref + rtol * abs(ex):
                    NOISE["ols_rows_nearer_to_exact_than_the_oracle"] += 1
                    NOISE["worst_oracle_error_rel"] = max(NOISE["worst_oracle_error_rel"], e_ref / max(abs(ex), 1e-300))
                    NOISE["worst_gpu_error_rel_on_those"] = max(NOISE["worst_gpu_error_rel_on_those"],
                                                                e_gpu / max(abs(ex), 1e-300))
                else:
                    still.append(j)
            bad = np.array(still, np.int64)
        if bad.size:
            rel = err[bad] / np.maximum(np.abs(b[bad]), 1e-300)
            j = bad[int(np.argmax(rel))]
            worst.append(f"{k}: {bad.size} mismatches, worst rel {rel.max():.2e} got {a[j]!r} want {b[j]!r} "
                         f"row {np.flatnonzero(m)[j]} r2 {ref['r2'][m][j]!r} n_used {ref['n_used'].ravel()[0]}")
    assert not worst, "; ".join(worst)


MISSING = 0xFFFFFFFF


def exact_ols(prob, s, y):
    """The regression of snarl s in 80-bit arithmetic (np.longdouble), built independently of the oracle: table
    (quantitative_table.cpp:22-104), merged identical columns and dropped last column (snarl_analyzer.cpp:620-684),
    covariate rows by filtered position (stats_test.cpp:450-452), then beta / se / t of the first dosage column and
    p = 2 T_df(|t|) with the reference's df = n - p + 1.  The arbiter for rows on which CUDA path and oracle differ by more
    than 1e-9: whose value is nearer to this one?"""
    from scipy import stats

    L = np.longdouble
    n_rows, n_cols = prob["n_rows"], prob["n_cols"]
    bits = np.unpackbits(prob["bits"], bitorder="little")[:n_rows * n_cols].reshape(n_rows, n_cols).astype(bool)
    P = prob["OP"]
    spo, peo = np.asarray(P.snarl_path_off, np.int64), np.asarray(P.path_edge_off, np.int64)
    er = np.asarray(P.edge_row, np.int64)
    n = n_cols // 2
    cols = []
    for pth in range(spo[s], spo[s + 1]):
        rows = er[peo[pth]:peo[pth + 1]]
        v = np.zeros(n_cols, bool) if (rows == MISSING).any() else np.logical_and.reduce(bits[rows], axis=0) if rows.size \
            else np.ones(n_cols, bool)
        c = v.reshape(n, 2).sum(1)
        if c.sum() > 0:
            cols.append(c)
    C = np.array(cols).T
    rs = C.sum(1)
    used = rs > 0
    C, rs = C[used], rs[used]
    X = C.astype(L) / rs.astype(L)[:, None]
    k = X.shape[1]
    merged, kept = np.zeros(k, bool), []
    for a in range(k):
        if merged[a]:
            continue
        v = X[:, a].copy()
        for b in range(a + 1, k):
            if not merged[b] and (C[:, b] == C[:, a]).all():
                merged[b] = True
                v = v + X[:, b]
        kept.append(v)
    D = [np.ones(X.shape[0], L)] + kept[:-1]
    if prob["covar"] is not None:
        D += [prob["covar"][:X.shape[0], c].astype(L) for c in range(prob["covar"].shape[1])]
    D = np.array(D).T
    yy = np.asarray(y)[used].astype(L)
    m = D.shape[1]
    M = np.concatenate([D.T @ D, np.eye(m, dtype=L)], 1)  # Gauss-Jordan with partial pivoting -> (X'X)^-1
    for i in range(m):
        piv = i + int(np.argmax(np.abs(M[i:, i])))
        M[[i, piv]] = M[[piv, i]]
        M[i] = M[i] / M[i, i]
        for j in range(m):
            if j != i:
                M[j] = M[j] - M[j, i] * M[i]
    inv = M[:, m:]
    beta = inv @ (D.T @ yy)
    res = yy - D @ beta
    df = max(int(X.shape[0]) - m + 1, 1)
    se = np.sqrt(inv[1, 1] * (res @ res) / df)
    t = float(beta[1] / se)
    return dict(beta=float(beta[1]), se=float(se), p_value=float(2 * stats.t.sf(abs(t), df)))


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
        assert_ols_equal(got, oracle.quantitative_batch(bits, n_cols, prob["y"], OP, covar=covar, **thr),
                         arbiter=lambda s: exact_ols(prob, s, prob["y"]), min_rows=2 + ncov)
    elif trait == "logistic":
        if n > 2600:
            return False
        use_cov = bool(ncov and rng.integers(0, 2))
        ctx.set_binary_phenotype(prob["is_case"])
        ctx.set_covariates(covar)
        got = ctx.binary_covar_batch(P, use_covariates=use_cov, **thr)
        ref = oracle.binary_covar_batch(bits, n_cols, prob["is_case"], OP, covar=covar, use_covariates=use_cov, **thr)
        try:
            assert_logit_equal(got, ref)
        except AssertionError:
            # a column next to separation (se of 15-25: weights at the clamp): the Hessian is ill conditioned and the IRLS
            # iterates carry their rounding noise into se at the 1e-9 level; the widths' test uses 1e-8 for the same reason
            assert_logit_equal(got, ref, rtol=1e-8)
            NOISE["logit_runs_beyond_1e-9_within_1e-8"] += 1
    else:
        expr = prob["expr"]
        if P.n_snarls * expr.shape[0] * n > 3e7:
            return False
        ctx.set_covariates(covar)
        got = ctx.eqtl_batch(P, expr, None, None, **thr)
        G = expr.shape[0]
        assert_ols_equal(got, oracle.eqtl_batch(bits, n_cols, expr, OP, None, None, covar=covar, **thr),
                         arbiter=lambda item: exact_ols(prob, item // G, expr[item % G]), min_rows=2 + ncov)
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
    print(f"fuzz: {n_ok} runs passed {counts}, {len(fails)} failed, seeds {seed0}..{seed - 1}; noise classes {NOISE}")
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
