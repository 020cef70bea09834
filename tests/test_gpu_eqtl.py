"""Parity of the CUDA eQTL path (one OLS per (snarl, gene of its window), EQTLSnarlAnalyzer::analyze_and_write_snarl,
src/snarl_analyzer.cpp:493-539) with the oracle."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _assert_pairs_equal(got, ref, rtol=1e-9):
    for k in ("allele_paths", "filtered", "n_used", "code"):
        np.testing.assert_array_equal(got[k], ref[k], err_msg=k)
    m = ref["code"] == 0
    atols = dict(p_value=1e-300, r2=1e-12, beta=1e-12, se=1e-300)
    for k in ("p_value", "r2", "beta", "se"):
        a, b = got[k][m], ref[k][m]
        ok = (np.isnan(a) & np.isnan(b)) | (np.abs(a - b) <= rtol * np.abs(b) + atols[k])
        bad = np.flatnonzero(~ok)
        assert bad.size == 0, f"{k}: {bad.size} mismatches, first got {a[bad[:3]]} want {b[bad[:3]]}"


def _problem(seed, n_snarls, n_samples, n_genes, max_paths=4):
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, n_snarls, n_samples, max_paths=max_paths)
    # expression as tests/simu/scripts/simulate_test_dataset.py:75-88: e_g * (1 + U(-1, 1)), some genes driven by a snarl
    base = rng.uniform(0, 8, size=(n_genes, 1))
    expr = base * (1.0 + rng.uniform(-1, 1, size=(n_genes, n_samples)))
    for g in range(0, n_genes, 5):
        s = int(rng.integers(0, n_snarls))
        expr[g] += 1.5 * dense[snarls[s][0][0]].reshape(n_samples, 2).sum(1)
    return dense, snarls, expr


def _run(ctx, oracle, dense, snarls, expr, off=None, idx=None, covar=None, **thr):
    from stoat_b200 import SnarlPaths

    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    ctx.set_matrix(bits, n_rows, n_cols)
    ctx.set_covariates(covar)
    got = ctx.eqtl_batch(SnarlPaths(spo, peo, er), expr, off, idx, **thr)
    ctx.set_covariates(None)
    ref = oracle.eqtl_batch(bits, n_cols, expr, oracle.Paths(spo, peo, er), off, idx, covar=covar, **thr)
    return got, ref


@pytest.mark.parametrize("seed,n_samples,n_genes", [(1, 80, 7), (2, 333, 12), (3, 1200, 5)])
def test_dense_pairs(gpu_ctx, oracle, seed, n_samples, n_genes):
    dense, snarls, expr = _problem(seed, 60, n_samples, n_genes)
    got, ref = _run(gpu_ctx, oracle, dense, snarls, expr)
    _assert_pairs_equal(got, ref)
    assert (ref["code"] == 0).sum() > 50


def test_gene_windows_csr(gpu_ctx, oracle):
    dense, snarls, expr = _problem(4, 90, 200, 20)
    rng = np.random.default_rng(4)
    counts = rng.integers(0, 6, size=len(snarls))  # some snarls have no gene in their window
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint64)
    idx = np.concatenate([rng.choice(20, size=c, replace=False) for c in counts] + [np.zeros(0, int)]).astype(np.uint32)
    got, ref = _run(gpu_ctx, oracle, dense, snarls, expr, off, idx)
    _assert_pairs_equal(got, ref)


def test_covariates_nondyadic_and_many_paths(gpu_ctx, oracle):
    rng = np.random.default_rng(5)
    dense, snarls = H.random_snarl_problem(rng, 80, 150, 40, max_paths=5)
    expr = rng.normal(size=(6, 150)) + 5
    covar = rng.normal(size=(150, 3))
    got, ref = _run(gpu_ctx, oracle, dense, snarls, expr, covar=covar, min_ind=2, min_hap=2, maf=0.01)
    _assert_pairs_equal(got, ref)


def test_dense_summary_matches_reduced_pairs(gpu_ctx, oracle):
    from stoat_b200 import SnarlPaths

    dense, snarls, expr = _problem(6, 70, 240, 16)
    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    gpu_ctx.set_matrix(bits, n_rows, n_cols)
    thr = 1e-2
    got = gpu_ctx.eqtl_dense_summary(SnarlPaths(spo, peo, er), expr, p_threshold=thr)
    ref = oracle.eqtl_batch(bits, n_cols, expr, oracle.Paths(spo, peo, er))
    S, G = len(snarls), expr.shape[0]
    p = ref["p_value"].reshape(S, G)
    ok = (ref["code"].reshape(S, G) == 0)
    for s in range(S):
        if ref["filtered"][s] or not ok[s].any():
            assert np.isnan(got["min_p"][s]) and got["n_below"][s] == 0
            continue
        ps = np.where(ok[s], p[s], np.inf)
        g = int(np.argmin(ps))
        assert abs(got["min_p"][s] - ps[g]) <= 1e-9 * ps[g] + 1e-300
        if got["argmin_gene"][s] != g:  # a tie within rounding
            assert abs(ps[got["argmin_gene"][s]] - ps[g]) <= 1e-9 * ps[g]
        near = np.abs(ps - thr) <= 1e-9 * thr
        lo, hi = int(((ps < thr) & ~near).sum()), int(((ps < thr) | near).sum())
        assert lo <= got["n_below"][s] <= hi
