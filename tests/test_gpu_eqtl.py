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


# ---------------------------------------------------------------------------------------------------
# the dense scan on the FP64 tensor cores (kernels_eqtl.cu) — two-path snarls, any number of tiles / gene groups
# ---------------------------------------------------------------------------------------------------
def _check_summary(got, ref, S, G, thr, rtol=1e-9, skip=()):
    p = ref["p_value"].reshape(S, G)
    ok = ref["code"].reshape(S, G) == 0
    n_checked = 0
    for s in range(S):
        if s in skip:
            continue
        if ref["filtered"][s] or not ok[s].any():
            assert np.isnan(got["min_p"][s]) and got["n_below"][s] == 0, s
            continue
        ps = np.where(ok[s], p[s], np.inf)
        g = int(np.argmin(ps))
        assert abs(got["min_p"][s] - ps[g]) <= rtol * ps[g] + 1e-300, (s, got["min_p"][s], ps[g])
        if got["argmin_gene"][s] != g:  # a tie within rounding
            assert abs(ps[got["argmin_gene"][s]] - ps[g]) <= rtol * ps[g], (s, got["argmin_gene"][s], g)
        near = np.abs(ps - thr) <= rtol * thr
        lo, hi = int(((ps < thr) & ~near).sum()), int(((ps < thr) | near).sum())
        assert lo <= got["n_below"][s] <= hi, (s, got["n_below"][s], lo, hi)
        n_checked += 1
    return n_checked


def _dense_problem(seed, n_snarls, n_samples, n_genes, **kw):
    from stoat_b200 import synth

    d = synth.numpy_dataset(n_snarls, n_samples, seed=seed, **kw)
    rng = np.random.default_rng(seed + 50)
    base = rng.uniform(0, 8, size=(n_genes, 1))
    expr = base * (1.0 + rng.uniform(-1, 1, size=(n_genes, n_samples)))
    dense = d["dense"]
    st = d["structure"]
    for g in range(0, n_genes, 4):  # genes driven by a snarl's alt dosage (simulate_test_dataset.py:75-88)
        s = int(rng.integers(0, n_snarls))
        if st.kind[s] == 2:
            continue
        expr[g] += rng.uniform(0.5, 3.0) * dense[st.row0[s] + 2].reshape(n_samples, 2).sum(1)
    return d, expr


@pytest.mark.parametrize("seed,n_snarls,n_samples,n_genes,group", [
    (11, 300, 100, 70, None),     # 3 snarl tiles, 2 gene tiles (the second one ragged)
    (12, 130, 517, 200, "64"),    # ragged sample count (9 stages), 4 gene groups of one tile
    (13, 64, 33, 5, None),        # less than one tile in every dimension
])
def test_dense_dmma_summary(oracle, monkeypatch, seed, n_snarls, n_samples, n_genes, group):
    import stoat_b200

    if group:
        monkeypatch.setenv("STOAT_GPU_EQTL_GROUP", group)
    d, expr = _dense_problem(seed, n_snarls, n_samples, n_genes, frac_missing_edge=0.02, frac_missing_hap=0.08)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    thr = 1e-3
    ref = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P)
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
        got = ctx.eqtl_dense_summary(d["paths"], expr, p_threshold=thr)
        t = ctx.timing()
        # resident mode gives the same answer
        ctx.stage_paths(d["paths"])
        ctx.set_expression(expr)
        ctx.eqtl_dense_run_staged(p_threshold=thr)
        again = ctx.eqtl_summary_fetch()
    assert _check_summary(got, ref, P.n_snarls, n_genes, thr) > n_snarls // 3
    for k in ("argmin_gene", "n_below"):
        np.testing.assert_array_equal(got[k], again[k])
    np.testing.assert_array_equal(np.isnan(got["min_p"]), np.isnan(again["min_p"]))
    # most snarls of the generator are two-path and dyadic: they must have taken the tensor-core path
    assert t["n_general"] < 0.2 * n_snarls + 8, t
    # and the general kernel alone (STOAT_GPU_EQTL=general) agrees
    monkeypatch.setenv("STOAT_GPU_EQTL", "general")
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
        gen = ctx.eqtl_dense_summary(d["paths"], expr, p_threshold=thr)
        assert ctx.timing()["n_general"] == n_snarls
    _check_summary(gen, ref, P.n_snarls, n_genes, thr)


def test_dense_dmma_degenerate_genes(oracle):
    """Near-perfect fits (explicit-residual branch), a gene that is an exact linear function of a dosage (rss = 0),
    thresholds above 1.  (A constant gene is not compared: the reference's uncentred statistic is then 0/0 rounding
    noise — it prints p ~ 1e-14 for pure noise — while the centred computation here yields an invalid t, p = 1.)"""
    import stoat_b200

    d, expr = _dense_problem(21, 140, 90, 9, frac_three=0.0, frac_node0=0.0, frac_missing_hap=0.03)
    st, dense, n = d["structure"], d["dense"], 90
    dos = dense[st.row0[3]].reshape(n, 2).sum(1) / 2.0
    dos5 = dense[st.row0[5]].reshape(n, 2).sum(1) / 2.0
    expr[0] = 2.0 + 3.0 * dos                                                # exact linear function of snarl 3's dosage
    expr[1] = 1.0 - 2.0 * dos5 + 1e-7 * np.random.default_rng(1).normal(size=n)  # near perfect for snarl 5
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    ref = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P)
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
        for thr in (1e-4, 0.5, 2.0):
            got = ctx.eqtl_dense_summary(d["paths"], expr, p_threshold=thr)
            # the exactly-linear gene has rss ~ 1e-30 in any implementation: compare p only where it is not rounding noise
            # snarl 3 x gene 0 has rss ~ 1e-30 in any implementation: its p is rounding noise, only its rank is checked
            _check_summary(got, ref, P.n_snarls, 9, thr, rtol=1e-6, skip={3})
            if not ref["filtered"][3]:
                assert got["argmin_gene"][3] == 0 and got["min_p"][3] < 1e-30
            if not ref["filtered"][5]:
                assert got["argmin_gene"][5] == 1


def test_dense_with_covariates_takes_general_path(gpu_ctx, oracle):
    d, expr = _dense_problem(31, 60, 80, 6)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    covar = np.random.default_rng(2).normal(size=(80, 2))
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_covariates(covar)
    got = gpu_ctx.eqtl_dense_summary(d["paths"], expr, p_threshold=1e-2)
    assert gpu_ctx.timing()["n_general"] == 60
    gpu_ctx.set_covariates(None)
    ref = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P, covar=covar)
    _check_summary(got, ref, P.n_snarls, 6, 1e-2)


@pytest.mark.parametrize("ncov", [1, 3, 5])
@pytest.mark.parametrize("kind,n_samples,n_genes,max_paths", [("partition", 333, 13, 2), ("partition", 1200, 70, 4),
                                                              ("random", 150, 9, 4), ("windows", 700, 30, 3)])
def test_covariates_sweep_kernels(gpu_ctx, oracle, monkeypatch, ncov, kind, n_samples, n_genes, max_paths):
    """E3C (csrc/quant_covar.cuh): with covariates a snarl of <= 4 paths takes one sweep for its design and one per batch of
    genes instead of the general kernel's three sweeps per gene.  Every pair against the oracle and against the general
    kernel on the same input (STOAT_GPU_COVAR_FAST=0); gene windows (CSR) and dense; more genes than one batch; shapes
    the sweep does not take (random problems: empty paths, identical columns, > 2 edges) come back through the flags."""
    rng = np.random.default_rng(7 * ncov + n_samples)
    if kind == "random":
        dense, snarls = H.random_snarl_problem(rng, 120, n_samples, 40, max_paths=max_paths, max_edges=3)
    else:
        dense, snarls = H.partition_problem(rng, 60, n_samples, max_paths=max_paths, p_missing_hap=0.05)
    base = rng.uniform(0, 8, size=(n_genes, 1))
    expr = base * (1.0 + rng.uniform(-1, 1, size=(n_genes, n_samples)))
    for g in range(0, n_genes, 4):
        s = int(rng.integers(0, len(snarls)))
        if snarls[s][0]:
            expr[g] += 1.5 * dense[snarls[s][0][0]].reshape(n_samples, 2).sum(1)
    covar = rng.normal(size=(n_samples, ncov)) * rng.uniform(0.5, 20.0, size=ncov) + rng.normal(size=ncov)
    expr[1] += 0.3 * covar[:, 0]
    off = idx = None
    if kind == "windows":
        counts = rng.integers(0, 12, size=len(snarls))
        off = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint64)
        idx = np.concatenate([rng.choice(n_genes, size=c, replace=False) for c in counts] + [np.zeros(0, int)]).astype(np.uint32)
    thr = dict(min_ind=2, min_hap=2, maf=0.01)
    got, ref = _run(gpu_ctx, oracle, dense, snarls, expr, off, idx, covar=covar, **thr)
    monkeypatch.setenv("STOAT_GPU_COVAR_FAST", "0")
    slow, _ = _run(gpu_ctx, oracle, dense, snarls, expr, off, idx, covar=covar, **thr)
    # a coefficient that is small against its standard error comes out of a cancellation (inv(X'X) X'y over terms of the
    # size of sum y): it is compared with 1e-9 |beta| + 1e-10 se — its uncertainty is se, its rounding noise ~1e-12 se
    # (scripts/fuzz_parity.py arbitrates such rows with an 80-bit regression: the CUDA value is the nearer one)
    for a_, b_ in ((slow, ref), (got, ref), (got, slow)):
        m = b_["code"] == 0
        fix = m & (np.abs(a_["beta"] - b_["beta"]) <= 1e-9 * np.abs(b_["beta"]) + 1e-10 * b_["se"])
        a2 = dict(a_)
        a2["beta"] = np.where(fix, b_["beta"], a_["beta"])
        _assert_pairs_equal(a2, b_)
    assert (ref["code"] == 0).sum() > 50
