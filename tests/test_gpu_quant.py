"""Parity of the CUDA quantitative path (K3 dosage_moments, K4 ols_solve, general path) with the oracle."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def run_both(ctx, oracle, dense, snarls, y, covar=None, **thr):
    from stoat_b200 import SnarlPaths

    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    ctx.set_matrix(bits, n_rows, n_cols)
    ctx.set_quantitative_phenotype(y)
    ctx.set_covariates(covar)
    got = ctx.quantitative_batch(SnarlPaths(spo, peo, er), **thr)
    ref = oracle.quantitative_batch(bits, n_cols, y, oracle.Paths(spo, peo, er), covar=covar, **thr)
    ctx.set_covariates(None)
    return got, ref


@pytest.mark.parametrize("seed,n_samples", [(1, 60), (2, 200), (3, 1001), (4, 5000)])
def test_partition_problems_fast_path(gpu_ctx, oracle, seed, n_samples):
    # every haplotype on <= 1 path: dosages in {0, 1/2, 1} -> the popcount fast path (<= 4 paths)
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, 300 if n_samples < 5000 else 50, n_samples, max_paths=4)
    y = rng.normal(size=n_samples) * 3 + 10
    # plant an effect on the first snarl
    y = y + dense[snarls[0][0][0]].reshape(n_samples, 2).sum(1)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y)
    H.assert_quant_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 10


@pytest.mark.parametrize("seed,n_samples", [(5, 40), (6, 333)])
def test_many_paths_general_path(gpu_ctx, oracle, seed, n_samples):
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, 150, n_samples, max_paths=9)
    y = rng.normal(size=n_samples)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=3, min_hap=5, maf=0.01)
    H.assert_quant_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 10


@pytest.mark.parametrize("seed,n_samples,n_rows", [(7, 30, 16), (8, 150, 40), (9, 700, 80)])
def test_random_problems_nondyadic(gpu_ctx, oracle, seed, n_samples, n_rows):
    # haplotypes matching several paths: row sums 3, 4, ... (thirds), shared rows, match-all and missing paths
    rng = np.random.default_rng(seed)
    dense, snarls = H.random_snarl_problem(rng, 250, n_samples, n_rows, max_paths=5)
    y = rng.normal(size=n_samples)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.01)
    H.assert_quant_equal(got, ref)


def test_covariates_and_position_quirk(gpu_ctx, oracle):
    """Covariate rows are taken by position in the FILTERED table (stats_test.cpp:450-452): with dropped samples
    they are misaligned; the device reproduces that.  C2' shape: 200 samples, covariates of the repo data set."""
    rng = np.random.default_rng(10)
    y = H.golden_quant_phenotype()
    cov = H.golden_quant_covariates()
    n = y.size
    assert n == 200 and cov.shape == (200, 4)
    dense, snarls = H.partition_problem(rng, 300, n, max_paths=3, p_missing_hap=0.15)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, covar=cov)
    H.assert_quant_equal(got, ref)
    assert (ref["n_used"][ref["filtered"] == 0] < n).any(), "the quirk is only visible when samples are dropped"
    got2, ref2 = run_both(gpu_ctx, oracle, dense, snarls, y, covar=cov[:, :1])
    H.assert_quant_equal(got2, ref2)


def test_merges_undefined_and_rank_deficient(gpu_ctx, oracle):
    n = 40
    H2 = 2 * n
    rng = np.random.default_rng(12)
    a = rng.random(H2) < 0.4
    b = ~a
    het = np.zeros(H2, bool)
    het[0::2] = True          # every sample heterozygous: a constant dosage column (collinear with intercept)
    dense = np.array([a, b, a.copy(), het, ~het, np.ones(H2, bool)])
    snarls = [
        [[0], [2], [1]],      # columns 0 and 1 identical -> merged (summed), 3 columns -> 2 -> 1 regression column
        [[0], [2]],           # two identical columns, no merge with < 3 columns
        [[0], [2], [0]],      # all three identical: merge leaves one column, drop-last leaves none -> UNDEF
        [[3], [4]],           # constant 1/2 column -> LDLT pivot ~ 0 -> pseudo-inverse branch
        [[5], [0]],           # match-all row + a: row sums 1 and 2
        [[0], [1], [3]],      # haplotypes on two paths: row sums up to 3 -> thirds
    ]
    y = rng.normal(size=n) + a.reshape(n, 2).sum(1)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=1, min_hap=1, maf=0.0)
    H.assert_quant_equal(got, ref)
    assert ref["code"][2] == oracle.P_UNDEF and ref["filtered"][2] == 0


def test_near_perfect_fit_uses_residual_pass(gpu_ctx, oracle):
    # R^2 -> 1: the moment expansion would cancel; such snarls are re-done with explicit residuals
    n = 500
    rng = np.random.default_rng(13)
    a = rng.random(2 * n) < 0.3
    dense = np.array([a, ~a])
    dos = a.reshape(n, 2).sum(1) / 2.0
    y = 3.0 + 2.0 * dos + rng.normal(size=n) * 1e-4
    got, ref = run_both(gpu_ctx, oracle, dense, [[[0], [1]]], y)
    assert ref["r2"][0] > 1 - 1e-6
    H.assert_quant_equal(got, ref)


def test_synthetic_generator_and_staged_mode(gpu_ctx, oracle):
    from stoat_b200 import synth

    d = synth.numpy_dataset(1200, 400, seed=9, frac_missing_edge=0.02)
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_quantitative_phenotype(d["y"])
    got = gpu_ctx.quantitative_batch(d["paths"])
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    ref = oracle.quantitative_batch(d["bits"], d["n_cols"], d["y"], P)
    H.assert_quant_equal(got, ref)
    gpu_ctx.stage_paths(d["paths"])
    gpu_ctx.quantitative_run_staged()
    gpu_ctx.quantitative_run_staged()  # re-running a staged batch must be idempotent
    again = gpu_ctx.quantitative_fetch()
    for k in ("allele_paths", "filtered", "n_used", "code"):
        np.testing.assert_array_equal(got[k], again[k], err_msg=k)
    for k in ("p_value", "beta", "se", "r2"):
        np.testing.assert_array_equal(np.nan_to_num(got[k], nan=-7.0), np.nan_to_num(again[k], nan=-7.0), err_msg=k)
