"""Parity of the CUDA logistic path (L1 logit_class_counts, L2 logit_irls, the general kernel's logistic
mode) with the oracle, and with rows written by the reference's own BinaryCovarSnarlAnalyzer (oracle/_ref)."""
import os

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def assert_logit_equal(got: dict, ref: dict, rtol=1e-9):
    """Integers and codes bit-exact; beta / se within 1e-9 relative (+ 1e-12 for values that are ~0); p = 2 (1 - cdf) is a
    multiple of 2^-52 in the reference, so it carries that quantum on top of the relative bound."""
    for k in ("allele_paths", "filtered", "n_used"):
        np.testing.assert_array_equal(got[k], ref[k], err_msg=k)
    keep = ref["filtered"] == 0
    np.testing.assert_array_equal(got["code"][keep], ref["code"][keep], err_msg="code")
    m = keep & (ref["code"] == 0)
    atols = dict(p_value=2.0 ** -51, beta=1e-12, se=1e-300)
    for k in ("p_value", "beta", "se"):
        a, b = got[k][m], ref[k][m]
        ok = (np.isnan(a) & np.isnan(b)) | (np.abs(a - b) <= rtol * np.abs(b) + atols[k])
        bad = np.flatnonzero(~ok)
        assert bad.size == 0, f"{k}: {bad.size} mismatches, first got {a[bad[:3]]} want {b[bad[:3]]} idx {np.flatnonzero(m)[bad[:3]]}"
    return int(m.sum())


def run_both(ctx, oracle, dense, snarls, is_case, covar=None, use_covariates=False, **thr):
    from stoat_b200 import SnarlPaths

    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    ctx.set_matrix(bits, n_rows, n_cols)
    ctx.set_binary_phenotype(is_case)
    ctx.set_covariates(covar)
    got = ctx.binary_covar_batch(SnarlPaths(spo, peo, er), use_covariates=use_covariates, **thr)
    ref = oracle.binary_covar_batch(bits, n_cols, is_case, oracle.Paths(spo, peo, er), covar=covar,
                                    use_covariates=use_covariates, **thr)
    ctx.set_covariates(None)
    return got, ref


def planted_phenotype(rng, dense, snarls, n_samples, effect=1.5):
    d = dense[snarls[0][0][0]].reshape(n_samples, 2).sum(1) / 2.0
    return (rng.random(n_samples) < 1.0 / (1.0 + np.exp(-(effect * d - 0.6)))).astype(np.uint8)


@pytest.mark.parametrize("seed,n_samples", [(1, 60), (2, 200), (3, 1001), (4, 5000)])
def test_two_path_class_counts(gpu_ctx, oracle, seed, n_samples):
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, 300 if n_samples < 5000 else 60, n_samples, max_paths=2)
    is_case = planted_phenotype(rng, dense, snarls, n_samples)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case)
    assert assert_logit_equal(got, ref) > 10
    assert gpu_ctx.timing()["n_general"] == 0


@pytest.mark.parametrize("seed,n_samples,p_missing", [(21, 90, 0.05), (22, 700, 0.0), (23, 2500, 0.3), (24, 5000, 0.01)])
def test_three_path_class_counts(gpu_ctx, oracle, monkeypatch, seed, n_samples, p_missing):
    """L3 (logit_three_path_kernel): a three-path snarl whose samples sit on one or two path-haplotypes is a regression over
    SIX weighted points (two dosage columns, Holm over their Wald tests).  Against the oracle and against the general
    kernel on the same snarls; two-path snarls in the same batch keep their own kernel."""
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, 200 if n_samples < 5000 else 60, n_samples, max_paths=3, p_missing_hap=p_missing)
    is_case = planted_phenotype(rng, dense, snarls, n_samples)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=2, min_hap=2, maf=0.01)
    assert assert_logit_equal(got, ref) > 20
    assert gpu_ctx.timing()["n_general"] <= len(snarls) // 20  # (an empty or a repeated column sends a snarl there)
    assert sum(len(s_) == 3 for s_ in snarls) > 20
    monkeypatch.setenv("STOAT_GPU_LOGIT", "general")
    slow, _ = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=2, min_hap=2, maf=0.01)
    assert_logit_equal(got, slow)


@pytest.mark.parametrize("seed,n_samples", [(5, 80), (6, 700)])
def test_many_paths_holm_general_kernel(gpu_ctx, oracle, seed, n_samples):
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, 200, n_samples, max_paths=6)
    is_case = planted_phenotype(rng, dense, snarls, n_samples)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=3, min_hap=5, maf=0.01)
    assert assert_logit_equal(got, ref) > 10
    assert gpu_ctx.timing()["n_general"] > 0


@pytest.mark.parametrize("seed,n_samples,n_rows", [(7, 30, 16), (8, 150, 40), (9, 700, 80)])
def test_random_problems_thirds_missing_edges(gpu_ctx, oracle, seed, n_samples, n_rows):
    rng = np.random.default_rng(seed)
    dense, snarls = H.random_snarl_problem(rng, 250, n_samples, n_rows, max_paths=4)
    is_case = (rng.random(n_samples) < 0.45).astype(np.uint8)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=2, min_hap=2, maf=0.01)
    assert_logit_equal(got, ref)


def test_separation_and_constant_phenotype(gpu_ctx, oracle):
    """Perfect separation (betas held only by the L2 term, weights clamped at 1e-8) and an all-case phenotype."""
    rng = np.random.default_rng(12)
    n = 400
    dense, snarls = H.partition_problem(rng, 40, n, max_paths=2, p_missing_hap=0.0)
    carrier = dense[snarls[0][0][0]].reshape(n, 2).any(1)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, carrier.astype(np.uint8))
    assert_logit_equal(got, ref)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, np.ones(n, np.uint8))
    assert_logit_equal(got, ref)


def test_class_count_path_equals_general_kernel(gpu_ctx, oracle, monkeypatch):
    rng = np.random.default_rng(13)
    n = 900
    dense, snarls = H.partition_problem(rng, 150, n, max_paths=2)
    is_case = planted_phenotype(rng, dense, snarls, n)
    fast, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case)
    monkeypatch.setenv("STOAT_GPU_LOGIT", "general")
    slow, _ = run_both(gpu_ctx, oracle, dense, snarls, is_case)
    assert gpu_ctx.timing()["n_general"] == 0  # the list is not used when every snarl takes the general kernel
    assert_logit_equal(slow, ref)
    assert_logit_equal(fast, slow)


def test_covariates_honoured_flag(gpu_ctx, oracle):
    """use_covariates=1: the design the reference meant (no reference output exists; oracle restatement only)."""
    rng = np.random.default_rng(14)
    n = 300
    dense, snarls = H.partition_problem(rng, 80, n, max_paths=3)
    covar = rng.normal(size=(n, 3))
    is_case = (rng.random(n) < 1.0 / (1.0 + np.exp(-covar[:, 0]))).astype(np.uint8)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, covar=covar, use_covariates=True)
    assert assert_logit_equal(got, ref) > 10
    ign, ref0 = run_both(gpu_ctx, oracle, dense, snarls, is_case, covar=covar, use_covariates=False)
    assert_logit_equal(ign, ref0)
    m = (ref["filtered"] == 0) & (ref["code"] == 0)
    assert not np.allclose(ref["beta"][m], ref0["beta"][m])


def test_against_reference_analyzer_rows(gpu_ctx, oracle):
    """Rows written by the reference's own BinaryCovarSnarlAnalyzer + writer (oracle/_ref), keyed by SNARL."""
    ref = pytest.importorskip("oracle.ref")
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    from stoat_b200 import SnarlPaths, format_value, synth

    d = synth.numpy_dataset(1500, 200, seed=8, frac_missing_edge=0.02)
    P = d["paths"]
    nodes = ref.nodes_from_structure(d["structure"])
    covar = np.random.default_rng(5).normal(size=(200, 4))
    out = ref.run("binary_covar", d["bits"], d["n_rows"], d["n_cols"], P.snarl_path_off, nodes, is_case=d["is_case"],
                  covar=covar, n_threads=4)
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_binary_phenotype(d["is_case"])
    gpu_ctx.set_covariates(covar)
    res = gpu_ctx.binary_covar_batch(SnarlPaths(P.snarl_path_off, P.path_edge_off, P.edge_row))
    gpu_ctx.set_covariates(None)
    spo = np.asarray(P.snarl_path_off, np.int64)
    kept = np.flatnonzero((res["filtered"] == 0) & (res["code"] != 5))
    undef = set(np.flatnonzero((res["filtered"] == 0) & (res["code"] == 5)).tolist())
    assert sorted(set(out["rows"]) - undef) == list(kept), "filter decisions differ"
    assert kept.size > 800
    for s in kept:
        (row,) = out["rows"][int(s)]
        assert row["ALLELE_PATHS"] == ",".join(str(v) for v in res["allele_paths"][spo[s]:spo[s + 1]]), (s, row)
        for col, key in (("P", "p_value"), ("BETA", "beta"), ("SE", "se")):
            want = "NA" if res["code"][s] == 1 else format_value(res[key][s], 0)
            if col == "BETA" and want != "NA" and abs(float(row[col])) < 1e-12 and abs(res[key][s]) < 1e-12:
                continue  # a coefficient that is mathematically zero: rounding noise on both sides
            assert H.printed_close(row[col], want), (s, col, row, want)
