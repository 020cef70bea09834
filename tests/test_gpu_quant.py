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


@pytest.mark.parametrize("ncov", [1, 3, 5])
@pytest.mark.parametrize("kind,n_samples,max_paths", [("partition", 700, 2), ("partition", 1300, 4), ("partition", 5000, 3),
                                                      ("random", 333, 4), ("synth", 1000, 0)])
def test_covariates_sweep_kernels(gpu_ctx, oracle, monkeypatch, ncov, kind, n_samples, max_paths):
    """K3C / K4C (csrc/quant_covar.cuh): with covariates a snarl of <= 4 paths takes one sweep over its samples (covariate
    rows by FILTERED position, stats_test.cpp:450-452; 1'C and C'C from prefix tables) instead of the general kernel.
    Parity with the oracle, and with the general kernel on the same input (STOAT_GPU_COVAR_FAST=0); shapes the sweep does
    not cover (identical columns, empty paths, > 2 checked edges, thirds next to a threshold) come back through the list."""
    from stoat_b200 import SnarlPaths, synth

    rng = np.random.default_rng(100 + ncov + n_samples)
    if kind == "partition":
        dense, snarls = H.partition_problem(rng, 150 if n_samples < 5000 else 40, n_samples, max_paths=max_paths,
                                            p_missing_hap=0.05)
    elif kind == "random":
        dense, snarls = H.random_snarl_problem(rng, 250, n_samples, 60, max_paths=max_paths, max_edges=3)
    else:
        d = synth.numpy_dataset(600, n_samples, seed=9, frac_three=0.2, frac_node0=0.1, frac_missing_edge=0.02)
    covar = rng.normal(size=(n_samples, ncov)) * rng.uniform(0.5, 20.0, size=ncov) + rng.normal(size=ncov)
    y = rng.normal(size=n_samples) * 2.0 + 7.0 + 0.5 * covar[:, 0]
    if kind == "synth":
        bits, n_rows, n_cols, P = d["bits"], d["n_rows"], d["n_cols"], d["paths"]
        spo, peo, er = P.snarl_path_off, P.path_edge_off, P.edge_row
    else:
        y = y + dense[snarls[0][0][0]].reshape(n_samples, 2).sum(1)
        bits, (n_rows, n_cols) = oracle.pack_matrix(dense), dense.shape
        spo, peo, er = H.csr_from_lists(snarls)
    thr = dict(min_ind=2, min_hap=2, maf=0.01)
    ref = oracle.quantitative_batch(bits, n_cols, y, oracle.Paths(spo, peo, er), covar=covar, **thr)
    gpu_ctx.set_matrix(bits, n_rows, n_cols)
    gpu_ctx.set_quantitative_phenotype(y)
    gpu_ctx.set_covariates(covar)
    try:
        got = gpu_ctx.quantitative_batch(SnarlPaths(spo, peo, er), **thr)
        n_general = gpu_ctx.timing()["n_general"]
        monkeypatch.setenv("STOAT_GPU_COVAR_FAST", "0")
        gpu_ctx.set_covariates(covar)
        slow = gpu_ctx.quantitative_batch(SnarlPaths(spo, peo, er), **thr)
    finally:
        gpu_ctx.set_covariates(None)
    H.assert_quant_equal(got, ref)
    H.assert_quant_equal(slow, ref)
    H.assert_quant_equal(got, slow)
    kept = int(((ref["filtered"] == 0) & (ref["code"] == 0)).sum())
    assert kept > 20
    if kind == "partition" and 1 + (max_paths - 1) + ncov <= 8:
        assert n_general <= max(4, len(spo) // 10), n_general  # the sweep kernels did the work


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


def _two_path_problem(rng, n_snarls, n_samples, frac_missing, overlap_every=0):
    """Two-path snarls with 1 or 2 checked edges per path; some haplotypes on neither path; every `overlap_every`-th
    snarl has haplotypes on BOTH paths (row sums of 3: dosages 1/3, 2/3, which the integer kernel hands on)."""
    Hn = 2 * n_samples
    dense = np.zeros((4 * n_snarls, Hn), dtype=bool)
    snarls = []
    for s in range(n_snarls):
        f = rng.uniform(0.05, 0.95)
        alt = rng.random(Hn) < f
        present = rng.random(Hn) >= frac_missing
        r = 4 * s
        dense[r] = dense[r + 1] = present & ~alt
        dense[r + 2] = dense[r + 3] = present & alt
        if overlap_every and s % overlap_every == 0:
            both = rng.random(Hn) < 0.02
            dense[r] |= both
            dense[r + 1] |= both
            dense[r + 2] |= both
            dense[r + 3] |= both
        e0 = [r, r + 1] if s % 3 else [r]
        e1 = [r + 2, r + 3] if s % 5 else [r + 3]
        snarls.append([e0, e1])
    return dense, snarls


@pytest.mark.parametrize("seed,n_samples,frac_missing,overlap", [(21, 700, 0.01, 0), (22, 1300, 0.45, 0),
                                                                (23, 500, 0.02, 7), (24, 7000, 0.01, 0)])
def test_two_path_integer_contraction(gpu_ctx, oracle, seed, n_samples, frac_missing, overlap):
    """K3M (csrc/quant_mma.cuh): X^T y as an exact int8 MMA over 8 phenotype slices.  Cases: plain; nearly half of the
    haplotypes on neither path (the FP64 side sums over unused samples); overlapping paths (replay on the general
    path); 14,000 haplotypes = 28 rounds > 24, i.e. two haplotype slices whose partial sums K4 adds."""
    rng = np.random.default_rng(seed)
    dense, snarls = _two_path_problem(rng, 96 if n_samples >= 7000 else 330, n_samples, frac_missing, overlap)
    y = rng.normal(size=n_samples) * 2.5 - 40.0
    y = y + 0.8 * dense[snarls[1][0][0]].reshape(n_samples, 2).sum(1)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.01)
    H.assert_quant_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 20


@pytest.mark.parametrize("ring", ["8x2", "11x2", "6x3"])
@pytest.mark.parametrize("n_samples,frac_missing,overlap", [(1300, 0.45, 0), (500, 0.02, 7), (7000, 0.01, 0)])
def test_two_path_contraction_ring_variant(gpu_ctx, oracle, monkeypatch, ring, n_samples, frac_missing, overlap):
    """The cp.async ring variant of K3M (STOAT_GPU_K3M_RING, rows through a per-warp shared-memory ring instead of
    registers) does the register kernel's arithmetic in its order: parity with the oracle AND bit-identical results,
    at one and at two haplotype slices (14,000 haplotypes), odd numbers of rounds included."""
    rng = np.random.default_rng(77)
    dense, snarls = _two_path_problem(rng, 96 if n_samples >= 7000 else 330, n_samples, frac_missing, overlap)
    y = rng.normal(size=n_samples) * 2.5 - 40.0
    y = y + 0.8 * dense[snarls[1][0][0]].reshape(n_samples, 2).sum(1)
    base, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.01)
    monkeypatch.setenv("STOAT_GPU_K3M_RING", ring)
    got, _ = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.01)
    H.assert_quant_equal(got, ref)
    for k in ("p_value", "beta", "se", "r2", "n_used", "filtered", "allele_paths"):
        np.testing.assert_array_equal(got[k], base[k], err_msg=k)


def test_two_path_integer_contraction_phenotype_range(gpu_ctx, oracle):
    """The slice split uses ONE fixed-point scale for the whole phenotype: a heavy-tailed phenotype (values from 1e-7 to
    1e5) must still give 1e-9 parity, and a constant phenotype (all slices zero) the reference's degenerate rows."""
    rng = np.random.default_rng(31)
    n = 900
    dense, snarls = _two_path_problem(rng, 200, n, 0.01)
    y = np.exp(rng.normal(size=n) * 4.0) * np.sign(rng.normal(size=n))
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.01)
    H.assert_quant_equal(got, ref)
    y0 = np.full(n, 3.25)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls[:40], y0, min_ind=2, min_hap=2, maf=0.01)
    np.testing.assert_array_equal(got["filtered"], ref["filtered"])
    np.testing.assert_array_equal(got["allele_paths"], ref["allele_paths"])


def test_sample_sharded_phases_single_gpu(gpu_ctx, oracle):
    """moments | (all-reduce) | solve on ONE GPU (world 1): the phase split, the deferred marking of snarls that need more
    than the two-path sums and their replay (stoat_b200/shard.py) must reproduce the ordinary batch call."""
    from stoat_b200 import shard, synth

    d = synth.numpy_dataset(1500, 448, seed=12, frac_missing_edge=0.02)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    ref = oracle.quantitative_batch(d["bits"], d["n_cols"], d["y"], P)
    got = shard.run_quantitative_sample_sharded(gpu_ctx, d["bits"], d["n_rows"], d["n_cols"], d["paths"], d["y"], 0, 1)
    H.assert_quant_equal(got, ref)
    # the deferred set is what the two-path kernel does not take: here the three-path snarls at least
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_quantitative_phenotype_shard(d["y"], *shard.phenotype_summary(d["y"]))
    gpu_ctx.stage_paths(d["paths"])
    gpu_ctx.quantitative_moments_staged()
    pi, ni, pd, nd = gpu_ctx.quantitative_partials_dev()
    assert ni == 7 * d["paths"].n_snarls and nd == 3 * d["paths"].n_snarls and pi and pd
    gpu_ctx.quantitative_solve_staged()
    res = gpu_ctx.quantitative_fetch()
    todo = shard.deferred_snarls(res)
    n_paths = np.diff(d["paths"].snarl_path_off.astype(np.int64))
    assert set(np.flatnonzero(n_paths != 2)) <= set(todo) and todo.size < d["paths"].n_snarls // 4


@pytest.mark.parametrize("n_shards", [2, 3])
def test_sample_sharded_column_split_on_one_gpu(gpu_ctx, oracle, n_shards):
    """The sample-sharded run with REAL column shards on a one-GPU box: n_shards contexts each hold a column range, the
    all-reduce is a sum over their partial buffers, the deferred snarls (three paths, thirds, missing edges: ~10 %) are
    finished on the device from the shards' gathered row slices (stoat_gpu_gather_rows_dev -> joined rows ->
    stoat_gpu_set_matrix_pitched_dev on a second context).  Same library calls as the NCCL driver."""
    import stoat_b200
    from stoat_b200 import shard, synth

    d = synth.numpy_dataset(3000, 640, seed=21, frac_missing_edge=0.02)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    ref = oracle.quantitative_batch(d["bits"], d["n_cols"], d["y"], P)
    ctxs = [gpu_ctx] + [stoat_b200.Context(0) for _ in range(n_shards - 1)]
    try:
        got = shard.run_quantitative_sample_sharded_local(ctxs, d["bits"], d["n_rows"], d["n_cols"], d["paths"], d["y"])
    finally:
        for c in ctxs[1:]:
            c.close()
    H.assert_quant_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 1000
    # the shard contexts keep their state: shard 0 still holds its column range
    assert gpu_ctx.row_pitch == (shard.plan_column_shards(d["n_cols"], n_shards)[0][1] + 255) // 256 * 32


def _three_path_problem(rng, n_snarls, n_samples, frac_missing):
    """Three-path snarls, 1 or 2 checked edges per path.  Every 11th snarl has a path nobody takes (the reference makes
    no column for it), every 13th two paths over the SAME rows (identical columns, which the reference merges), every
    17th a few haplotypes on two paths at once (row sums of 3)."""
    Hn = 2 * n_samples
    dense = np.zeros((6 * n_snarls, Hn), dtype=bool)
    snarls = []
    for s in range(n_snarls):
        f = rng.dirichlet([2.0, 1.5, 1.0])
        which = rng.choice(3, size=Hn, p=f)
        present = rng.random(Hn) >= frac_missing
        r = 6 * s
        for j in range(3):
            dense[r + 2 * j] = dense[r + 2 * j + 1] = present & (which == j)
        if s % 11 == 5:
            dense[r + 2] = dense[r + 3] = False
        if s % 17 == 3:
            both = rng.random(Hn) < 0.01
            dense[r] |= both
            dense[r + 1] |= both
            dense[r + 4] |= both
            dense[r + 5] |= both
        paths = [[r, r + 1] if s % 3 else [r + 1], [r + 2, r + 3] if s % 5 else [r + 2], [r + 4, r + 5] if s % 2 else [r + 4]]
        if s % 13 == 7:
            paths[1] = list(paths[0])
        snarls.append(paths)
    return dense, snarls


@pytest.mark.parametrize("seed,n_samples,frac_missing", [(41, 650, 0.01), (42, 1500, 0.3), (43, 5000, 0.005),
                                                         (44, 7000, 0.01)])
def test_three_path_integer_contraction(gpu_ctx, oracle, seed, n_samples, frac_missing):
    """K3M3 (csrc/quant_mma.cuh): both regression columns of a three-path snarl from one int8 MMA, X^T X from class
    popcounts.  Empty paths, merged columns and row sums of 3 are handed to the warp-per-snarl kernel; 14,000 haplotypes
    need two haplotype slices, where every three-path snarl takes that kernel."""
    rng = np.random.default_rng(seed)
    dense, snarls = _three_path_problem(rng, 60 if n_samples >= 5000 else 240, n_samples, frac_missing)
    y = rng.normal(size=n_samples) * 1.7 + 12.0
    y = y + 0.6 * dense[snarls[2][0][0]].reshape(n_samples, 2).sum(1) - 0.4 * dense[snarls[2][1][0]].reshape(n_samples, 2).sum(1)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.01)
    H.assert_quant_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 20
    # default thresholds filter some of them: the filter decisions must agree too
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, maf=0.2)
    H.assert_quant_equal(got, ref)


def test_thirds_exact_ties_at_the_maf_threshold(gpu_ctx, oracle):
    """A path without checked edges matches everybody, so a carrier of the other path has dosage 1/3 (one haplotype) or
    1/2 (both).  With n3 / 3 + n2 / 2 = 0.05 * n_samples the allele-frequency filter sits on an exact tie and its outcome
    depends on how the reference's running sums round (src/snarl_analyzer.cpp:558-576): K3M2x replays them in order
    (lean_exact_filtered) instead of sending the snarl to the general path.  The filter flags must be identical,
    whichever way each tie falls (at 800 samples they fall both ways)."""
    rng = np.random.default_rng(77)
    n = 800
    Hn = 2 * n
    combos = [(120, 0), (114, 4), (108, 8), (90, 20), (60, 40), (30, 60), (6, 76), (0, 80)]
    n_snarls = 48
    dense = np.zeros((2 * n_snarls, Hn), dtype=bool)
    snarls = []
    for s in range(n_snarls):
        n3, n2 = combos[s % len(combos)]
        if s >= 32:  # off the tie by one carrier, either side
            n3 = max(n3 + (1 if s % 2 else -1), 0)
        who = rng.permutation(n)
        one, both = who[:n3], who[n3:n3 + n2]
        row = np.zeros(Hn, dtype=bool)
        row[2 * one + rng.integers(0, 2, size=n3)] = True
        row[2 * both] = True
        row[2 * both + 1] = True
        dense[2 * s] = dense[2 * s + 1] = row
        snarls.append([[2 * s, 2 * s + 1] if s % 2 else [2 * s], []])
    y = rng.normal(size=n)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, y, min_ind=2, min_hap=2, maf=0.05)
    np.testing.assert_array_equal(got["filtered"], ref["filtered"])
    H.assert_quant_equal(got, ref)
    assert 4 < int(ref["filtered"][:32].sum()) < 28  # the ties fall both ways
    assert gpu_ctx.timing()["n_general"] == 0        # ... and none of them needed the general path
