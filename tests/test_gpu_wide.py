"""Parity at the widths and batch sizes config C5 (10M snarls x 50,000 haplotypes) reaches and C3 does not:

  * K1's column-chunked stages (rows wider than 2 KB; forced at small widths with STOAT_GPU_K1_CHUNK), including the
    hand-over of per-path counts between the chunks of a non-biallelic snarl;
  * 50,000 haplotypes: Fisher tables of ~850 recurrence steps, K3M's haplotype slices (gridDim.y > 1);
  * >= 65,536 haplotypes: the reductions that no longer pack two counts into one 32-bit word;
  * >= 65,536 snarls: the overlapped run (count pass of chunk c + 1 beside the Fisher list of chunk c).

Same bars as everywhere: tables / flags bit-exact, p-values 1e-9 relative (Fisher additionally 1e-12)."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def binary_both(ctx, oracle, dense, snarls, is_case, **thr):
    from stoat_b200 import SnarlPaths

    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    ctx.set_matrix(bits, n_rows, n_cols)
    ctx.set_binary_phenotype(is_case)
    got = ctx.binary_batch(SnarlPaths(spo, peo, er), **thr)
    ref = oracle.binary_batch(bits, n_cols, is_case, oracle.Paths(spo, peo, er), **thr)
    return got, ref


def quant_both(ctx, oracle, dense, snarls, y, **thr):
    from stoat_b200 import SnarlPaths

    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    ctx.set_matrix(bits, n_rows, n_cols)
    ctx.set_quantitative_phenotype(y)
    ctx.set_covariates(None)
    got = ctx.quantitative_batch(SnarlPaths(spo, peo, er), **thr)
    ref = oracle.quantitative_batch(bits, n_cols, y, oracle.Paths(spo, peo, er), **thr)
    return got, ref


@pytest.mark.parametrize("chunk", [256, 272, 512])
def test_column_chunks_forced_at_small_width(gpu_ctx, oracle, monkeypatch, chunk):
    """Every snarl shape through the chunked K1 (2 - 5 chunks per row at these widths): adversarial lists (shared rows,
    match-all and missing paths, > 4 checked edges) and partitions with up to 6 distinct rows."""
    monkeypatch.setenv("STOAT_GPU_K1_CHUNK", str(chunk))
    rng = np.random.default_rng(chunk)
    for n_samples, n_rows in ((1300, 60), (2500, 90), (4100, 40)):
        dense, snarls = H.random_snarl_problem(rng, 200, n_samples, n_rows)
        is_case = (rng.random(n_samples) < 0.45).astype(np.uint8)
        got, ref = binary_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=2, min_hap=3, maf=0.01)
        H.assert_binary_equal(got, ref)
    dense, snarls = H.partition_problem(rng, 150, 3000, max_paths=3)
    is_case = (rng.random(3000) < 0.5).astype(np.uint8)
    got, ref = binary_both(gpu_ctx, oracle, dense, snarls, is_case)
    H.assert_binary_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 10


@pytest.mark.parametrize("n_samples", [25000, 33000])
def test_generator_at_c5_width(gpu_ctx, oracle, n_samples):
    """The bench generator at 50,000 haplotypes (C5; 6,272-byte rows, four column chunks, packed reductions) and at
    66,000 (8,256-byte rows, unpacked reductions), binary and quantitative, against the oracle."""
    from stoat_b200 import synth

    d = synth.numpy_dataset(240, n_samples, seed=9, frac_missing_edge=0.03)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_binary_phenotype(d["is_case"])
    got = gpu_ctx.binary_batch(d["paths"])
    ref = oracle.binary_batch(d["bits"], d["n_cols"], d["is_case"], P)
    H.assert_binary_equal(got, ref)
    m = (ref["filtered"] == 0) & (ref["fisher_code"] == 0)
    assert m.sum() > 100
    np.testing.assert_allclose(got["p_fisher"][m], ref["p_fisher"][m], rtol=1e-12, atol=0)
    gpu_ctx.set_quantitative_phenotype(d["y"])
    gpu_ctx.set_covariates(None)
    gq = gpu_ctx.quantitative_batch(d["paths"])
    rq = oracle.quantitative_batch(d["bits"], d["n_cols"], d["y"], P)
    H.assert_quant_equal(gq, rq)
    assert (rq["filtered"] == 0).sum() > 100
    if n_samples == 25000:
        # three-path and thirds snarls stay on the integer contraction at this width (slice table read from global
        # memory); before, all ~7 % of them fell to the general kernel, 212 of the 220 ms of a C5 quantitative step
        assert gpu_ctx.timing()["n_general"] <= 12  # missing-edge and four-path snarls of the generator


def test_long_fisher_tables_at_c5_width(gpu_ctx, oracle, monkeypatch):
    """2x2 tables over 50,000 haplotypes: ~850 steps of the recurrence per table, up to several thousand for the skewed
    ones.  Default quotient (N * rcp(D)) within 1e-12 of the oracle; STOAT_GPU_FISHER=exact bit-identical."""
    rng = np.random.default_rng(50)
    n = 25000
    tables = []
    for i in range(60):
        f = rng.uniform(0.02, 0.98)
        shift = rng.choice([0.0, 0.0, 0.01, 0.03])
        a = rng.binomial(n, f)
        b = rng.binomial(n, min(f + shift, 0.995))
        tables.append(([n - a, a], [n - b, b]))
    is_case = np.array([0] * (n // 2) + [1] * (n // 2), np.uint8)
    dense, snarls = H.matrix_realising_tables(tables, is_case)
    got, ref = binary_both(gpu_ctx, oracle, dense, snarls, is_case)
    H.assert_binary_equal(got, ref)
    m = ref["fisher_code"] == 0
    assert m.sum() > 30
    np.testing.assert_allclose(got["p_fisher"][m], ref["p_fisher"][m], rtol=1e-12, atol=0)
    monkeypatch.setenv("STOAT_GPU_FISHER", "exact")
    got, ref = binary_both(gpu_ctx, oracle, dense, snarls, is_case)
    np.testing.assert_array_equal(got["p_fisher"][m], ref["p_fisher"][m])


def test_partitions_at_c5_width_quantitative(gpu_ctx, oracle):
    """Two-, three- and four-path snarls at 50,000 haplotypes: K3M runs in haplotype slices, the other shapes on whatever
    the width still allows (K3 / the general kernel)."""
    rng = np.random.default_rng(51)
    n = 25000
    dense, snarls = H.partition_problem(rng, 24, n, max_paths=4)
    y = rng.normal(size=n) * 2 + 5
    y = y + dense[snarls[0][0][0]].reshape(n, 2).sum(1)
    got, ref = quant_both(gpu_ctx, oracle, dense, snarls, y)
    H.assert_quant_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 10
    is_case = (rng.random(n) < 0.5).astype(np.uint8)
    gb, rb = binary_both(gpu_ctx, oracle, dense, snarls, is_case)
    H.assert_binary_equal(gb, rb)


@pytest.mark.parametrize("overlap", ["1", "4", "8"])
def test_overlapped_run_many_snarls(gpu_ctx, oracle, monkeypatch, overlap):
    """>= 65,536 snarls: the run is cut into chunks whose Fisher lists run beside the next chunk's count pass; the result
    must not depend on the number of chunks."""
    from stoat_b200 import synth

    monkeypatch.setenv("STOAT_GPU_OVERLAP", overlap)
    d = synth.numpy_dataset(70000, 48, seed=12, frac_missing_edge=0.02)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_binary_phenotype(d["is_case"])
    got = gpu_ctx.binary_batch(d["paths"])
    ref = oracle.binary_batch(d["bits"], d["n_cols"], d["is_case"], P)
    H.assert_binary_equal(got, ref)
    # resident mode, twice in a row without reading anything back in between (the runs only enqueue)
    gpu_ctx.stage_paths(d["paths"])
    gpu_ctx.binary_run_staged()
    gpu_ctx.binary_run_staged()
    again = gpu_ctx.binary_fetch()
    for k in got:
        np.testing.assert_array_equal(got[k], again[k], err_msg=k)
    t = gpu_ctx.timing()
    assert t["n_fisher"] > 0
    m = gpu_ctx.mean_timing(2)
    assert m["table_ms"] > 0 and m["stats_ms"] > 0


def test_chunked_upload_matches_single_copy(gpu_ctx, oracle, monkeypatch):
    """set_matrix sends the reference-layout bytes up in chunks of whole rows (copy of chunk k + 1 beside the repack of
    chunk k); rows must come out the same whatever the chunk size, also when rows are not byte aligned (n_cols % 8 != 0)."""
    rng = np.random.default_rng(5)
    for n_cols in (402, 4098):
        n_rows = 1000
        dense = rng.random((n_rows, n_cols)) < 0.5
        bits = oracle.pack_matrix(dense)
        monkeypatch.delenv("STOAT_GPU_UPLOAD_CHUNK_ROWS", raising=False)
        gpu_ctx.set_matrix(bits, n_rows, n_cols)
        want = [gpu_ctx.read_row(r).copy() for r in range(0, n_rows, 37)] + [gpu_ctx.read_row(n_rows - 1).copy()]
        monkeypatch.setenv("STOAT_GPU_UPLOAD_CHUNK_ROWS", "128")  # 8 chunks, the last one short
        gpu_ctx.set_matrix(bits, n_rows, n_cols)
        got = [gpu_ctx.read_row(r).copy() for r in range(0, n_rows, 37)] + [gpu_ctx.read_row(n_rows - 1).copy()]
        for a, b in zip(got, want):
            np.testing.assert_array_equal(a, b)
        for r in (0, 127, 128, 129, 511, 512, 999):
            row = np.unpackbits(gpu_ctx.read_row(r), bitorder="little")
            np.testing.assert_array_equal(row[:n_cols].astype(bool), dense[r])


@pytest.mark.parametrize("n_samples", [25000, 33000])
def test_logistic_at_c5_width_and_beyond(gpu_ctx, oracle, n_samples):
    """The logistic path at 50,000 haplotypes and at 66,000 (class-count reductions no longer packed two per word; the
    general kernel's sweeps over ~1,000 sample blocks).  Tolerance 1e-8 instead of 1e-9: at convergence the Newton step is
    H^-1 times the rounding noise of the gradient sum, and the reference's (and the oracle's) sequential sum over 33,000
    samples ordered controls-first passes through partial sums of ~8,000, so its own beta is defined to ~5e-12 absolute only
    (numpy's pairwise sum, the oracle and the GPU's two exact-count formulations differ by that much; the two GPU paths agree
    with each other to 1e-14)."""
    import test_gpu_logit
    from stoat_b200 import synth

    d = synth.numpy_dataset(120, n_samples, seed=14, frac_missing_edge=0.03)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_binary_phenotype(d["is_case"])
    gpu_ctx.set_covariates(None)
    got = gpu_ctx.binary_covar_batch(d["paths"])
    ref = oracle.binary_covar_batch(d["bits"], d["n_cols"], d["is_case"], P)
    assert test_gpu_logit.assert_logit_equal(got, ref, rtol=1e-8) > 50


def test_dense_eqtl_tcgen05_at_c5_width(oracle):
    """The int8 tcgen05 contraction at 25,000 samples (196 blocks of 128 samples per gene tile, INT32 sums up to 3.2e6) and
    with a gene count that is not a multiple of the 32-gene tile."""
    import stoat_b200
    from stoat_b200 import synth

    n = 25000
    d = synth.numpy_dataset(150, n, seed=15, frac_three=0.0, frac_node0=0.0)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    rng = np.random.default_rng(15)
    G = 37
    expr = rng.uniform(0, 8, size=(G, 1)) * (1.0 + rng.uniform(-1, 1, size=(G, n)))
    expr[3] += 0.2 * d["dense"][d["structure"].row0[5]].reshape(n, 2).sum(1) if "dense" in d else 0.0
    ref = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P)
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
        got = ctx.eqtl_dense_summary(d["paths"], expr, p_threshold=1e-2)
        assert ctx.timing()["n_general"] <= 4
    pv = np.where(ref["code"].reshape(-1, G) == 0, ref["p_value"].reshape(-1, G), np.inf)
    rows = np.flatnonzero((ref["filtered"] == 0) & np.isfinite(pv.min(axis=1)))
    assert rows.size > 100
    want = pv[rows].min(axis=1)
    np.testing.assert_allclose(got["min_p"][rows], want, rtol=1e-9, atol=1e-300)
    near = np.abs(pv[rows] - 1e-2) <= 1e-11
    lo, hi = ((pv[rows] < 1e-2) & ~near).sum(axis=1), ((pv[rows] < 1e-2) | near).sum(axis=1)
    assert np.all((lo <= got["n_below"][rows]) & (got["n_below"][rows] <= hi))
    assert np.isnan(got["min_p"][ref["filtered"] == 1]).all()
