"""Parity of the CUDA binary path (K0 repack, K1 path_and_popc, K2 binary_stats) with the oracle, through the C ABI.

Bit-exact: g0/g1/total_sum/filtered/codes.  P-values: 1e-9 relative (+ the 2^-53 quantum of the reference's own
`1.0 - cdf`), and string-exact against the reference's golden TSV rows."""
import json
import os

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def run_both(ctx, oracle, dense, snarls, is_case, **thr):
    n_rows, n_cols = dense.shape
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    from stoat_b200 import SnarlPaths

    ctx.set_matrix(bits, n_rows, n_cols)
    ctx.set_binary_phenotype(is_case)
    got = ctx.binary_batch(SnarlPaths(spo, peo, er), **thr)
    ref = oracle.binary_batch(bits, n_cols, is_case, oracle.Paths(spo, peo, er), **thr)
    return got, ref


@pytest.mark.parametrize("n_cols", [2, 14, 126, 128, 130, 402, 1000, 4098])
def test_repack_rows(gpu_ctx, oracle, n_cols):
    # K0 against the reference addressing (src/matrix.cpp:67-84), including rows that start mid-byte
    rng = np.random.default_rng(n_cols)
    n_rows = 37
    dense = rng.random((n_rows, n_cols)) < 0.5
    bits = oracle.pack_matrix(dense)
    gpu_ctx.set_matrix(bits, n_rows, n_cols)
    pitch = gpu_ctx.row_pitch
    assert pitch == ((n_cols + 255) // 256) * 32
    for r in (0, 1, 17, n_rows - 1):
        row = gpu_ctx.read_row(r)
        got = np.unpackbits(row, bitorder="little")
        np.testing.assert_array_equal(got[:n_cols].astype(bool), dense[r])
        assert not got[n_cols:].any(), "padding bits must be zero"


def test_reference_unit_case_identify_path(gpu_ctx, oracle):
    # tests/unittest/snarl_analyzer_unit.cpp:100-118 has 3 columns; the ABI needs an even haplotype count, so the
    # same rows are laid out over 4 columns with column 3 empty
    dense = np.zeros((2, 4), bool)
    dense[0, 0] = dense[1, 0] = dense[0, 2] = dense[1, 2] = True
    got, ref = run_both(gpu_ctx, oracle, dense, [[[0, 1]]], np.array([0, 1], np.uint8), min_ind=0, min_hap=0, maf=0.0)
    assert got["g0"][0] + got["g1"][0] == 2
    H.assert_binary_equal(got, ref)


@pytest.mark.parametrize("seed,n_samples,n_rows", [(1, 10, 12), (2, 101, 40), (3, 200, 64), (4, 1000, 200), (5, 2500, 90)])
def test_random_problems(gpu_ctx, oracle, seed, n_samples, n_rows):
    rng = np.random.default_rng(seed)
    dense, snarls = H.random_snarl_problem(rng, 300, n_samples, n_rows)
    is_case = (rng.random(n_samples) < 0.45).astype(np.uint8)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=2, min_hap=3, maf=0.01)
    H.assert_binary_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 20


@pytest.mark.parametrize("seed,n_samples", [(11, 200), (12, 1300), (13, 5000)])
def test_partition_problems(gpu_ctx, oracle, seed, n_samples):
    rng = np.random.default_rng(seed)
    dense, snarls = H.partition_problem(rng, 400 if n_samples < 5000 else 60, n_samples)
    is_case = (rng.random(n_samples) < 0.5).astype(np.uint8)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case)
    H.assert_binary_equal(got, ref)
    assert (ref["filtered"] == 0).sum() > 10


def test_edge_cases(gpu_ctx, oracle):
    rng = np.random.default_rng(7)
    n_samples = 64
    dense = rng.random((20, 2 * n_samples)) < 0.5
    dense[3] = False
    dense[4] = True
    is_case = (np.arange(n_samples) % 2).astype(np.uint8)
    many_rows = [[0, 1], [2, 5], [6, 7], [8, 9], [10, 11], [12, 13], [14, 15]]       # > 6 distinct rows: LDG path
    long_paths = [[int(x) for x in rng.integers(0, 20, 30)], [int(x) for x in rng.integers(0, 20, 30)]]
    snarls = [
        [],                                  # a snarl without paths
        [[]],                                # one path, no checked edge: everybody
        [[], []],                            # two match-all paths
        [[H.MISSING], [0]],                  # missing edge: nobody
        [[3], [4]],                          # all-zero row, all-one row
        [[0, 0, 0], [1]],                    # the same row several times
        [[0], [1], [2], [3], [4], [5]],      # 6 columns -> 2xN with an empty column
        many_rows,
        long_paths,
        [[4], [4]],                          # identical columns
        [[5, H.MISSING, 6], [7, 8]],
    ]
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=1, min_hap=1, maf=0.0)
    H.assert_binary_equal(got, ref)
    # empty batch
    from stoat_b200 import SnarlPaths

    out = gpu_ctx.binary_batch(SnarlPaths(np.zeros(1, np.uint64), np.zeros(1, np.uint64), np.zeros(0, np.uint32)))
    assert out["filtered"].size == 0


def test_golden_rows_through_the_gpu(gpu_ctx, oracle):
    """C1': the 1,463 (GROUP_PATHS -> P_FISHER, P_CHI2) rows of tests/expected_output/vcf/binary/binary_table_vcf.tsv.
    A matrix realising each table is pushed through K0+K1+K2; the formatted strings must equal the reference's."""
    from stoat_b200 import format_value

    rows = H.golden_binary_rows()
    is_case = H.golden_binary_phenotype()
    tables = [H.parse_group_paths(r["GROUP_PATHS"]) for r in rows]
    dense, snarls = H.matrix_realising_tables(tables, is_case)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case)
    H.assert_binary_equal(got, ref)
    bad = []
    k = 0
    for i, r in enumerate(rows):
        g0, g1 = tables[i]
        p0 = k
        k += len(g0)
        assert list(got["g0"][p0:k]) == g0 and list(got["g1"][p0:k]) == g1
        assert got["filtered"][i] == 0, r
        sc = format_value(got["p_chi2"][i], got["chi2_code"][i])
        sf = format_value(got["p_fisher"][i], got["fisher_code"][i])
        if sc != r["P_CHI2"] or sf != r["P_FISHER"]:
            bad.append((r, sc, sf))
    assert not bad, bad[:5]


def test_unit_known_answers_through_the_gpu(gpu_ctx, oracle):
    # tests/unittest/binary_test_unit.cpp known answers, including the NA / "1" / 2xN cases
    from stoat_b200 import format_value

    with open(os.path.join(H.GOLDEN, "unit_known_answers.json")) as f:
        unit = json.load(f)
    cases = [c for c in unit["binary_unit"] if sum(c["g0"]) + sum(c["g1"]) > 0]
    n_ctrl = max(sum(c["g0"]) for c in cases)
    n_case = max(sum(c["g1"]) for c in cases)
    n_ctrl += n_ctrl % 2
    n_case += n_case % 2
    is_case = np.array([0] * (n_ctrl // 2) + [1] * (n_case // 2), np.uint8)
    dense, snarls = H.matrix_realising_tables([(c["g0"], c["g1"]) for c in cases], is_case)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case, min_ind=0, min_hap=0, maf=-1.0)
    H.assert_binary_equal(got, ref)
    for i, c in enumerate(cases):
        nonempty = sum(1 for a, b in zip(c["g0"], c["g1"]) if a + b)
        if nonempty < 2:
            assert got["filtered"][i] == 1  # the analyzer never calls the tests with < 2 columns
            continue
        sc = format_value(got["p_chi2"][i], got["chi2_code"][i])
        if nonempty == len(c["g0"]):  # no column dropped: the unit test's table is what reaches fisher_khi2
            want = c.get("chi2_2x2", c.get("chi2_2xN"))
            assert sc == want, (c, sc)
            if "fisher" in c:
                assert format_value(got["p_fisher"][i], got["fisher_code"][i]) == c["fisher"], c


def test_large_tables_fisher_parity(gpu_ctx, oracle, monkeypatch):
    """Fisher's recurrence in the reference's order of operations at 10,000 haplotypes (~400 steps per table): the
    default kernel takes the quotient as N * rcp(D) (1 ulp per step) and must stay within 1e-12; with
    STOAT_GPU_FISHER=exact (correctly rounded division) the recurrence is bit-identical to the reference's."""
    rng = np.random.default_rng(21)
    n_samples = 5000
    tables = []
    for _ in range(120):
        f = rng.uniform(0.05, 0.95)
        shift = rng.choice([0.0, 0.0, 0.02, 0.05])
        a = rng.binomial(5000, f)
        b = rng.binomial(5000, min(f + shift, 0.99))
        tables.append(([5000 - a, a], [5000 - b, b]))
    is_case = np.array([0] * 2500 + [1] * 2500, np.uint8)
    dense, snarls = H.matrix_realising_tables(tables, is_case)
    got, ref = run_both(gpu_ctx, oracle, dense, snarls, is_case)
    H.assert_binary_equal(got, ref)
    m = ref["fisher_code"] == 0
    np.testing.assert_allclose(got["p_fisher"][m], ref["p_fisher"][m], rtol=1e-12, atol=0)
    import stoat_b200

    monkeypatch.setenv("STOAT_GPU_FISHER", "exact")
    with stoat_b200.Context(0) as ctx:
        got, ref = run_both(ctx, oracle, dense, snarls, is_case)
    np.testing.assert_array_equal(got["p_fisher"][m], ref["p_fisher"][m])


def test_ldg_variant_matches(oracle, monkeypatch):
    """The straight-from-global kernel (used for snarls that do not fit a TMA stage) on everything."""
    import stoat_b200

    monkeypatch.setenv("STOAT_GPU_K1", "ldg")
    ctx = stoat_b200.Context(0)
    try:
        rng = np.random.default_rng(31)
        dense, snarls = H.partition_problem(rng, 200, 700)
        is_case = (rng.random(700) < 0.5).astype(np.uint8)
        got, ref = run_both(ctx, oracle, dense, snarls, is_case)
        H.assert_binary_equal(got, ref)
    finally:
        ctx.close()


def test_synthetic_c3_shape_and_properties(gpu_ctx, oracle):
    """The bench generator at reduced size vs the oracle, then size-independent properties at a larger size:
    a haplotype is on at most one of {ref, alt, third} => sum(g0+g1) <= n_cols per snarl, and the resident
    (staged) path returns the same bytes as the one-shot call."""
    from stoat_b200 import synth

    d = synth.numpy_dataset(1500, 500, seed=5, frac_missing_edge=0.02)
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_binary_phenotype(d["is_case"])
    got = gpu_ctx.binary_batch(d["paths"])
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    ref = oracle.binary_batch(d["bits"], d["n_cols"], d["is_case"], P)
    H.assert_binary_equal(got, ref)
    assert 0.5 < (ref["filtered"] == 0).mean() < 1.0
    # staged / resident mode
    gpu_ctx.stage_paths(d["paths"])
    gpu_ctx.binary_run_staged()
    again = gpu_ctx.binary_fetch()
    for k in got:
        np.testing.assert_array_equal(got[k], again[k], err_msg=k)
    st = d["structure"]
    normal = np.flatnonzero(st.kind <= 1)  # not the match-all / missing-edge variants
    tot = got["total_sum"][normal]
    assert (tot <= d["n_cols"]).all() and (tot >= 0.9 * d["n_cols"]).all()


def test_invalid_row_index_is_reported(gpu_ctx, oracle):
    import stoat_b200

    dense = np.ones((2, 8), bool)
    gpu_ctx.set_matrix(oracle.pack_matrix(dense), 2, 8)
    gpu_ctx.set_binary_phenotype(np.array([0, 1, 0, 1], np.uint8))
    spo, peo, er = H.csr_from_lists([[[0], [7]]])
    with pytest.raises(stoat_b200.StoatGpuError):
        gpu_ctx.binary_batch(stoat_b200.SnarlPaths(spo, peo, er))


def test_reused_context_big_snarl_after_biallelic_batch(oracle):
    """Regression: the path-program buffer is reused between batches; a snarl that takes the LDG path must not leave the
    previous batch's shape class / first-path index where the TMA kernel reads them (it used to write two stale counts)."""
    import stoat_b200
    from stoat_b200 import SnarlPaths

    rng = np.random.default_rng(77)
    S, n_samples = 6000, 40
    H2 = 2 * n_samples
    dense = rng.random((4 * S + 8, H2)) < 0.7
    is_case = (rng.random(n_samples) < 0.5).astype(np.uint8)
    bial = [[[4 * s, 4 * s + 1], [4 * s + 2, 4 * s + 3]] for s in range(S)]
    big_last = bial[:-1] + [[[4 * S - 4 + k for k in range(8)]]]  # one path, 8 distinct rows: over the 6-row stage limit
    bits = oracle.pack_matrix(dense)
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(bits, dense.shape[0], H2)
        ctx.set_binary_phenotype(is_case)
        for snarls in (bial, big_last, bial, big_last):
            spo, peo, er = H.csr_from_lists(snarls)
            got = ctx.binary_batch(SnarlPaths(spo, peo, er), min_ind=0, min_hap=0, maf=0.0)
            ref = oracle.binary_batch(bits, H2, is_case, oracle.Paths(spo, peo, er), min_ind=0, min_hap=0, maf=0.0)
            H.assert_binary_equal(got, ref)


def test_matrix_from_set_bits(gpu_ctx, oracle):
    """stoat_gpu_set_matrix_pairs (the device half of make_edge_matrix): the (row, column) pairs of the set
    bits, in any order and with repeats, give the same device matrix and the same results as the dense upload."""
    import stoat_b200
    from stoat_b200 import SnarlPaths

    rng = np.random.default_rng(77)
    dense, snarls = H.partition_problem(rng, 120, 333)  # 666 columns: rows are not byte aligned in the reference layout
    n_rows, n_cols = dense.shape
    is_case = (rng.random(333) < 0.5).astype(np.uint8)
    bits = oracle.pack_matrix(dense)
    spo, peo, er = H.csr_from_lists(snarls)
    gpu_ctx.set_matrix(bits, n_rows, n_cols)
    want_rows = [gpu_ctx.read_row(r).copy() for r in range(n_rows)]
    gpu_ctx.set_binary_phenotype(is_case)
    want = gpu_ctx.binary_batch(SnarlPaths(spo, peo, er))
    r, c = np.nonzero(dense)
    order = rng.permutation(r.size)
    r = np.concatenate([r[order], r[:50]]).astype(np.uint32)  # shuffled, with repeats
    c = np.concatenate([c[order], c[:50]]).astype(np.uint32)
    gpu_ctx.set_matrix_pairs(r, c, n_rows, n_cols)
    for k in range(n_rows):
        np.testing.assert_array_equal(gpu_ctx.read_row(k), want_rows[k], err_msg=f"row {k}")
    gpu_ctx.set_binary_phenotype(is_case)
    got = gpu_ctx.binary_batch(SnarlPaths(spo, peo, er))
    for k in want:
        np.testing.assert_array_equal(got[k], want[k], err_msg=k)
    with pytest.raises(stoat_b200.StoatGpuError):
        gpu_ctx.set_matrix_pairs(np.array([n_rows], np.uint32), np.array([0], np.uint32), n_rows, n_cols)
    gpu_ctx.set_matrix_pairs(np.zeros(0, np.uint32), np.zeros(0, np.uint32), n_rows, n_cols)  # empty: all-zero matrix
    assert not gpu_ctx.read_row(0).any()
