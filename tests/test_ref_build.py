"""oracle/_ref (the reference's OWN hot-path sources, compiled in place by oracle/ref_build.py) against
  (1) the reference's golden vectors — this pins the Boost stand-ins the build links instead of the absent system Boost;
  (2) the oracle restatement on seeded problems — this pins the restatement, OLS included, on the reference's own code.
CPU only.  Skipped where the build is absent (a box without /root/reference that did not receive oracle/_ref/).
"""
import json
import os

import numpy as np
import pytest

import helpers as H

ref = pytest.importorskip("oracle.ref")
pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built and /root/reference not mounted")


def _close(a: str, b: str, rel=2e-4) -> bool:
    """Printed values carry 5 significant digits; a 1e-12 relative difference may flip the last one."""
    if a == b:
        return True
    try:
        x, y = float(a), float(b)
    except ValueError:
        return False
    if np.isnan(x) and np.isnan(y):
        return True
    return abs(x - y) <= rel * max(abs(x), abs(y))


def test_reference_code_reproduces_its_golden_binary_rows():
    # tests/expected_output/vcf/binary/binary_table_vcf.tsv: real Boost produced these strings
    rows = H.golden_binary_rows()
    assert len(rows) == 1463
    for r in rows:
        g0, g1 = H.parse_group_paths(r["GROUP_PATHS"])
        chi2, fisher = ref.fisher_khi2(g0, g1)
        assert chi2 == r["P_CHI2"], r
        assert fisher == r["P_FISHER"], r


def test_reference_code_unit_known_answers():
    with open(os.path.join(H.GOLDEN, "unit_known_answers.json")) as f:
        unit = json.load(f)
    n = 0
    for c in unit["binary_unit"]:
        g0, g1 = c["g0"], c["g1"]
        if "chi2_2x2" in c:
            assert ref.chi2_2x2(g0[0], g0[1], g1[0], g1[1]) == c["chi2_2x2"], c
            n += 1
        if "fisher" in c:
            assert ref.fisher_2x2(g0[0], g0[1], g1[0], g1[1]) == c["fisher"], c
            n += 1
        if "chi2_2xN" in c:
            assert ref.fisher_khi2(g0, g1)[0] == c["chi2_2xN"], c
            n += 1
    assert n == 19
    for c in unit["set_precision"]:
        assert ref.set_precision(c["value"]) == c["expect"]
    for r in unit["graph_rows"]:
        g0, g1 = H.parse_group_paths(r["group_paths"])
        assert ref.fisher_khi2(g0, g1) == (r["p_chi2"], r["p_fisher"])


def test_oracle_linear_regression_matches_reference_eigen(oracle):
    """LinearRegression::linear_regression (src/stats_test.cpp:424-506, vendored Eigen LDLT / JacobiSVD) vs the oracle's
    restatement: well-conditioned, with covariates, collinear (pseudo-inverse branch) and tiny-n designs."""
    rng = np.random.default_rng(7)
    n_checked = 0
    for trial in range(120):
        n = int(rng.integers(4, 80))
        k = int(rng.integers(1, 4))
        df = rng.choice([0.0, 0.5, 1.0], size=(n, k), p=[0.5, 0.3, 0.2])
        if trial % 7 == 3 and k > 1:
            df[:, 1] = df[:, 0]                       # rank-deficient: pseudo-inverse path
        if trial % 11 == 5:
            df[:, 0] = 1.0                            # collinear with the intercept
        covar = rng.normal(size=(n, int(rng.integers(1, 4)))) if trial % 3 == 0 else None
        if n < k + (0 if covar is None else covar.shape[1]) + 4:
            continue  # saturated fit: rss is rounding noise in any implementation, nothing to compare
        y = rng.normal(size=n) + (df[:, 0] * rng.normal())
        want = ref.linear_regression(df, y, covar)
        got = oracle.linear_regression(df, y, covar)
        got_s = [oracle.set_precision(got[q]) for q in ("p", "beta", "se", "r2")]
        assert all(_close(a, b) for a, b in zip(got_s, want)), (trial, n, k, got_s, want)
        n_checked += 1
    assert n_checked >= 100


# ---------------------------------------------------------------------------------------------------
# whole analyzers: the reference's analyze_and_write_snarl loop + writer vs rows formatted from the oracle's arrays
# ---------------------------------------------------------------------------------------------------
def _binary_want(oracle, res, spo, s):
    g0, g1 = res["g0"][spo[s]:spo[s + 1]], res["g1"][spo[s]:spo[s + 1]]
    keep = (g0 + g1) > 0
    return dict(GROUP_PATHS=",".join(f"{a}:{b}" for a, b in zip(g0[keep], g1[keep])),
                P_FISHER=oracle.format_value(res["p_fisher"][s], res["fisher_code"][s]),
                P_CHI2=oracle.format_value(res["p_chi2"][s], res["chi2_code"][s]))


def _check_binary(oracle, bits, n_rows, n_cols, P, nodes, is_case, threads):
    res = oracle.binary_batch(bits, n_cols, is_case, P)
    out = ref.run("binary", bits, n_rows, n_cols, P.snarl_path_off, nodes, is_case=is_case, n_threads=threads)
    spo = P.snarl_path_off.astype(np.int64)
    kept = np.flatnonzero(res["filtered"] == 0)
    assert sorted(out["rows"]) == list(kept), "filter decisions differ"
    assert out["n_filtered"] == P.n_snarls - kept.size
    for s in kept:
        (row,) = out["rows"][int(s)]
        want = _binary_want(oracle, res, spo, s)
        assert row["GROUP_PATHS"] == want["GROUP_PATHS"], (s, row, want)   # tables: bit-exact
        assert _close(row["P_FISHER"], want["P_FISHER"]) and _close(row["P_CHI2"], want["P_CHI2"]), (s, row, want)
    return kept.size


def _check_quant(oracle, bits, n_rows, n_cols, P, nodes, y, covar, threads):
    res = oracle.quantitative_batch(bits, n_cols, y, P, covar=covar)
    out = ref.run("quantitative", bits, n_rows, n_cols, P.snarl_path_off, nodes, y=y, covar=covar, n_threads=threads)
    spo = P.snarl_path_off.astype(np.int64)
    kept = np.flatnonzero(res["filtered"] == 0)
    assert sorted(out["rows"]) == list(kept), "filter decisions differ"
    for s in kept:
        (row,) = out["rows"][int(s)]
        assert row["ALLELE_PATHS"] == ",".join(str(v) for v in res["allele_paths"][spo[s]:spo[s + 1]]), (s, row)
        for col, key in (("P", "p_value"), ("RSQUARE", "r2"), ("BETA", "beta"), ("SE", "se")):
            assert _close(row[col], oracle.set_precision(res[key][s])), (s, col, row, res[key][s])
    return kept.size


def test_reference_analyzers_vs_oracle_synthetic(oracle):
    """The generator of the bench (every snarl kind: 2 / 3 paths, node-0 path, missing edge), natural node paths."""
    from stoat_b200 import synth

    d = synth.numpy_dataset(400, 150, seed=5, frac_missing_edge=0.03)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    nodes = ref.nodes_from_structure(d["structure"])
    assert _check_binary(oracle, d["bits"], d["n_rows"], d["n_cols"], P, nodes, d["is_case"], threads=4) > 200
    assert _check_quant(oracle, d["bits"], d["n_rows"], d["n_cols"], P, nodes, d["y"], None, threads=4) > 200
    covar = np.random.default_rng(3).normal(size=(150, 3))
    assert _check_quant(oracle, d["bits"], d["n_rows"], d["n_cols"], P, nodes, d["y"], covar, threads=1) > 200
    # both node-path encodings of the same CSR give the same rows
    alt = ref.nodes_from_csr(P.snarl_path_off, P.path_edge_off, P.edge_row, d["n_rows"])
    assert _check_binary(oracle, d["bits"], d["n_rows"], d["n_cols"], P, alt, d["is_case"], threads=2) > 200


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_reference_analyzers_vs_oracle_adversarial(oracle, seed):
    """Random CSR: rows shared between paths and snarls, empty edge lists, missing edges, all-zero rows, haplotypes on
    several paths (non-dyadic dosages, identical columns, rank-deficient designs)."""
    rng = np.random.default_rng(100 + seed)
    n_samples, n_rows = 60, 40
    dense, snarls = H.random_snarl_problem(rng, 150, n_samples, n_rows)
    spo, peo, er = H.csr_from_lists(snarls)
    bits = oracle.pack_matrix(dense)
    P = oracle.Paths(spo, peo, er)
    nodes = ref.nodes_from_csr(spo, peo, er, n_rows)
    is_case = (rng.random(n_samples) < 0.5).astype(np.uint8)
    y = rng.normal(size=n_samples)
    _check_binary(oracle, bits, n_rows, 2 * n_samples, P, nodes, is_case, threads=2)
    _check_quant(oracle, bits, n_rows, 2 * n_samples, P, nodes, y, None, threads=2)
    dense, snarls = H.partition_problem(rng, 120, n_samples)
    spo, peo, er = H.csr_from_lists(snarls)
    bits = oracle.pack_matrix(dense)
    P = oracle.Paths(spo, peo, er)
    nodes = ref.nodes_from_csr(spo, peo, er, dense.shape[0])
    assert _check_binary(oracle, bits, dense.shape[0], 2 * n_samples, P, nodes, is_case, threads=2) > 30
    assert _check_quant(oracle, bits, dense.shape[0], 2 * n_samples, P, nodes, y, rng.normal(size=(n_samples, 2)), threads=2) > 30


def test_reference_eqtl_analyzer_vs_oracle(oracle):
    """EQTLSnarlAnalyzer (src/snarl_analyzer.cpp:493-539): gene windows by found_gene_snarl (:471-491), one OLS per gene."""
    from stoat_b200 import synth

    n_samples, G, window = 90, 12, 40
    d = synth.numpy_dataset(160, n_samples, seed=9, frac_missing_edge=0.02)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    nodes = ref.nodes_from_structure(d["structure"])
    rng = np.random.default_rng(21)
    expr = rng.uniform(0, 8, size=(G, 1)) * (1 + rng.uniform(-1, 1, size=(G, n_samples)))
    gene_pos = np.stack([100 * np.arange(G), 100 * np.arange(G) + 30], axis=1).astype(np.uint64)
    start = rng.integers(0, 100 * G, size=P.n_snarls)
    snarl_pos = np.stack([start, start + rng.integers(1, 50, size=P.n_snarls)], axis=1).astype(np.uint64)
    # the window rule of found_gene_snarl, restated for the CSR the batched ABI takes
    off, idx = [0], []
    for s in range(P.n_snarls):
        lo = max(int(snarl_pos[s, 0]) - window, 0)
        hi = int(snarl_pos[s, 1]) + window
        idx += [g for g in range(G) if not (gene_pos[g, 1] < lo or gene_pos[g, 0] > hi)]
        off.append(len(idx))
    off, idx = np.array(off, np.uint64), np.array(idx, np.uint32)
    assert 0 < idx.size < P.n_snarls * G
    for covar in (None, rng.normal(size=(n_samples, 2))):
        res = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P, off, idx, covar=covar)
        out = ref.run("eqtl", d["bits"], d["n_rows"], d["n_cols"], P.snarl_path_off, nodes, covar=covar, expression=expr,
                      gene_pos=gene_pos, snarl_pos=snarl_pos, window=window, n_threads=3)
        spo = P.snarl_path_off.astype(np.int64)
        n_rows_checked = 0
        for s in range(P.n_snarls):
            rows = {r["GENE"]: r for r in out["rows"].get(s, [])}
            if res["filtered"][s]:
                assert not rows
                continue
            assert sorted(rows) == sorted(f"g{g}" for g in idx[int(off[s]):int(off[s + 1])])
            for k in range(int(off[s]), int(off[s + 1])):
                row = rows[f"g{idx[k]}"]
                assert row["ALLELE_PATHS"] == ",".join(str(v) for v in res["allele_paths"][spo[s]:spo[s + 1]])
                for col, key in (("P", "p_value"), ("RSQUARE", "r2"), ("BETA", "beta"), ("SE", "se")):
                    assert _close(row[col], oracle.set_precision(res[key][k])), (s, k, col, row, res[key][k])
                n_rows_checked += 1
        assert n_rows_checked > 100


# ---------------------------------------------------------------------------------------------------
# logistic regression with covariates: LogisticRegression + BinaryCovarSnarlAnalyzer
# ---------------------------------------------------------------------------------------------------
def _logit_strings(oracle, got):
    if got["code"] == 1:
        return ["NA", "NA", "NA"]
    return [oracle.set_precision(got[q]) for q in ("p", "beta", "se")]


def test_oracle_logistic_regression_matches_reference_eigen(oracle):
    """LogisticRegression::logistic_regression (src/stats_test.cpp:49-176: IRLS, L2 1e-4, Wald z, Holm over the dosage
    columns) vs the oracle's restatement; covariates are passed and — as in the reference — have no effect."""
    rng = np.random.default_rng(11)
    n_checked = n_multi = n_na = 0
    for trial in range(160):
        n = int(rng.integers(12, 200))
        k = int(rng.integers(1, 4))
        df = rng.choice([0.0, 1.0 / 3.0, 0.5, 1.0], size=(n, k), p=[0.5, 0.05, 0.25, 0.2])
        eff = rng.normal(scale=2.0)
        y = (rng.random(n) < 1.0 / (1.0 + np.exp(-(eff * df[:, 0] - 0.3)))).astype(np.uint8)
        if trial % 9 == 4:
            y = (df[:, 0] > 0.4).astype(np.uint8)       # perfect separation: large betas held by the L2 term
        if trial % 13 == 6:
            y[:] = 1                                    # constant phenotype
        covar = rng.normal(size=(n, 2)) if trial % 2 == 0 else None
        want = ref.logistic_regression(df, y, covar)
        got = oracle.logistic_regression(df, y, covar)
        got_s = _logit_strings(oracle, got)
        assert all(_close(a, b) for a, b in zip(got_s, want)), (trial, n, k, got_s, want)
        with_cov = oracle.logistic_regression(df, y, covar, use_covariates=False)
        assert with_cov == got or (np.isnan(got["p"]) and np.isnan(with_cov["p"]))
        n_checked += 1
        n_multi += k > 1
        n_na += want[0] == "NA"
    assert n_checked == 160 and n_multi > 60


def _check_binary_covar(oracle, bits, n_rows, n_cols, P, nodes, is_case, covar, threads):
    res = oracle.binary_covar_batch(bits, n_cols, is_case, P, covar=covar)
    out = ref.run("binary_covar", bits, n_rows, n_cols, P.snarl_path_off, nodes, is_case=is_case, covar=covar, n_threads=threads)
    spo = P.snarl_path_off.astype(np.int64)
    kept = np.flatnonzero((res["filtered"] == 0) & (res["code"] != 5))  # code 5: the reference reads out of bounds
    undef = set(np.flatnonzero((res["filtered"] == 0) & (res["code"] == 5)).tolist())
    assert sorted(set(out["rows"]) - undef) == list(kept), "filter decisions differ"
    for s in kept:
        (row,) = out["rows"][int(s)]
        assert row["ALLELE_PATHS"] == ",".join(str(v) for v in res["allele_paths"][spo[s]:spo[s + 1]]), (s, row)
        want = ["NA"] * 3 if res["code"][s] == 1 else [oracle.set_precision(res[k][s]) for k in ("p_value", "beta", "se")]
        for col, w in zip(("P", "BETA", "SE"), want):
            assert _close(row[col], w), (s, col, row, w)
    return kept.size


def test_reference_binary_covar_analyzer_vs_oracle(oracle):
    from stoat_b200 import synth

    d = synth.numpy_dataset(400, 150, seed=6, frac_missing_edge=0.03)
    P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
    nodes = ref.nodes_from_structure(d["structure"])
    covar = np.random.default_rng(4).normal(size=(150, 3))
    assert _check_binary_covar(oracle, d["bits"], d["n_rows"], d["n_cols"], P, nodes, d["is_case"], covar, threads=4) > 200
    assert _check_binary_covar(oracle, d["bits"], d["n_rows"], d["n_cols"], P, nodes, d["is_case"], None, threads=1) > 200
    rng = np.random.default_rng(17)
    n_samples = 60
    dense, snarls = H.partition_problem(rng, 150, n_samples)
    spo, peo, er = H.csr_from_lists(snarls)
    bits = oracle.pack_matrix(dense)
    P = oracle.Paths(spo, peo, er)
    nodes = ref.nodes_from_csr(spo, peo, er, dense.shape[0])
    is_case = (rng.random(n_samples) < 0.5).astype(np.uint8)
    assert _check_binary_covar(oracle, bits, dense.shape[0], 2 * n_samples, P, nodes, is_case, None, threads=2) > 30
    dense, snarls = H.random_snarl_problem(rng, 150, n_samples, 40)
    spo, peo, er = H.csr_from_lists(snarls)
    bits = oracle.pack_matrix(dense)
    P = oracle.Paths(spo, peo, er)
    nodes = ref.nodes_from_csr(spo, peo, er, 40)
    _check_binary_covar(oracle, bits, 40, 2 * n_samples, P, nodes, is_case, None, threads=2)


def test_shim_builds_against_reference_types():
    """host/stoat_gpu_shim.cpp compiled with -DSTOAT_GPU_SHIM_USE_REFERENCE_TYPES against the reference's own headers (plus the
    three accessors of INTEGRATION.md §2a, patched into a copy of src/matrix.hpp at build time) and linked with the
    reference's matrix.cpp / snarl_data_t.cpp objects: the mode a maintainer compiles.  Running it needs a GPU
    (tests/test_gpu_shim.py); here: it builds, and it binds both libraries."""
    import subprocess

    from oracle import ref_build

    exe = ref_build.build_shim_reftypes()
    if exe is None:
        pytest.skip("libstoat_gpu.so or oracle/_ref not built")
    assert os.path.exists(exe)
    needed = subprocess.run(["readelf", "-d", exe], capture_output=True, text=True).stdout
    assert "libstoat_ref.so" in needed and "libstoat_gpu.so" in needed
    syms = subprocess.run(["nm", "-C", "--undefined-only", exe], capture_output=True, text=True).stdout
    assert "stoat_vcf::EdgeBySampleMatrix::find_edge" in syms and "stoat_vcf::EdgeBySampleMatrix::push_matrix" in syms
    assert "stoat_gpu_binary_batch" in syms
