"""The oracle against every golden vector the reference's own tests hold for the hot path (CPU only).

Sources: see tests/golden/make_golden.py.  String-exact, as the reference's tests compare strings.
"""
import json
import os

import numpy as np
import pytest

from helpers import GOLDEN, golden_binary_rows, golden_quant_rows, parse_group_paths


@pytest.fixture(scope="module")
def unit():
    with open(os.path.join(GOLDEN, "unit_known_answers.json")) as f:
        return json.load(f)


def test_binary_unit_known_answers(oracle, unit):
    # tests/unittest/binary_test_unit.cpp:7-116
    n = 0
    for c in unit["binary_unit"]:
        g0, g1 = c["g0"], c["g1"]
        if "chi2_2x2" in c:
            assert oracle.format_value(*oracle.chi2_2x2(g0[0], g0[1], g1[0], g1[1])) == c["chi2_2x2"], c
            n += 1
        if "fisher" in c:
            assert oracle.format_value(*oracle.fisher_2x2(g0[0], g0[1], g1[0], g1[1])) == c["fisher"], c
            n += 1
        if "chi2_2xN" in c:
            assert oracle.format_value(*oracle.chi2_2xN(g0, g1)) == c["chi2_2xN"], c
            n += 1
    assert n == 19


def test_set_precision_known_answers(oracle, unit):
    # tests/unittest/utils_unit.cpp:9-23
    for c in unit["set_precision"]:
        assert oracle.set_precision(c["value"]) == c["expect"]
    assert oracle.format_value(0.0, oracle.P_NA) == "NA"
    assert oracle.format_value(0.0, oracle.P_DBLMAX) == "1.798e+308"


def test_graph_system_rows(oracle, unit):
    # tests/system/graph_simu_test.cpp:165-168,199,336-337 (same fisher_khi2 as the vcf path)
    for r in unit["graph_rows"]:
        g0, g1 = parse_group_paths(r["group_paths"])
        (cv, ck), (fv, fk) = oracle.fisher_khi2(g0, g1)
        assert oracle.format_value(cv, ck) == r["p_chi2"]
        assert oracle.format_value(fv, fk) == r["p_fisher"]


def test_golden_binary_table(oracle):
    # tests/expected_output/vcf/binary/binary_table_vcf.tsv: 1,463 rows
    rows = golden_binary_rows()
    assert len(rows) == 1463
    n_big = n_zero = n_one = 0
    for r in rows:
        g0, g1 = parse_group_paths(r["GROUP_PATHS"])
        (cv, ck), (fv, fk) = oracle.fisher_khi2(g0, g1)
        assert oracle.format_value(cv, ck) == r["P_CHI2"], r
        assert oracle.format_value(fv, fk) == r["P_FISHER"], r
        n_zero += r["P_CHI2"] == "0"
        n_one += r["P_FISHER"] == "1"
    assert n_zero >= 5 and n_one >= 100  # the cancellation and centre-mass-zero branches are exercised


def test_identify_path_and_filters(oracle, unit):
    # tests/unittest/snarl_analyzer_unit.cpp:100-221, tests/unittest/quantitative_test_unit.cpp:15-83
    ip = unit["identify_path"]
    bits = np.zeros(oracle.matrix_bytes(ip["n_rows"], ip["n_cols"]) + 8, np.uint8)
    for r, c in ip["set"]:
        oracle.lib().orc_matrix_set(bits.ctypes.data_as(__import__("ctypes").POINTER(__import__("ctypes").c_uint8)),
                                    ip["n_cols"], r, c)
    assert list(oracle.identify_path(bits, ip["n_cols"], ip["path_rows"])) == ip["expect"]
    # empty edge list => every column; missing edge => nobody (src/snarl_analyzer.cpp:334-354)
    assert list(oracle.identify_path(bits, ip["n_cols"], [])) == [0, 1, 2]
    assert list(oracle.identify_path(bits, ip["n_cols"], [0, oracle.MISSING_EDGE])) == []
    for c in unit["filtration_binary"]:
        assert oracle.filtration_binary(c["g0"], c["g1"], c["total"], c["min_ind"], c["min_hap"], c["maf"]) == c["expect"]
    for c in unit["filtration_quant"]:
        assert oracle.filtration_quant(np.array(c["df"], float), c["min_ind"], c["min_hap"], c["maf"]) == c["expect"]
    for c in unit["remove_empty_columns_quant"]:
        np.testing.assert_array_equal(oracle.remove_empty_columns_quant(c["df"]), np.array(c["expect"], float))
    # combine_identical_columns sums duplicates, only with >= 3 columns (src/snarl_analyzer.cpp:620-674)
    df = np.array([[0.5, 0.5, 0.0], [0.25, 0.25, 0.5], [0.0, 0.0, 1.0]])
    np.testing.assert_array_equal(oracle.combine_identical_columns(df), np.array([[1.0, 0.0], [0.5, 0.5], [0.0, 1.0]]))
    df2 = np.array([[0.5, 0.5], [0.5, 0.5]])
    np.testing.assert_array_equal(oracle.combine_identical_columns(df2), df2)


def test_matrix_layout(oracle):
    # src/matrix.cpp:67-84: bit = row * cols + col, LSB first, rows not byte aligned
    dense = np.zeros((3, 13), bool)
    dense[1, 0] = dense[2, 12] = dense[0, 7] = True
    bits = oracle.pack_matrix(dense)
    assert bits[0] == 0x80 and bits[1] == (1 << (13 - 8)) and bits[(2 * 13 + 12) // 8] == 1 << ((2 * 13 + 12) % 8)
    np.testing.assert_array_equal(oracle.unpack_matrix(bits, 3, 13), dense)


def test_ols_golden_row_identities(oracle):
    """OLS is pinned by the reference only through its expected TSVs, whose VCF input is a missing blob.  What the
    rows DO pin, for 2-path snarls where every haplotype is assigned (n_used = 200 samples, one regression column):
      (beta/se)^2 = df_res * R^2 / (1 - R^2) with df_res = n - p + 1 = 199  (the reference's sic, stats_test.cpp:467)
      P           = 2 * StudentT_sf(|beta/se|, df_res)
    Checked at the 4-significant-digit precision of the file."""
    rows = [r for r in golden_quant_rows() if r["ALLELE_PATHS"].count(",") == 1
            and sum(map(int, r["ALLELE_PATHS"].split(","))) == 400]
    assert len(rows) > 800
    n = 200
    quirk_closer = checked = 0
    for r in rows:
        r2, b, se, p = float(r["RSQUARE"]), float(r["BETA"]), float(r["SE"]), float(r["P"])
        if se == 0 or r2 < 1e-3:
            continue
        t2 = (b / se) ** 2
        q = (n - 1) * r2 / (1 - r2)
        honest = (n - 2) * r2 / (1 - r2)
        checked += 1
        quirk_closer += abs(q / t2 - 1) < abs(honest / t2 - 1)
        assert abs(q / t2 - 1) < 2e-3
        assert abs(2 * oracle.student_t_sf(n - 1, abs(b / se)) - p) <= 2e-2 * p
    assert checked > 300 and quirk_closer > 0.95 * checked


def test_bh_adjust_reference_unit_vectors(oracle):
    """tests/unittest/post_processing_unit.cpp:12-46 (known answers), :54-66 (clamp), :48-52 (empty)."""
    import numpy as np

    got = oracle.bh_adjust([0.01, 0.04, 0.03, 0.002, 0.05])
    np.testing.assert_allclose(got, [0.025, 0.05, 0.05, 0.01, 0.05], rtol=1e-12)
    assert (oracle.bh_adjust([0.9, 0.95, 0.99]) <= 1.0).all()
    assert oracle.bh_adjust([]).size == 0
    # monotone in p and never below p
    rng = np.random.default_rng(1)
    p = rng.random(1000) ** 3
    adj = oracle.bh_adjust(p)
    o = np.argsort(p)
    assert (np.diff(adj[o]) >= 0).all() and (adj >= p).all()
