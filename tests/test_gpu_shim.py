"""The C++ host shim (host/stoat_gpu_shim.cpp: the batched stand-ins of the reference's Binary / Quantitative / EQTL
SnarlAnalyzer classes) driven by host/shim_selftest on a B200.

* golden: the reference's expected `stoat vcf` binary output (tests/expected_output/vcf/binary/binary_table_vcf.tsv,
  committed as tests/golden/binary_table_vcf.golden.tsv) must come back byte for byte.
* synth: rows of all four analyzers against rows formatted from the oracle's results on the dumped problem.
"""
import os
import subprocess

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def selftest():
    from stoat_b200 import build

    build.build()
    return build.build_host()


def test_golden_binary_tsv_byte_for_byte(selftest, tmp_path):
    out = tmp_path / "binary.tsv"
    golden = os.path.join(H.GOLDEN, "binary_table_vcf.golden.tsv")
    r = subprocess.run([selftest, "golden", golden, os.path.join(H.GOLDEN, "binary_phenotype.tsv"), str(out)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    got = open(out).read().splitlines()
    want = open(golden).read().splitlines()
    assert len(got) == len(want) == 1464
    bad = [(g, w) for g, w in zip(got, want) if g != w]
    assert not bad, f"{len(bad)} rows differ, first: {bad[0]}"


def test_golden_through_the_reference_types_build(tmp_path):
    """The same golden file through the shim compiled against the REFERENCE'S OWN Snarl_data_t / Edge_t /
    EdgeBySampleMatrix (oracle/_ref/shim_reftypes_selftest, built by oracle/ref_build.py with the INTEGRATION.md §2a
    accessors): the build a maintainer would make."""
    exe = os.path.join(ROOT, "oracle", "_ref", "shim_reftypes_selftest")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/shim_reftypes_selftest not built (needs /root/reference at build time)")
    out = tmp_path / "binary.tsv"
    golden = os.path.join(H.GOLDEN, "binary_table_vcf.golden.tsv")
    r = subprocess.run([exe, "golden", golden, os.path.join(H.GOLDEN, "binary_phenotype.tsv"), str(out)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert open(out).read() == open(golden).read()


def _rows(path):
    with open(path) as f:
        return [ln.rstrip("\n").split("\t") for ln in f][1:]


def _close(a: str, b: str) -> bool:
    """Printed values carry 5 significant digits; a 1e-9 relative difference may flip the last one."""
    if a == b:
        return True
    try:
        x, y = float(a), float(b)
    except ValueError:
        return False
    return abs(x - y) <= 2e-4 * max(abs(x), abs(y))


def test_synthetic_chunk_all_analyzers(selftest, oracle, tmp_path):
    r = subprocess.run([selftest, "synth", "11", "120", "90", str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    d = str(tmp_path)
    meta = open(os.path.join(d, "meta.txt")).read().splitlines()
    n_rows, n_cols, S, G = (int(x) for x in meta[0].split())
    info = [ln.split("\t") for ln in meta[1:]]
    ld = lambda name, t: np.fromfile(os.path.join(d, name), dtype=t)  # noqa: E731
    bits = ld("bits.u8", np.uint8)
    P = oracle.Paths(ld("snarl_path_off.u64", np.uint64), ld("path_edge_off.u64", np.uint64), ld("edge_row.u32", np.uint32))
    is_case, y = ld("is_case.u8", np.uint8), ld("y.f64", np.float64)
    covar = ld("covar.f64", np.float64).reshape(-1, 3)
    expr = ld("expression.f64", np.float64).reshape(G, n_cols // 2)
    goff, gidx = ld("gene_off.u64", np.uint64), ld("gene_idx.u32", np.uint32)
    fmt = oracle.format_value
    spo = P.snarl_path_off.astype(np.int64)
    prefix = lambda s: ["chr7", info[s][0], info[s][1], info[s][2], info[s][3]]  # noqa: E731

    # binary
    ref = oracle.binary_batch(bits, n_cols, is_case, P)
    want = []
    for s in range(S):
        if ref["filtered"][s]:
            continue
        g0, g1 = ref["g0"][spo[s]:spo[s + 1]], ref["g1"][spo[s]:spo[s + 1]]
        keep = (g0 + g1) > 0
        gp = ",".join(f"{a}:{b}" for a, b in zip(g0[keep], g1[keep]))
        want.append(prefix(s) + [fmt(ref["p_fisher"][s], ref["fisher_code"][s]), fmt(ref["p_chi2"][s], ref["chi2_code"][s]),
                                 gp, info[s][4]])
    got = _rows(os.path.join(d, "binary.tsv"))
    assert len(got) == len(want) > 20
    for g, w in zip(got, want):
        assert g[:5] == w[:5] and g[7:] == w[7:] and _close(g[5], w[5]) and _close(g[6], w[6]), (g, w)

    # quantitative, without and with covariates
    for name, cv in (("quantitative.tsv", None), ("quantitative_covar.tsv", covar)):
        ref = oracle.quantitative_batch(bits, n_cols, y, P, covar=cv)
        want = []
        for s in range(S):
            if ref["filtered"][s]:
                continue
            ap = ",".join(str(v) for v in ref["allele_paths"][spo[s]:spo[s + 1]])
            want.append(prefix(s) + [fmt(ref[k][s], 0) for k in ("p_value", "r2", "beta", "se")] + [ap, info[s][4]])
        got = _rows(os.path.join(d, name))
        assert len(got) == len(want) > 20
        for g, w in zip(got, want):
            assert g[:5] == w[:5] and g[9:] == w[9:] and all(_close(a, b) for a, b in zip(g[5:9], w[5:9])), (g, w)

    # eQTL: one row per (snarl, gene of its window)
    ref = oracle.eqtl_batch(bits, n_cols, expr, P, goff, gidx)
    want = []
    for s in range(S):
        if ref["filtered"][s]:
            continue
        ap = ",".join(str(v) for v in ref["allele_paths"][spo[s]:spo[s + 1]])
        for k in range(int(goff[s]), int(goff[s + 1])):
            want.append(prefix(s) + [f"GENE{gidx[k]}"] + [fmt(ref[q][k], 0) for q in ("p_value", "r2", "beta", "se")] +
                        [ap, info[s][4]])
    got = _rows(os.path.join(d, "eqtl.tsv"))
    assert len(got) == len(want) > 20
    for g, w in zip(got, want):
        assert g[:6] == w[:6] and g[10:] == w[10:] and all(_close(a, b) for a, b in zip(g[6:10], w[6:10])), (g, w)

    # logistic regression (GpuBinaryCovarSnarlAnalyzer): covariates passed, without effect, as in the reference
    ref = oracle.binary_covar_batch(bits, n_cols, is_case, P, covar=covar)
    want = []
    for s in range(S):
        if ref["filtered"][s]:
            continue
        ap = ",".join(str(v) for v in ref["allele_paths"][spo[s]:spo[s + 1]])
        vals = ["NA"] * 3 if ref["code"][s] == 1 else [fmt(ref[k][s], 0) for k in ("p_value", "beta", "se")]
        want.append(prefix(s) + vals + [ap, info[s][4]])
    got = _rows(os.path.join(d, "binary_covar.tsv"))
    assert len(got) == len(want) > 20
    for g, w in zip(got, want):
        assert g[:5] == w[:5] and g[8:] == w[8:] and all(_close(a, b) for a, b in zip(g[5:8], w[5:8])), (g, w)

    # the chunk cut over three contexts on three host threads: the same file
    assert open(os.path.join(d, "binary_multi.tsv")).read() == open(os.path.join(d, "binary.tsv")).read()
    assert open(os.path.join(d, "quantitative_multi.tsv")).read() == open(os.path.join(d, "quantitative.tsv")).read()


def test_device_formatted_columns_equal_the_c_library(gpu_ctx):
    """stoat_gpu_format_values (device) against stoat_gpu_format_value (snprintf "%.4e" / "%.4g", the reference's
    set_precision): every string identical, over magnitudes from denormals to 1e308, exact decimal ties and the codes."""
    from stoat_b200 import format_value

    rng = np.random.default_rng(3)
    v = np.concatenate([
        10.0 ** rng.uniform(-320, 300, 200000) * rng.choice([-1.0, 1.0], 200000, p=[0.2, 0.8]),
        rng.uniform(0, 1, 100000), rng.uniform(0.09, 0.11, 20000), rng.normal(size=50000) * 100,
        np.round(rng.uniform(0, 1e5, 20000)),                       # integers: "%g" drops the zeros, switches at 1e4
        np.array([0.0, -0.0, 1.0, 0.1, 0.09999999999999999, 0.1 - 1e-17, 9.9995e-5, 9.99949999e-5, 0.099995, 0.99995,
                  9999.5, 9999.4999, 99995.0, 1e-300, 5e-324, 1.7976931348623157e308, np.inf, -np.inf, np.nan,
                  0.5, 0.25, 0.125, 0.0625, 0.03125, 0.015625, 1.00005, 1.00015, 123450.0, 1234.5, 12.345, 1.2345e-3]),
        (np.arange(1, 20001) * 2 + 1) / 2.0 * 1e-4,                 # k + 0.5 in the 5th digit: rounding boundaries
    ])
    codes = np.zeros(v.size, np.uint8)
    codes[:40] = np.tile(np.array([0, 1, 2, 3, 4], np.uint8), 8)
    got = gpu_ctx.format_values(v, codes)
    bad = [(float(x), int(c), g, format_value(float(x), int(c))) for x, c, g in zip(v, codes, got)
           if g != format_value(float(x), int(c))]
    assert not bad, f"{len(bad)} of {v.size} strings differ, first: {bad[:5]}"
    assert gpu_ctx.format_values(np.array([0.0123456]), None) == ["1.2346e-02"]
