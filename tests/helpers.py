"""Shared builders for the parity tests: seeded inputs that both the CUDA path and the oracle consume."""
from __future__ import annotations

import csv
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MISSING = 0xFFFFFFFF


def parse_group_paths(gp: str):
    cols = [c.split(":") for c in gp.split(",")]
    return [int(a) for a, _ in cols], [int(b) for _, b in cols]


def golden_binary_rows():
    with open(os.path.join(GOLDEN, "binary_table_vcf.golden.tsv")) as f:
        return list(csv.DictReader(f, delimiter="\t"))


def golden_quant_rows(name="quantitative"):
    with open(os.path.join(GOLDEN, f"{name}_table_vcf.golden.tsv")) as f:
        return list(csv.DictReader(f, delimiter="\t"))


def golden_binary_phenotype():
    """data/binary/phenotype.tsv: 1 = control -> 0, 2 = case -> 1 (src/arg_parser.cpp:20-95)."""
    with open(os.path.join(GOLDEN, "binary_phenotype.tsv")) as f:
        rows = list(csv.DictReader(f, delimiter="\t"))
    return np.array([1 if r["PHENO"] == "2" else 0 for r in rows], dtype=np.uint8)


def golden_quant_phenotype():
    with open(os.path.join(GOLDEN, "quantitative_phenotype.tsv")) as f:
        rows = list(csv.DictReader(f, delimiter="\t"))
    return np.array([float(r["PHENO"]) for r in rows])


def golden_quant_covariates():
    with open(os.path.join(GOLDEN, "quantitative_covariate.tsv")) as f:
        rows = list(csv.DictReader(f, delimiter="\t"))
    return np.array([[float(r[k]) for k in ("SEX", "PC1", "PC2", "PC3")] for r in rows])


def csr_from_lists(snarls):
    """snarls: list of snarls, each a list of paths, each a list of row indices -> (spo, peo, edge_row)."""
    spo, peo, rows = [0], [0], []
    for paths in snarls:
        for p in paths:
            rows.extend(p)
            peo.append(len(rows))
        spo.append(len(peo) - 1)
    return (np.array(spo, np.uint64), np.array(peo, np.uint64), np.array(rows, np.uint32))


def matrix_realising_tables(tables, is_case):
    """One single-edge path per table column.  tables: list of (g0 list, g1 list).  Returns a dense bool
    matrix [rows, H] in which path j of table t is matched by exactly g0[j] control and g1[j] case
    haplotypes (disjoint between the columns of a table), and the snarl lists for csr_from_lists."""
    is_case_hap = np.repeat(is_case.astype(bool), 2)
    ctrl = np.flatnonzero(~is_case_hap)
    case = np.flatnonzero(is_case_hap)
    H = is_case_hap.size
    n_rows = sum(len(g0) for g0, _ in tables)
    dense = np.zeros((n_rows, H), dtype=bool)
    snarls, r = [], 0
    for g0, g1 in tables:
        assert sum(g0) <= ctrl.size and sum(g1) <= case.size, "table does not fit the phenotype split"
        c0 = c1 = 0
        paths = []
        for a, b in zip(g0, g1):
            dense[r, ctrl[c0:c0 + a]] = True
            dense[r, case[c1:c1 + b]] = True
            c0 += a
            c1 += b
            paths.append([r])
            r += 1
        snarls.append(paths)
    return dense, snarls


def random_snarl_problem(rng, n_snarls, n_samples, n_rows, density=0.5, max_paths=5, max_edges=4,
                         p_missing_edge=0.05, p_empty_path=0.03, p_zero_row=0.1, correlated=True):
    """Adversarial small problem: random rows shared between paths / snarls, empty edge lists, missing edges,
    all-zero rows, haplotypes matching several paths (non-dyadic dosages).  Dense bool + snarl lists."""
    H = 2 * n_samples
    dens = rng.uniform(0.05, 0.98, size=(n_rows, 1)) if correlated else density
    dense = rng.random((n_rows, H)) < dens
    dense[rng.random(n_rows) < p_zero_row] = False
    snarls = []
    for _ in range(n_snarls):
        npth = int(rng.integers(1, max_paths + 1))
        paths = []
        for _ in range(npth):
            if rng.random() < p_empty_path:
                paths.append([])
                continue
            ne = int(rng.integers(1, max_edges + 1))
            p = [int(x) for x in rng.integers(0, n_rows, size=ne)]
            if rng.random() < p_missing_edge:
                p[int(rng.integers(0, ne))] = MISSING
            paths.append(p)
        snarls.append(paths)
    return dense, snarls


def partition_problem(rng, n_snarls, n_samples, max_paths=4, p_missing_hap=0.05):
    """Realistic snarls: every haplotype takes exactly one path (or none), 1..2 private rows per path."""
    H = 2 * n_samples
    rows, snarls = [], []
    for _ in range(n_snarls):
        npth = int(rng.integers(2, max_paths + 1))
        probs = rng.dirichlet(np.ones(npth) * rng.uniform(0.3, 3.0))
        allele = rng.choice(npth, size=H, p=probs)
        allele[rng.random(H) < p_missing_hap] = -1
        paths = []
        for j in range(npth):
            ne = int(rng.integers(1, 3))
            p = []
            for _ in range(ne):
                rows.append(allele == j)
                p.append(len(rows) - 1)
            paths.append(p)
        snarls.append(paths)
    return np.array(rows, dtype=bool).reshape(len(rows), H), snarls


def assert_binary_equal(got: dict, ref: dict, rtol=1e-9):
    """Bit-exact tables / flags / codes; p-values within rtol relative plus the 2^-53 quantum that the
    reference's own `1.0 - cdf` subtraction carries."""
    for k in ("g0", "g1", "total_sum", "filtered"):
        np.testing.assert_array_equal(got[k], ref[k], err_msg=k)
    keep = ref["filtered"] == 0
    for k in ("chi2_code", "fisher_code"):
        np.testing.assert_array_equal(got[k][keep], ref[k][keep], err_msg=k)
    q = 2.0 ** -52
    for pk, ck in (("p_chi2", "chi2_code"), ("p_fisher", "fisher_code")):
        m = keep & (ref[ck] == 0)
        a, b = got[pk][m], ref[pk][m]
        err = np.abs(a - b)
        tol = rtol * np.abs(b) + (q if pk == "p_chi2" else 0.0)
        bad = np.flatnonzero(~(err <= tol))
        assert bad.size == 0, f"{pk}: {bad.size} mismatches, first got {a[bad[:3]]} want {b[bad[:3]]}"


def assert_quant_equal(got: dict, ref: dict, rtol=1e-9):
    """Integer outputs bit-exact; p / beta / se within 1e-9 relative; r2 (= 1 - rss/tss, a difference of O(1)
    numbers) and beta additionally get an absolute 1e-12 floor for values that are themselves ~0."""
    atols = dict(p_value=1e-300, r2=1e-12, beta=1e-12, se=1e-300)
    for k in ("allele_paths", "filtered", "n_used"):
        np.testing.assert_array_equal(got[k], ref[k], err_msg=k)
    keep = ref["filtered"] == 0
    np.testing.assert_array_equal(got["code"][keep], ref["code"][keep], err_msg="code")
    m = keep & (ref["code"] == 0)
    for k in ("p_value", "r2", "beta", "se"):
        a, b = got[k][m], ref[k][m]
        both_nan = np.isnan(a) & np.isnan(b)
        err = np.abs(a - b)
        ok = both_nan | (err <= rtol * np.abs(b) + atols[k])
        bad = np.flatnonzero(~ok)
        assert bad.size == 0, f"{k}: {bad.size} mismatches, first got {a[bad[:3]]} want {b[bad[:3]]} idx {np.flatnonzero(m)[bad[:3]]}"


# ---------------------------------------------------------------------------------------------------
# result arrays (oracle or CUDA path) against TSV rows written by the reference's own analyzers (oracle/_ref)
# ---------------------------------------------------------------------------------------------------
def printed_close(a: str, b: str, rel=2e-4) -> bool:
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


def assert_binary_rows(fmt, res: dict, ref_rows: dict, snarl_path_off) -> int:
    """fmt(value, code) -> the reference's string.  Tables bit-exact, filter decisions identical, p-values as printed."""
    spo = np.asarray(snarl_path_off, np.int64)
    kept = np.flatnonzero(res["filtered"] == 0)
    assert sorted(ref_rows) == list(kept), "filter decisions differ"
    for s in kept:
        (row,) = ref_rows[int(s)]
        g0, g1 = res["g0"][spo[s]:spo[s + 1]], res["g1"][spo[s]:spo[s + 1]]
        keep = (g0 + g1) > 0
        gp = ",".join(f"{a}:{b}" for a, b in zip(g0[keep], g1[keep]))
        assert row["GROUP_PATHS"] == gp, (s, row, gp)
        pf, pc = fmt(res["p_fisher"][s], res["fisher_code"][s]), fmt(res["p_chi2"][s], res["chi2_code"][s])
        assert printed_close(row["P_FISHER"], pf) and printed_close(row["P_CHI2"], pc), (s, row, pf, pc)
    return kept.size


def assert_quant_rows(fmt, res: dict, ref_rows: dict, snarl_path_off) -> int:
    spo = np.asarray(snarl_path_off, np.int64)
    kept = np.flatnonzero(res["filtered"] == 0)
    assert sorted(ref_rows) == list(kept), "filter decisions differ"
    for s in kept:
        (row,) = ref_rows[int(s)]
        assert row["ALLELE_PATHS"] == ",".join(str(v) for v in res["allele_paths"][spo[s]:spo[s + 1]]), (s, row)
        for col, key in (("P", "p_value"), ("RSQUARE", "r2"), ("BETA", "beta"), ("SE", "se")):
            assert printed_close(row[col], fmt(res[key][s], 0)), (s, col, row, res[key][s])
    return kept.size
