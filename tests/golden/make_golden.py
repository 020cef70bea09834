#!/usr/bin/env python3
"""Extract golden vectors for the hot path from the reference's OWN tests and expected outputs.

Run once in the build container (where /root/reference is mounted); the outputs are committed so
that nothing under tests/ needs /root/reference at run time.

Sources (relative to the reference root):
  tests/unittest/binary_test_unit.cpp:7-116              chi2_2x2 / chi2_2xN / fastFishersExactTest strings
  tests/unittest/utils_unit.cpp:9-30                     set_precision strings
  tests/unittest/snarl_analyzer_unit.cpp:100-221         identify_path + filters (transcribed as data)
  tests/unittest/quantitative_test_unit.cpp:15-83        quantitative column surgery (transcribed as data)
  tests/system/graph_simu_test.cpp:165-168,199,336-337   GROUP_PATHS -> P_FISHER,P_CHI2 rows
  tests/expected_output/vcf/binary/binary_table_vcf.tsv  1,463 rows GROUP_PATHS -> P_FISHER,P_CHI2
  tests/expected_output/vcf/quantitative{,_covar}/quantitative_table_vcf.tsv   P,RSQUARE,BETA,SE,ALLELE_PATHS
  data/quantitative/phenotype.tsv, covariate.tsv         the phenotype those rows were computed on
"""
import csv
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def read(p):
    with open(os.path.join(REF, p)) as f:
        return f.read()


def binary_unit():
    """Parse SECTION blocks of binary_test_unit.cpp into {g0, g1, chi2?, fisher?, chi2_2xN?}."""
    src = read("tests/unittest/binary_test_unit.cpp")
    cases = []
    for m in re.finditer(r'SECTION\("([^"]*)"\)\s*\{(.*?)\n    \}', src, re.S):
        name, body = m.group(1), m.group(2)
        g0 = [int(x) for x in re.search(r"g0 = \{([^}]*)\}", body).group(1).split(",")]
        g1 = [int(x) for x in re.search(r"g1 = \{([^}]*)\}", body).group(1).split(",")]
        case = {"name": name, "g0": g0, "g1": g1}
        for fn, key in (("chi2_2x2", "chi2_2x2"), ("fastFishersExactTest", "fisher"), ("chi2_2xN", "chi2_2xN")):
            mm = re.search(r'REQUIRE\(fk\.%s\([^)]*\)\s*==\s*"([^"]*)"\)' % fn, body)
            if mm:
                case[key] = mm.group(1)
        cases.append(case)
    return cases


def set_precision_unit():
    src = read("tests/unittest/utils_unit.cpp")
    out = []
    for m in re.finditer(r'REQUIRE\(set_precision\(([^)]*)\)\s*==\s*"([^"]*)"\)', src):
        out.append({"value": float(m.group(1)), "expect": m.group(2)})
    return out


def graph_rows():
    src = read("tests/system/graph_simu_test.cpp")
    rows = set()
    for m in re.finditer(r'emplace_back\("path0\\t[^"]*"\)', src):
        f = m.group(0)[len('emplace_back("'):-2].split("\\t")
        # CHR START END SNARL PATH_LENGTHS P_FISHER P_CHI2 GROUP_PATHS DEPTH
        if f[7] != "NA":
            rows.add((f[7], f[5], f[6]))
    return [{"group_paths": g, "p_fisher": pf, "p_chi2": pc} for g, pf, pc in sorted(rows)]


def tsv_columns(path, cols):
    out = []
    with open(os.path.join(REF, path)) as f:
        rd = csv.reader(f, delimiter="\t")
        header = next(rd)
        header[0] = header[0].lstrip("#")
        idx = [header.index(c) for c in cols]
        for row in rd:
            out.append([row[i] for i in idx])
    return out


def write_tsv(name, header, rows):
    with open(os.path.join(OUT, name), "w") as f:
        f.write("\t".join(header) + "\n")
        for r in rows:
            f.write("\t".join(r) + "\n")


def main():
    unit = {
        "source": "tests/unittest/binary_test_unit.cpp:7-116, tests/unittest/utils_unit.cpp:9-30, "
                  "tests/system/graph_simu_test.cpp:165-168,199,336-337,367",
        "binary_unit": binary_unit(),
        "set_precision": set_precision_unit(),
        "graph_rows": graph_rows(),
        # transcribed by hand from tests/unittest/snarl_analyzer_unit.cpp:100-221 (inputs -> expected)
        "identify_path": {"n_rows": 2, "n_cols": 3, "set": [[0, 0], [1, 0], [0, 2], [1, 2]],
                          "path_rows": [0, 1], "expect": [0, 2]},
        "filtration_binary": [
            {"g0": [5, 1], "g1": [1, 5], "total": 12, "min_ind": 2, "min_hap": 4, "maf": 0.1, "expect": False},
            {"g0": [5, 1], "g1": [1, 5], "total": 2, "min_ind": 2, "min_hap": 4, "maf": 0.1, "expect": True},
            {"g0": [1, 1], "g1": [1, 1], "total": 4, "min_ind": 2, "min_hap": 10, "maf": 0.1, "expect": True},
            {"g0": [10, 1], "g1": [0, 10], "total": 21, "min_ind": 2, "min_hap": 4, "maf": 0.45, "expect": True},
        ],
        "remove_empty_columns_binary": {"g0": [1, 0, 2, 0], "g1": [0, 0, 1, 0], "expect_g0": [1, 2], "expect_g1": [0, 1]},
        "filtration_quant": [
            {"df": [[1.0, 1.0], [0.5, 0.5]], "min_ind": 2, "min_hap": 2, "maf": 0.1, "expect": False},
            {"df": [[1.0, 1.0]], "min_ind": 2, "min_hap": 2, "maf": 0.1, "expect": True},
            {"df": [[0.1, 0.1], [0.1, 0.1]], "min_ind": 2, "min_hap": 2, "maf": 0.1, "expect": True},
            {"df": [[1.9, 0.1], [1.9, 0.1]], "min_ind": 2, "min_hap": 2, "maf": 0.4, "expect": True},
            # tests/unittest/quantitative_test_unit.cpp:38-60
            {"df": [[0.5, 0], [0.5, 0], [1, 0], [1, 0], [1, 0], [1, 0], [1, 0], [1, 0], [1, 0], [0, 1]],
             "min_ind": 3, "min_hap": 5, "maf": 0.01, "expect": False},
            {"df": [[0.5, 0], [0.5, 0], [1, 0], [1, 0], [1, 0], [1, 0], [1, 0], [1, 0], [1, 0], [0, 1]],
             "min_ind": 3, "min_hap": 5, "maf": 0.2, "expect": True},
        ],
        "remove_empty_columns_quant": [
            {"df": [[1.0, 0.0, 0.0, 0.0], [0.5, 0.0, 0.5, 0.0]], "expect": [[1.0, 0.0], [0.5, 0.5]]},
            {"df": [[0.5, 0, 0.5], [0.5, 0, 0.5], [1, 0, 0], [1, 0, 0], [0, 0, 1]],
             "expect": [[0.5, 0.5], [0.5, 0.5], [1, 0], [1, 0], [0, 1]]},
        ],
    }
    with open(os.path.join(OUT, "unit_known_answers.json"), "w") as f:
        json.dump(unit, f, indent=1)

    # every column: tests/test_gpu_shim.py reproduces the file byte for byte through the C++ host shim
    bcols = ["CHR", "START_POS", "END_POS", "SNARL", "PATH_LENGTHS", "P_FISHER", "P_CHI2", "GROUP_PATHS", "DEPTH"]
    rows = tsv_columns("tests/expected_output/vcf/binary/binary_table_vcf.tsv", bcols)
    write_tsv("binary_table_vcf.golden.tsv", ["#CHR"] + bcols[1:], rows)

    for sub in ("quantitative", "quantitative_covar"):
        rows = tsv_columns("tests/expected_output/vcf/%s/quantitative_table_vcf.tsv" % sub,
                           ["SNARL", "P", "RSQUARE", "BETA", "SE", "ALLELE_PATHS"])
        write_tsv("%s_table_vcf.golden.tsv" % sub, ["SNARL", "P", "RSQUARE", "BETA", "SE", "ALLELE_PATHS"], rows)

    # phenotype / covariates of the quantitative data set (200 samples), for the C2' shaped tests
    ph = tsv_columns("data/quantitative/phenotype.tsv", ["IID", "PHENO"])
    write_tsv("quantitative_phenotype.tsv", ["IID", "PHENO"], ph)
    cv = tsv_columns("data/quantitative/covariate.tsv", ["IID", "SEX", "PC1", "PC2", "PC3"])
    write_tsv("quantitative_covariate.tsv", ["IID", "SEX", "PC1", "PC2", "PC3"], cv)
    bp = tsv_columns("data/binary/phenotype.tsv", ["IID", "PHENO"])
    write_tsv("binary_phenotype.tsv", ["IID", "PHENO"], bp)
    print("wrote golden vectors to", OUT)


if __name__ == "__main__":
    main()
