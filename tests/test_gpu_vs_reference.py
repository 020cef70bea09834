"""The CUDA path against the REFERENCE'S OWN analyzers (oracle/_ref: src/snarl_analyzer.cpp & co. compiled in place, the
prebuilt .so travels to the GPU box).  Rows are compared the way the reference's system test compares TSVs: keyed by
SNARL, order free (tests/system/compare_files_utils.hpp:16-18); tables bit-exact, printed p / beta / se / r2 equal."""
import numpy as np
import pytest

import helpers as H
import stoat_b200

pytestmark = pytest.mark.gpu
ref = pytest.importorskip("oracle.ref")


@pytest.fixture(scope="module")
def refbuild():
    if not ref.available():
        pytest.skip("oracle/_ref was not shipped to this box")
    return ref


def test_binary_and_quantitative_rows_match_reference(gpu_ctx, oracle, refbuild):
    from stoat_b200 import synth

    d = synth.numpy_dataset(1500, 200, seed=13, frac_missing_edge=0.03)   # C1'/C2' shape: ~1,500 snarls x 400 haplotypes
    nodes = ref.nodes_from_structure(d["structure"])
    spo = d["paths"].snarl_path_off
    gpu_ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
    gpu_ctx.set_binary_phenotype(d["is_case"])
    got = gpu_ctx.binary_batch(d["paths"])
    want = ref.run("binary", d["bits"], d["n_rows"], d["n_cols"], spo, nodes, is_case=d["is_case"], n_threads=4)
    assert H.assert_binary_rows(stoat_b200.format_value, got, want["rows"], spo) > 800

    covar = np.random.default_rng(5).normal(size=(200, 3))
    for cv in (None, covar):
        gpu_ctx.set_quantitative_phenotype(d["y"])
        gpu_ctx.set_covariates(cv)
        got = gpu_ctx.quantitative_batch(d["paths"])
        want = ref.run("quantitative", d["bits"], d["n_rows"], d["n_cols"], spo, nodes, y=d["y"], covar=cv, n_threads=4)
        assert H.assert_quant_rows(stoat_b200.format_value, got, want["rows"], spo) > 800
    gpu_ctx.set_covariates(None)


@pytest.mark.parametrize("seed", [0, 1])
def test_adversarial_rows_match_reference(gpu_ctx, oracle, refbuild, seed):
    from stoat_b200 import SnarlPaths

    rng = np.random.default_rng(300 + seed)
    n_samples, n_rows = 70, 48
    dense, snarls = H.random_snarl_problem(rng, 200, n_samples, n_rows)
    spo, peo, er = H.csr_from_lists(snarls)
    bits = oracle.pack_matrix(dense)
    nodes = ref.nodes_from_csr(spo, peo, er, n_rows)
    is_case = (rng.random(n_samples) < 0.5).astype(np.uint8)
    y = rng.normal(size=n_samples)
    paths = SnarlPaths(spo, peo, er)
    gpu_ctx.set_matrix(bits, n_rows, 2 * n_samples)
    gpu_ctx.set_binary_phenotype(is_case)
    got = gpu_ctx.binary_batch(paths)
    want = ref.run("binary", bits, n_rows, 2 * n_samples, spo, nodes, is_case=is_case, n_threads=2)
    H.assert_binary_rows(stoat_b200.format_value, got, want["rows"], spo)
    gpu_ctx.set_quantitative_phenotype(y)
    got = gpu_ctx.quantitative_batch(paths)
    want = ref.run("quantitative", bits, n_rows, 2 * n_samples, spo, nodes, y=y, n_threads=2)
    H.assert_quant_rows(stoat_b200.format_value, got, want["rows"], spo)
