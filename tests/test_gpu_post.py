"""Benjamini-Hochberg adjustment on the device (stoat_gpu_bh_adjust) against the oracle's restatement of
adjust_pvalues_with_BH (src/post_processing.cpp:7-38) and the reference's own unit-test vectors."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_bh_known_answers(gpu_ctx):
    # tests/unittest/post_processing_unit.cpp:12-46
    got = gpu_ctx.bh_adjust([0.01, 0.04, 0.03, 0.002, 0.05])
    np.testing.assert_allclose(got, [0.025, 0.05, 0.05, 0.01, 0.05], rtol=1e-12)
    assert (gpu_ctx.bh_adjust([0.9, 0.95, 0.99]) <= 1.0).all()  # :54-66
    assert gpu_ctx.bh_adjust([]).size == 0  # :48-52
    assert gpu_ctx.bh_adjust([0.25])[0] == 0.25


@pytest.mark.parametrize("n,seed", [(1000, 1), (65537, 2), (1_000_000, 3)])
def test_bh_bit_exact_against_oracle(gpu_ctx, oracle, n, seed):
    rng = np.random.default_rng(seed)
    p = rng.random(n) ** 4  # many small values
    p[rng.integers(0, n, n // 20)] = 0.0                       # exact zeros ("0" rows)
    p[rng.integers(0, n, n // 20)] = 1.0
    tied = rng.integers(0, n, n // 10)
    p[tied] = p[tied[0]]                                        # a large tie group
    p[rng.integers(0, n, 5)] = 5e-324                           # denormal
    got = gpu_ctx.bh_adjust(p)
    np.testing.assert_array_equal(got, oracle.bh_adjust(p))


def test_bh_rejects_nan(gpu_ctx):
    import stoat_b200

    with pytest.raises(stoat_b200.StoatGpuError):
        gpu_ctx.bh_adjust([0.1, float("nan")])
