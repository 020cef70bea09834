"""stoat_gpu_bh_adjust_static: the Benjamini-Hochberg adjustment with its scratch in a file-scope device array of the
library (no cudaMalloc; no counterpart in the reference), called through the
C-ABI and compared with the existing entry point stoat_gpu_bh_adjust — bit-exact, as tests/test_gpu_post.py compares that
one with the oracle.  Collected last on purpose (file name).

The context is this file's own (not the session fixture), so the test is the same alone in a fresh process and after the
rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ctx():
    import stoat_b200

    c = stoat_b200.Context(0)
    yield c
    c.close()


def _column(n, seed):
    rng = np.random.default_rng(seed)
    p = rng.random(n) ** 4
    p[rng.integers(0, n, max(1, n // 20))] = 0.0
    p[rng.integers(0, n, max(1, n // 20))] = 1.0
    tied = rng.integers(0, n, max(1, n // 10))
    p[tied] = p[tied[0]]
    p[rng.integers(0, n, 3)] = 5e-324
    return p


def test_static_scratch_matches_allocating_entry_point_on_two_calls_in_a_row(ctx):
    # two calls in a row on the SAME static scratch: a long column, then a shorter one that must not see what the first
    # left behind (sorted keys, scan arrays, the validity flag), then the long one again
    first, second = _column(300_000, 11), _column(4_097, 12)
    want_first, want_second = ctx.bh_adjust(first), ctx.bh_adjust(second)
    got_first = ctx.bh_adjust_static(first)
    got_second = ctx.bh_adjust_static(second)
    np.testing.assert_array_equal(got_first, want_first)
    np.testing.assert_array_equal(got_second, want_second)
    np.testing.assert_array_equal(ctx.bh_adjust_static(first), want_first)
    assert ctx.bh_adjust_static([]).size == 0
    assert ctx.bh_adjust_static([0.25])[0] == 0.25


def test_static_scratch_at_its_maximum_and_beyond(ctx):
    import stoat_b200

    n = ctx.BH_STATIC_MAX_N
    p = _column(n, 13)
    np.testing.assert_array_equal(ctx.bh_adjust_static(p), ctx.bh_adjust(p))
    with pytest.raises(stoat_b200.StoatGpuError, match="exceeds the static scratch"):
        ctx.bh_adjust_static(np.full(n + 1, 0.5))
    # a rejected call leaves the entry point usable
    np.testing.assert_array_equal(ctx.bh_adjust_static(p[:1000]), ctx.bh_adjust(p[:1000]))


def test_static_scratch_validates_on_the_device_and_clears_its_flag(ctx):
    import stoat_b200

    p = _column(5000, 14)
    for bad in (-0.1, float("nan"), float("inf")):
        q = p.copy()
        q[1234] = bad
        out = np.full_like(q, -7.0)
        rc = ctx._lib.stoat_gpu_bh_adjust_static(ctx._ctx, C.c_void_p(q.ctypes.data), q.size, C.c_void_p(out.ctypes.data))
        assert rc != 0
        assert (out == -7.0).all()  # nothing written
        with pytest.raises(stoat_b200.StoatGpuError, match="finite and >= 0"):
            ctx.bh_adjust_static(q)
        # the flag of the failed call must not leak into the next one
        np.testing.assert_array_equal(ctx.bh_adjust_static(p), ctx.bh_adjust(p))


def test_static_scratch_with_device_pointers(ctx):
    torch = pytest.importorskip("torch")
    p = _column(70_000, 15)
    want = ctx.bh_adjust(p)
    d_p = torch.as_tensor(p, device="cuda:0")
    d_out = torch.empty_like(d_p)
    torch.cuda.synchronize()
    for _ in range(2):
        d_out.zero_()
        torch.cuda.synchronize()
        ctx.bh_adjust_static_dev(d_p.data_ptr(), p.size, d_out.data_ptr())
        np.testing.assert_array_equal(d_out.cpu().numpy(), want)
