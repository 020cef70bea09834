"""The oracle's own tail functions (its stand-ins for Boost.Math) against mpmath at 60 digits, and its OLS
against an independent numpy / scipy evaluation (CPU only)."""
import numpy as np
import pytest

mp = pytest.importorskip("mpmath")


def test_gamma_q_against_mpmath(oracle):
    mp.mp.dps = 60
    rng = np.random.default_rng(1)
    worst = 0
    for df in (1, 2, 3, 4, 5, 7, 10, 63):
        for x in list(rng.uniform(1e-3, 84, 25)) + list(rng.uniform(85.01, 1200, 25)):
            q = mp.gammainc(mp.mpf(df) / 2, mp.mpf(x) / 2, mp.inf, regularized=True)
            got = oracle.gamma_q(df / 2, x / 2)
            worst = max(worst, abs((mp.mpf(got) - q) / q))
    assert worst < 1e-14


def test_chi2_pvalue_follows_reference_rule(oracle):
    # stat <= 85: 1.0 - cdf in double (multiples of 2^-53); > 85: the true upper tail
    mp.mp.dps = 60
    for stat in (70.0, 80.0, 84.9):
        p = oracle.chi2_pvalue(stat, 1)
        assert p == 0.0 or p >= 2.0 ** -53
        assert (p / 2.0 ** -53) == round(p / 2.0 ** -53)
    true = mp.gammainc(mp.mpf(1) / 2, mp.mpf(90) / 2, mp.inf, regularized=True)
    assert abs(mp.mpf(oracle.chi2_pvalue(90.0, 1)) - true) / true < 1e-14


def test_student_t_against_mpmath(oracle):
    mp.mp.dps = 60
    rng = np.random.default_rng(2)
    worst = 0
    for df in (1, 2, 3, 10, 50, 199, 4998, 24999):
        for t in list(rng.uniform(0, 3, 12)) + list(rng.uniform(3, 30, 12)) + [1e-8]:
            z = mp.mpf(df) / (mp.mpf(df) + mp.mpf(t) ** 2)
            sf = mp.betainc(mp.mpf(df) / 2, mp.mpf(1) / 2, 0, z, regularized=True) / 2
            if sf < mp.mpf("1e-290"):
                continue
            worst = max(worst, abs((mp.mpf(oracle.student_t_sf(df, t)) - sf) / sf))
    assert worst < 1e-13


def test_ols_against_numpy(oracle):
    stats = pytest.importorskip("scipy.stats")
    rng = np.random.default_rng(3)
    done = 0
    for _ in range(150):
        n, nv, nc = int(rng.integers(5, 80)), int(rng.integers(1, 4)), int(rng.integers(0, 4))
        df_ = rng.choice([0, 0.5, 1.0], size=(n, nv))
        y = rng.normal(size=n) + df_[:, 0]
        cov = rng.normal(size=(n, nc))
        X = np.column_stack([np.ones(n), df_] + ([cov] if nc else []))
        if np.linalg.matrix_rank(X) < X.shape[1] or n - X.shape[1] < 2:
            continue
        r = oracle.linear_regression(df_, y, cov if nc else None)
        inv = np.linalg.inv(X.T @ X)
        beta = inv @ X.T @ y
        res = y - X @ beta
        rss = res @ res
        dfr = max(n - X.shape[1] + 1, 1)  # the reference's n - p + 1 (stats_test.cpp:467)
        se = np.sqrt(inv[1, 1] * rss / dfr)
        p = 2 * stats.t.sf(abs(beta[1] / se), dfr)
        r2 = 1 - rss / ((y - y.mean()) @ (y - y.mean()))
        for k, v in (("p", p), ("beta", beta[1]), ("se", se), ("r2", r2)):
            assert abs(r[k] - v) <= 1e-9 * abs(v) + 1e-13, (k, r[k], v)
        done += 1
    assert done > 80


def test_rank_deficient_goes_through_pseudo_inverse(oracle):
    # a constant dosage column is collinear with the intercept: |D_i| < 1e-10 => pinv (stats_test.cpp:408-416)
    n = 12
    df_ = np.full((n, 1), 0.5)
    y = np.arange(n, dtype=float)
    X = np.column_stack([np.ones(n), df_])
    inv, used_pinv = oracle.xtx_inverse(X.T @ X)
    assert used_pinv
    np.testing.assert_allclose(inv, np.linalg.pinv(X.T @ X, rcond=1e-6 / np.linalg.norm(X.T @ X, 2)), rtol=1e-9, atol=1e-12)
    r = oracle.linear_regression(df_, y, None)
    assert np.isfinite(r["beta"])
    # zero regression columns: the reference indexes out of bounds; the oracle reports it
    r0 = oracle.linear_regression(np.zeros((n, 0)), y, None)
    assert r0["code"] == oracle.P_UNDEF
