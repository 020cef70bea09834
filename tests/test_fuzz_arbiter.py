"""The arbiter of scripts/fuzz_parity.py (an 80-bit regression built independently of the oracle) against the oracle on
well-conditioned rows: the campaign's verdicts on rows beyond 1e-9 rest on it.  CPU only."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))


def test_extended_precision_regression_agrees_with_the_oracle(oracle):
    import fuzz_parity as F

    checked = 0
    for seed in (1003, 1017, 1042, 1110):
        prob, _ = F.draw_case(seed)
        if prob["n"] < 30:
            continue
        ref = oracle.quantitative_batch(prob["bits"], prob["n_cols"], prob["y"], prob["OP"], covar=prob["covar"], **prob["thr"])
        with np.errstate(invalid="ignore"):
            ok = (ref["filtered"] == 0) & (ref["code"] == 0) & np.isfinite(ref["r2"]) & (ref["r2"] < 1 - 1e-6) & \
                 (ref["se"] > 0) & (ref["n_used"] > 8 + prob["ncov"])
        for s in np.flatnonzero(ok)[:6]:
            ex = F.exact_ols(prob, int(s), prob["y"])
            # the oracle follows the reference's order of operations in double precision: its own rounding error is what
            # separates it from the 80-bit value (1e-7 covers the worst conditioned designs of these seeds)
            assert abs(ex["beta"] - ref["beta"][s]) <= 1e-7 * abs(ref["beta"][s]) + 1e-9 * ref["se"][s], (seed, s)
            assert abs(ex["se"] - ref["se"][s]) <= 1e-7 * ref["se"][s], (seed, s)
            assert abs(ex["p_value"] - ref["p_value"][s]) <= 1e-6 * ref["p_value"][s] + 1e-300, (seed, s)
            checked += 1
    assert checked >= 8
