import sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import oracle, stoat_b200
import test_gpu_eqtl as T
d, expr = T._dense_problem(21, 140, 90, 9, frac_three=0.0, frac_node0=0.0, frac_missing_hap=0.03)
st, dense, n = d["structure"], d["dense"], 90
dos = dense[st.row0[3]].reshape(n, 2).sum(1) / 2.0
dos5 = dense[st.row0[5]].reshape(n, 2).sum(1) / 2.0
expr[0] = 2.0 + 3.0 * dos
expr[1] = 1.0 - 2.0 * dos5 + 1e-7 * np.random.default_rng(1).normal(size=n)
expr[2] = 4.25
P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
ref = oracle.eqtl_batch(d["bits"], d["n_cols"], expr, P)
p = ref["p_value"].reshape(140, 9)
for variant in ("all", "drop0", "drop1", "drop2"):
    e = expr.copy()
    if variant == "drop0": e[0] = expr[3]
    if variant == "drop1": e[1] = expr[3]
    if variant == "drop2": e[2] = expr[3]
    r2 = oracle.eqtl_batch(d["bits"], d["n_cols"], e, P)
    pp = r2["p_value"].reshape(140, 9)
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"])
        got = ctx.eqtl_dense_summary(d["paths"], e, p_threshold=1e-4)
    bad = [s for s in range(140) if not r2["filtered"][s] and abs(got["min_p"][s] - np.nanmin(pp[s])) > 1e-6 * np.nanmin(pp[s])]
    print(variant, "bad snarls:", bad[:20], len(bad))
    for s in bad[:3]:
        print("  s", s, "got", got["min_p"][s], got["argmin_gene"][s], "want", pp[s], "n_used", r2["n_used"][s])
