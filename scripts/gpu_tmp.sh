python - <<'PY' 2>/dev/null
import sys, os, numpy as np
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import oracle, stoat_b200
from stoat_b200 import synth
d = synth.numpy_dataset(120, 33000, seed=14, frac_missing_edge=0.03)
P = oracle.Paths(d["paths"].snarl_path_off, d["paths"].path_edge_off, d["paths"].edge_row)
ref = oracle.binary_covar_batch(d["bits"], d["n_cols"], d["is_case"], P)
m=(ref["filtered"]==0)&(ref["code"]==0)
npaths=np.diff(d["paths"].snarl_path_off.astype(np.int64))
for mode in ("", "general"):
    if mode: os.environ["STOAT_GPU_LOGIT"]=mode
    with stoat_b200.Context(0) as ctx:
        ctx.set_matrix(d["bits"], d["n_rows"], d["n_cols"]); ctx.set_binary_phenotype(d["is_case"])
        got = ctx.binary_covar_batch(d["paths"])
    for k in ("beta","se","p_value"):
        a,b=got[k],ref[k]; e=np.where(m,np.abs(a-b),0); rel=e/np.maximum(np.abs(b),1e-300)
        idx=np.argsort(-rel)[:3]
        print(mode or "fast", k, [(int(i), int(npaths[i]), float(a[i]), float(b[i]), float(rel[i])) for i in idx])
PY
