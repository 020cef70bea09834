#!/usr/bin/env python3
"""Which snarls of the C3 generator at a given seed take the general quantitative path (K3G), and what they look like.
python scripts/diag_general.py 43"""
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["STOAT_GPU_DEBUG_GENERAL"] = "1"
import torch  # noqa: E402
import stoat_b200  # noqa: E402
from stoat_b200 import synth  # noqa: E402

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 43
S, N = 1_000_000, 5_000
d = synth.torch_dataset(S, N, torch.device("cuda:0"), seed=seed)
st = d["structure"]
with stoat_b200.Context(0) as ctx:
    bits = d["bits_dev"].cpu().numpy()
    ctx.set_matrix(bits, d["n_rows"], d["n_cols"])
    ctx.set_quantitative_phenotype(np.asarray(d["y"]))
    sys.stderr.flush()
    saved = os.dup(2)
    tmp = open("/tmp/diag_stderr.txt", "w+")
    os.dup2(tmp.fileno(), 2)
    res = ctx.quantitative_batch(d["paths"])
    os.dup2(saved, 2)
    tmp.seek(0)
    text = tmp.read()
m = re.search(r"general path: (\d+) snarls:([ \d]*)", text)
print("kinds", np.bincount(st.kind).tolist(), "kept", int((np.asarray(res["filtered"]) == 0).sum()))
if not m:
    print("no snarl on the general path;", text[-300:])
    sys.exit(0)
ids = [int(t) for t in m.group(2).split()]
print(m.group(1), "snarls on the general path")
spo = np.asarray(d["paths"].snarl_path_off).astype(np.int64)
for s in ids:
    ap = np.asarray(res["allele_paths"])[spo[s]:spo[s + 1]].tolist()
    print(f"snarl {s}: kind {int(st.kind[s])} freq {st.freq[s]:.5f} n_used {int(res['n_used'][s])} filtered {int(res['filtered'][s])} "
          f"alleles {ap} p {res['p_value'][s]!r} beta {res['beta'][s]!r} se {res['se'][s]!r} r2 {res['r2'][s]!r}")
