"""ctypes front end of oracle/_ref/libstoat_ref.so — the REFERENCE'S OWN hot-path code (TEST INFRASTRUCTURE).

Built by oracle/ref_build.py in the container that mounts /root/reference; the .so travels to the GPU box with the
snapshot (oracle/_ref/ is git-ignored, not gpurun-ignored).  Used by tests/ to validate the oracle restatement (and the
CUDA path) against the reference itself, and by bench.py as the `"kind": "reference"` CPU baseline.  Never imported by
the product (stoat_b200/).
"""
from __future__ import annotations

import csv
import ctypes as C
import os
import tempfile

import numpy as np

from . import ref_build

MISSING = 0xFFFFFFFF
_lib = None


def available() -> bool:
    """True when the reference build exists (or can be made: /root/reference mounted)."""
    return os.path.exists(ref_build.LIB) or ref_build.available()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if ref_build.available():
            ref_build.build()
        if not os.path.exists(ref_build.LIB):
            raise RuntimeError("oracle/_ref is not built (needs /root/reference; run python oracle/ref_build.py)")
        L = C.CDLL(ref_build.LIB)
        u8p, u64p, f64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint64, C.c_double))
        L.ref_run.restype = C.c_double
        L.ref_run.argtypes = [C.c_int, u8p, C.c_uint64, C.c_uint64, u64p, C.c_uint64, u64p, u64p, u64p, u8p, u64p, u8p, f64p,
                              f64p, C.c_uint64, f64p, C.c_uint64, u64p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_double,
                              C.c_int, C.c_char_p, u64p]
        L.ref_linear_regression.argtypes = [f64p, C.c_uint64, C.c_uint64, f64p, f64p, C.c_uint64, C.c_char_p, C.c_uint64]
        L.ref_logistic_regression.argtypes = [f64p, C.c_uint64, C.c_uint64, u8p, f64p, C.c_uint64, C.c_char_p, C.c_uint64]
        L.ref_fisher_khi2.argtypes = [u64p, u64p, C.c_uint64, C.c_char_p, C.c_uint64]
        L.ref_chi2_2x2.argtypes = [C.c_uint64] * 4 + [C.c_char_p, C.c_uint64]
        L.ref_fisher_2x2.argtypes = [C.c_uint64] * 4 + [C.c_char_p, C.c_uint64]
        L.ref_set_precision.argtypes = [C.c_double, C.c_char_p, C.c_uint64]
        _lib = L
    return _lib


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def _str(fn, *args) -> str:
    buf = C.create_string_buffer(512)
    n = fn(*args, buf, 512)
    if n < 0:
        raise RuntimeError("oracle/_ref: string buffer too small")
    return buf.value.decode()


def set_precision(v: float) -> str:
    return _str(lib().ref_set_precision, float(v))


def chi2_2x2(a, b, c, d) -> str:
    return _str(lib().ref_chi2_2x2, int(a), int(b), int(c), int(d))


def fisher_2x2(a, b, c, d) -> str:
    return _str(lib().ref_fisher_2x2, int(a), int(b), int(c), int(d))


def fisher_khi2(g0, g1):
    """-> (chi2 string, fisher string): FisherKhi2::fisher_khi2 (src/stats_test.cpp:362-380)."""
    a = np.ascontiguousarray(g0, np.uint64)
    b = np.ascontiguousarray(g1, np.uint64)
    s = _str(lib().ref_fisher_khi2, _p(a, C.c_uint64), _p(b, C.c_uint64), a.size)
    chi2, fisher = s.split("\t")
    return chi2, fisher


def linear_regression(df, y, covar=None):
    """-> (p, beta, se, r2) strings: LinearRegression::linear_regression (src/stats_test.cpp:424-506)."""
    df = np.ascontiguousarray(df, np.float64)
    y = np.ascontiguousarray(y, np.float64)
    cv = None if covar is None else np.ascontiguousarray(covar, np.float64)
    s = _str(lib().ref_linear_regression, _p(df, C.c_double), df.shape[0], df.shape[1], _p(y, C.c_double),
             _p(cv, C.c_double), 0 if cv is None else cv.shape[1])
    return tuple(s.split("\t"))


def logistic_regression(df, y, covar=None):
    """-> (p, beta, se) strings: LogisticRegression::logistic_regression (src/stats_test.cpp:49-176)."""
    df = np.ascontiguousarray(df, np.float64)
    y = np.ascontiguousarray(y, np.uint8)
    cv = None if covar is None else np.ascontiguousarray(covar, np.float64)
    s = _str(lib().ref_logistic_regression, _p(df, C.c_double), df.shape[0], df.shape[1], _p(y, C.c_uint8),
             _p(cv, C.c_double), 0 if cv is None else cv.shape[1])
    return tuple(s.split("\t"))


# ---------------------------------------------------------------------------------------------------
# CSR (edge rows) -> node paths + edge header, the form the reference's analyzers consume
# ---------------------------------------------------------------------------------------------------
def nodes_from_csr(snarl_path_off, path_edge_off, edge_row, n_rows):
    """Any CSR.  Row r is the edge >(2r+1)>(2r+2); a path lists its rows' node pairs separated by node 0, whose edges
    identify_path skips (src/snarl_analyzer.cpp:328-330); a missing edge is a node pair absent from the header; a path
    without checked edges is the single node 0."""
    peo = np.asarray(path_edge_off, np.int64)
    er = np.asarray(edge_row, np.int64)
    miss_a = 2 * n_rows + 1
    ids, off = [], [0]
    for p in range(peo.size - 1):
        rows = er[peo[p]:peo[p + 1]]
        if rows.size == 0:
            ids.append(0)
        for k, r in enumerate(rows):
            if k:
                ids.append(0)
            if r == MISSING:
                ids += [miss_a, miss_a + 1]
            else:
                ids += [2 * int(r) + 1, 2 * int(r) + 2]
        off.append(len(ids))
    node_id = np.array(ids, np.uint64)
    edge_desc = np.zeros((n_rows, 4), np.uint64)
    edge_desc[:, 0] = 2 * np.arange(n_rows, dtype=np.uint64) + 1
    edge_desc[:, 2] = 2 * np.arange(n_rows, dtype=np.uint64) + 2
    return dict(path_node_off=np.array(off, np.uint64), node_id=node_id, node_rev=np.zeros(node_id.size, np.uint8),
                edge_desc=edge_desc)


def nodes_from_structure(st):
    """The natural node paths of stoat_b200.synth.make_structure (>a>b>d, >a>c>d, >a>e>d; node-0 alt path >a>0>d;
    missing-edge alt path >a>c>x with x not in the matrix): what `stoat vcf` would read from a paths file.  Vectorised."""
    S = st.n_snarls
    spo = st.paths.snarl_path_off.astype(np.int64)
    n_paths_of = np.diff(spo)
    P = int(spo[-1])
    snarl_of = np.repeat(np.arange(S), n_paths_of)
    j = np.arange(P) - np.repeat(spo[:-1], n_paths_of)
    base = 6 * snarl_of + 1                      # a = base, b.. = base + 1 + j, d = base + 4... see below
    kind = st.kind[snarl_of]
    a, d = base, base + 4
    mid = base + 1 + j                           # b (j=0), c (j=1), e (j=2)
    mid = np.where((kind == 2) & (j == 1), 0, mid)
    last = np.where((kind == 3) & (j == 1), base + 5, d)
    node_id = np.stack([a, mid, last], axis=1).reshape(-1).astype(np.uint64)
    path_node_off = (3 * np.arange(P + 1)).astype(np.uint64)
    # header: row0 + 2j + k  <->  edge k of path j
    edge_desc = np.zeros((st.n_rows, 4), np.uint64)
    for jj in range(3):
        has = np.flatnonzero(st.n_rows_of >= 2 * jj + 2)
        r = st.row0[has] + 2 * jj
        b0 = (6 * has + 1).astype(np.uint64)
        edge_desc[r, 0], edge_desc[r, 2] = b0, b0 + 1 + jj
        edge_desc[r + 1, 0], edge_desc[r + 1, 2] = b0 + 1 + jj, b0 + 4
    return dict(path_node_off=path_node_off, node_id=node_id, node_rev=np.zeros(node_id.size, np.uint8), edge_desc=edge_desc)


# ---------------------------------------------------------------------------------------------------
# the analyzers
# ---------------------------------------------------------------------------------------------------
def run(trait: str, bits, n_rows, n_cols, snarl_path_off, nodes: dict, *, is_case=None, y=None, covar=None, expression=None,
        gene_pos=None, snarl_pos=None, window=0, min_ind=3, min_hap=5, maf=0.05, n_threads=0, tsv_path=None, parse=True):
    """Run the reference's own analyzer loop.  -> dict(seconds, n_filtered, rows: {snarl index: [row dicts]})."""
    L = lib()
    t = {"binary": 0, "quantitative": 1, "eqtl": 2, "binary_covar": 3}[trait]
    bits = np.ascontiguousarray(bits, np.uint8)
    need = (int(n_rows) * int(n_cols) + 7) // 8
    assert bits.size >= need
    spo = np.ascontiguousarray(snarl_path_off, np.uint64)
    S = spo.size - 1
    keep = [np.ascontiguousarray(nodes[k], dt) for k, dt in (("edge_desc", np.uint64), ("path_node_off", np.uint64),
                                                             ("node_id", np.uint64), ("node_rev", np.uint8))]
    ic = None if is_case is None else np.ascontiguousarray(is_case, np.uint8)
    yy = None if y is None else np.ascontiguousarray(y, np.float64)
    cv = None if covar is None else np.ascontiguousarray(covar, np.float64)
    ex = None if expression is None else np.ascontiguousarray(expression, np.float64)
    gp = None if gene_pos is None else np.ascontiguousarray(gene_pos, np.uint64)
    sp = None if snarl_pos is None else np.ascontiguousarray(snarl_pos, np.uint64)
    own = tsv_path is None
    if own:
        fd, tsv_path = tempfile.mkstemp(suffix=".tsv", prefix="stoat_ref_")
        os.close(fd)
    nf = C.c_uint64(0)
    secs = L.ref_run(t, _p(bits, C.c_uint8), int(n_rows), int(n_cols), _p(keep[0], C.c_uint64), S, _p(spo, C.c_uint64),
                     _p(keep[1], C.c_uint64), _p(keep[2], C.c_uint64), _p(keep[3], C.c_uint8), _p(sp, C.c_uint64),
                     _p(ic, C.c_uint8), _p(yy, C.c_double), _p(cv, C.c_double), 0 if cv is None else cv.shape[1],
                     _p(ex, C.c_double), 0 if ex is None else ex.shape[0], _p(gp, C.c_uint64), int(window), int(min_ind),
                     int(min_hap), float(maf), int(n_threads), tsv_path.encode(), C.byref(nf))
    if secs < 0:
        raise RuntimeError("oracle/_ref: ref_run failed")
    out = dict(seconds=secs, n_filtered=int(nf.value), tsv_path=None if own else tsv_path)
    if parse:
        rows = {}
        with open(tsv_path) as f:
            for r in csv.DictReader(f, delimiter="\t"):
                rows.setdefault(int(r["SNARL"].split("_")[0]), []).append(r)
        out["rows"] = rows
    if own:
        os.unlink(tsv_path)
    return out


def num_threads() -> int:
    return int(lib().ref_num_threads())
