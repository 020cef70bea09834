"""CPU oracle of the `stoat vcf` hot path — TEST INFRASTRUCTURE, not product code.

ctypes front end of ``oracle/libstoat_oracle.so`` (built from ``oracle/stoat_oracle.cpp`` by
``oracle/Makefile`` / ``__graft_entry__.build()``).  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this package; the
product (``stoat_b200``) never does.

Array conventions are those of ``include/stoat_gpu.h``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libstoat_oracle.so")

MISSING_EDGE = 0xFFFFFFFF
P_VALUE, P_NA, P_ZERO, P_ONE, P_DBLMAX, P_UNDEF = range(6)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "stoat_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _declare(_lib)
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _declare(L):
    u8p, u32p, u64p, f64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint32, C.c_uint64, C.c_double))
    L.orc_matrix_get.restype = C.c_int
    L.orc_matrix_get.argtypes = [u8p, C.c_uint64, C.c_uint64, C.c_uint64]
    L.orc_matrix_set.restype = None
    L.orc_matrix_set.argtypes = [u8p, C.c_uint64, C.c_uint64, C.c_uint64]
    L.orc_matrix_set_many.restype = None
    L.orc_matrix_set_many.argtypes = [u8p, C.c_uint64, u64p, u64p, C.c_uint64]
    L.orc_identify_path.restype = C.c_uint64
    L.orc_identify_path.argtypes = [u8p, C.c_uint64, u32p, C.c_uint64, u64p]
    L.orc_set_precision.restype = C.c_int
    L.orc_set_precision.argtypes = [C.c_double, C.c_char_p, C.c_uint64]
    L.orc_format_value.restype = C.c_int
    L.orc_format_value.argtypes = [C.c_double, C.c_uint8, C.c_char_p, C.c_uint64]
    L.orc_chi2_2x2.restype = None
    L.orc_chi2_2x2.argtypes = [C.c_uint64] * 4 + [f64p, u8p]
    L.orc_fisher_2x2.restype = None
    L.orc_fisher_2x2.argtypes = [C.c_uint64] * 4 + [f64p, u8p]
    L.orc_chi2_2xN.restype = None
    L.orc_chi2_2xN.argtypes = [u64p, u64p, C.c_uint64, f64p, u8p]
    L.orc_fisher_khi2.restype = None
    L.orc_fisher_khi2.argtypes = [u64p, u64p, C.c_uint64, f64p, u8p, f64p, u8p]
    L.orc_chi2_pvalue.restype = C.c_double
    L.orc_chi2_pvalue.argtypes = [C.c_double, C.c_uint64]
    for f in (L.orc_gamma_q, L.orc_gamma_p, L.orc_student_t_sf):
        f.restype = C.c_double
        f.argtypes = [C.c_double, C.c_double]
    L.orc_filtration_binary.restype = C.c_int
    L.orc_filtration_binary.argtypes = [u64p, u64p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_double]
    L.orc_filtration_quant.restype = C.c_int
    L.orc_filtration_quant.argtypes = [f64p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_double]
    L.orc_remove_empty_columns_quant.restype = C.c_uint64
    L.orc_remove_empty_columns_quant.argtypes = [f64p, C.c_uint64, C.c_uint64]
    L.orc_combine_identical_columns.restype = C.c_uint64
    L.orc_combine_identical_columns.argtypes = [f64p, C.c_uint64, C.c_uint64]
    L.orc_linear_regression.restype = None
    L.orc_linear_regression.argtypes = [f64p, C.c_uint64, C.c_uint64, f64p, f64p, C.c_uint64,
                                        f64p, f64p, f64p, f64p, u8p]
    L.orc_xtx_inverse.restype = C.c_int
    L.orc_xtx_inverse.argtypes = [f64p, C.c_int, f64p]
    L.orc_binary_batch.restype = C.c_double
    L.orc_binary_batch.argtypes = [u8p, C.c_uint64, u8p, C.c_uint64, u64p, u64p, u32p,
                                   C.c_uint64, C.c_uint64, C.c_double,
                                   u32p, u32p, u64p, u8p, f64p, f64p, u8p, u8p, C.c_int, C.c_char_p]
    L.orc_quantitative_batch.restype = C.c_double
    L.orc_quantitative_batch.argtypes = [u8p, C.c_uint64, f64p, f64p, C.c_uint64, C.c_uint64, u64p, u64p, u32p,
                                         C.c_uint64, C.c_uint64, C.c_double,
                                         u32p, u8p, u32p, f64p, f64p, f64p, f64p, u8p, C.c_int]
    L.orc_binary_covar_batch.restype = C.c_double
    L.orc_binary_covar_batch.argtypes = [u8p, C.c_uint64, u8p, f64p, C.c_uint64, C.c_int, C.c_uint64, u64p, u64p, u32p,
                                         C.c_uint64, C.c_uint64, C.c_double,
                                         u32p, u8p, u32p, f64p, f64p, f64p, u8p, C.c_int]
    L.orc_logistic_regression.restype = None
    L.orc_logistic_regression.argtypes = [f64p, C.c_uint64, C.c_uint64, u8p, f64p, C.c_uint64, C.c_int,
                                          f64p, f64p, f64p, u8p]
    L.orc_eqtl_batch.restype = C.c_double
    L.orc_eqtl_batch.argtypes = [u8p, C.c_uint64, f64p, C.c_uint64, u64p, u32p, f64p, C.c_uint64,
                                 C.c_uint64, u64p, u64p, u32p, C.c_uint64, C.c_uint64, C.c_double,
                                 u32p, u8p, u32p, f64p, f64p, f64p, f64p, u8p, C.c_int]
    L.orc_quantitative_table.restype = C.c_uint64
    L.orc_quantitative_table.argtypes = [u8p, C.c_uint64, u64p, u32p, C.c_uint64, C.c_uint64,
                                         C.c_uint64, C.c_uint64, C.c_double, u32p, u64p, f64p, u64p, u8p]
    L.orc_num_threads.restype = C.c_int
    L.orc_num_threads.argtypes = []


# ------------------------------------------------------------------------------------------------
# data holders
# ------------------------------------------------------------------------------------------------
@dataclass
class Paths:
    """Flattened snarl paths (``stoat_gpu_paths`` of include/stoat_gpu.h)."""
    snarl_path_off: np.ndarray  # uint64 [S+1]
    path_edge_off: np.ndarray   # uint64 [P+1]
    edge_row: np.ndarray        # uint32 [sum R]

    def __post_init__(self):
        self.snarl_path_off = np.ascontiguousarray(self.snarl_path_off, dtype=np.uint64)
        self.path_edge_off = np.ascontiguousarray(self.path_edge_off, dtype=np.uint64)
        self.edge_row = np.ascontiguousarray(self.edge_row, dtype=np.uint32)
        if self.edge_row.size == 0:  # keep a valid pointer
            self.edge_row = np.zeros(1, dtype=np.uint32)[:0].copy()

    @property
    def n_snarls(self) -> int:
        return len(self.snarl_path_off) - 1

    @property
    def n_paths(self) -> int:
        return int(self.snarl_path_off[-1])

    def slice(self, s0: int, s1: int) -> "Paths":
        """Snarls [s0, s1) as a self-contained batch (shard for one rank)."""
        p0, p1 = int(self.snarl_path_off[s0]), int(self.snarl_path_off[s1])
        e0, e1 = int(self.path_edge_off[p0]), int(self.path_edge_off[p1])
        return Paths(self.snarl_path_off[s0:s1 + 1] - np.uint64(p0),
                     self.path_edge_off[p0:p1 + 1] - np.uint64(e0),
                     self.edge_row[e0:e1].copy())


def matrix_bytes(n_rows: int, n_cols: int) -> int:
    return (n_rows * n_cols + 7) // 8


def pack_matrix(dense: np.ndarray) -> np.ndarray:
    """dense bool [rows, cols] -> reference layout bytes (src/matrix.cpp:67-84)."""
    flat = np.ascontiguousarray(dense, dtype=np.uint8).reshape(-1)
    return np.packbits(flat, bitorder="little")


def unpack_matrix(bits: np.ndarray, n_rows: int, n_cols: int) -> np.ndarray:
    return np.unpackbits(bits, bitorder="little")[: n_rows * n_cols].reshape(n_rows, n_cols).astype(bool)


# ------------------------------------------------------------------------------------------------
# scalar entry points
# ------------------------------------------------------------------------------------------------
def set_precision(v: float) -> str:
    buf = C.create_string_buffer(64)
    lib().orc_set_precision(float(v), buf, 64)
    return buf.value.decode()


def format_value(v: float, code: int) -> str:
    buf = C.create_string_buffer(64)
    lib().orc_format_value(float(v), int(code), buf, 64)
    return buf.value.decode()


def _vc():
    return C.c_double(), C.c_uint8()


def chi2_2x2(a, b, c, d):
    v, k = _vc()
    lib().orc_chi2_2x2(a, b, c, d, C.byref(v), C.byref(k))
    return v.value, k.value


def fisher_2x2(a, b, c, d):
    v, k = _vc()
    lib().orc_fisher_2x2(a, b, c, d, C.byref(v), C.byref(k))
    return v.value, k.value


def chi2_2xN(g0, g1):
    g0 = np.ascontiguousarray(g0, dtype=np.uint64)
    g1 = np.ascontiguousarray(g1, dtype=np.uint64)
    v, k = _vc()
    lib().orc_chi2_2xN(_p(g0, C.c_uint64), _p(g1, C.c_uint64), len(g0), C.byref(v), C.byref(k))
    return v.value, k.value


def fisher_khi2(g0, g1):
    """-> ((chi2 value, code), (fisher value, code)) like FisherKhi2::fisher_khi2."""
    g0 = np.ascontiguousarray(g0, dtype=np.uint64)
    g1 = np.ascontiguousarray(g1, dtype=np.uint64)
    cv, ck = _vc()
    fv, fk = _vc()
    lib().orc_fisher_khi2(_p(g0, C.c_uint64), _p(g1, C.c_uint64), len(g0),
                          C.byref(cv), C.byref(ck), C.byref(fv), C.byref(fk))
    return (cv.value, ck.value), (fv.value, fk.value)


def chi2_pvalue(stat: float, df: int) -> float:
    return lib().orc_chi2_pvalue(float(stat), int(df))


def gamma_q(a, x):
    return lib().orc_gamma_q(float(a), float(x))


def gamma_p(a, x):
    return lib().orc_gamma_p(float(a), float(x))


def student_t_sf(df, t):
    return lib().orc_student_t_sf(float(df), float(t))


def identify_path(bits, n_cols, rows):
    rows = np.ascontiguousarray(rows, dtype=np.uint32)
    out = np.zeros(max(int(n_cols), 1), dtype=np.uint64)
    rp = _p(rows, C.c_uint32) if rows.size else C.POINTER(C.c_uint32)()
    n = lib().orc_identify_path(_p(bits, C.c_uint8), n_cols, rp, rows.size, _p(out, C.c_uint64))
    return out[:n].copy()


def filtration_binary(g0, g1, total_sum, min_ind, min_hap, maf) -> bool:
    g0 = np.ascontiguousarray(g0, dtype=np.uint64)
    g1 = np.ascontiguousarray(g1, dtype=np.uint64)
    return bool(lib().orc_filtration_binary(_p(g0, C.c_uint64), _p(g1, C.c_uint64), len(g0),
                                            total_sum, min_ind, min_hap, maf))


def filtration_quant(df, min_ind, min_hap, maf) -> bool:
    df = np.ascontiguousarray(df, dtype=np.float64)
    if df.ndim != 2:
        df = df.reshape(len(df), -1)
    return bool(lib().orc_filtration_quant(_p(df, C.c_double), df.shape[0], df.shape[1], min_ind, min_hap, maf))


def remove_empty_columns_quant(df) -> np.ndarray:
    df = np.array(df, dtype=np.float64, order="C")
    nc = lib().orc_remove_empty_columns_quant(_p(df, C.c_double), df.shape[0], df.shape[1])
    return df.reshape(-1)[: df.shape[0] * nc].reshape(df.shape[0], nc).copy()


def combine_identical_columns(df) -> np.ndarray:
    df = np.array(df, dtype=np.float64, order="C")
    nc = lib().orc_combine_identical_columns(_p(df, C.c_double), df.shape[0], df.shape[1])
    return df.reshape(-1)[: df.shape[0] * nc].reshape(df.shape[0], nc).copy()


def linear_regression(df, y, covar=None):
    """-> dict(p, beta, se, r2, code), LinearRegression::linear_regression."""
    df = np.ascontiguousarray(df, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    n, nv = df.shape
    if covar is None or np.size(covar) == 0:
        cov_p, ncov = C.POINTER(C.c_double)(), 0
    else:
        covar = np.ascontiguousarray(covar, dtype=np.float64)
        cov_p, ncov = _p(covar, C.c_double), covar.shape[1]
    p, b, se, r2 = (C.c_double() for _ in range(4))
    code = C.c_uint8()
    lib().orc_linear_regression(_p(df, C.c_double), n, nv, _p(y, C.c_double), cov_p, ncov,
                                C.byref(p), C.byref(b), C.byref(se), C.byref(r2), C.byref(code))
    return dict(p=p.value, beta=b.value, se=se.value, r2=r2.value, code=code.value)


def logistic_regression(df, y, covar=None, use_covariates=False):
    """-> dict(p, beta, se, code), LogisticRegression::logistic_regression (src/stats_test.cpp:49-176).
    use_covariates=False is the reference as written (its covariate block shadows its own variables, :59-62)."""
    df = np.ascontiguousarray(df, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.uint8)
    n, nv = df.shape
    if covar is None or np.size(covar) == 0:
        cov_p, ncov = C.POINTER(C.c_double)(), 0
    else:
        covar = np.ascontiguousarray(covar, dtype=np.float64)
        cov_p, ncov = _p(covar, C.c_double), covar.shape[1]
    p, b, se = (np.zeros(1) for _ in range(3))
    code = np.zeros(1, np.uint8)
    lib().orc_logistic_regression(_p(df, C.c_double), n, nv, _p(y, C.c_uint8), cov_p, ncov, int(bool(use_covariates)),
                                  _p(p, C.c_double), _p(b, C.c_double), _p(se, C.c_double), _p(code, C.c_uint8))
    return dict(p=float(p[0]), beta=float(b[0]), se=float(se[0]), code=int(code[0]))


def xtx_inverse(xtx):
    a = np.asfortranarray(xtx, dtype=np.float64)
    out = np.zeros_like(a, order="F")
    used_pinv = lib().orc_xtx_inverse(_p(a, C.c_double), a.shape[0], _p(out, C.c_double))
    return np.array(out), bool(used_pinv)


# ------------------------------------------------------------------------------------------------
# batched analyzers
# ------------------------------------------------------------------------------------------------
def binary_batch(bits, n_cols, is_case, paths: Paths, min_ind=3, min_hap=5, maf=0.05,
                 n_threads=0, tsv_path=None):
    S, P = paths.n_snarls, paths.n_paths
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    is_case = np.ascontiguousarray(is_case, dtype=np.uint8)
    out = dict(g0=np.zeros(P, np.uint32), g1=np.zeros(P, np.uint32), total_sum=np.zeros(S, np.uint64),
               filtered=np.zeros(S, np.uint8), p_chi2=np.zeros(S), p_fisher=np.zeros(S),
               chi2_code=np.zeros(S, np.uint8), fisher_code=np.zeros(S, np.uint8))
    secs = lib().orc_binary_batch(
        _p(bits, C.c_uint8), n_cols, _p(is_case, C.c_uint8), S,
        _p(paths.snarl_path_off, C.c_uint64), _p(paths.path_edge_off, C.c_uint64), _p(paths.edge_row, C.c_uint32),
        min_ind, min_hap, maf,
        _p(out["g0"], C.c_uint32), _p(out["g1"], C.c_uint32), _p(out["total_sum"], C.c_uint64),
        _p(out["filtered"], C.c_uint8), _p(out["p_chi2"], C.c_double), _p(out["p_fisher"], C.c_double),
        _p(out["chi2_code"], C.c_uint8), _p(out["fisher_code"], C.c_uint8),
        n_threads, tsv_path.encode() if tsv_path else None)
    out["seconds"] = secs
    return out


def _quant_out(S, P, n_pairs=None):
    n = S if n_pairs is None else n_pairs
    return dict(allele_paths=np.zeros(P, np.uint32), filtered=np.zeros(S, np.uint8), n_used=np.zeros(S, np.uint32),
                p_value=np.zeros(n), r2=np.zeros(n), beta=np.zeros(n), se=np.zeros(n), code=np.zeros(n, np.uint8))


def quantitative_batch(bits, n_cols, y, paths: Paths, covar=None, min_ind=3, min_hap=5, maf=0.05, n_threads=0):
    S, P = paths.n_snarls, paths.n_paths
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    y = np.ascontiguousarray(y, dtype=np.float64)
    if covar is None or np.size(covar) == 0:
        cov_p, ncov = C.POINTER(C.c_double)(), 0
    else:
        covar = np.ascontiguousarray(covar, dtype=np.float64)
        cov_p, ncov = _p(covar, C.c_double), covar.shape[1]
    out = _quant_out(S, P)
    secs = lib().orc_quantitative_batch(
        _p(bits, C.c_uint8), n_cols, _p(y, C.c_double), cov_p, ncov, S,
        _p(paths.snarl_path_off, C.c_uint64), _p(paths.path_edge_off, C.c_uint64), _p(paths.edge_row, C.c_uint32),
        min_ind, min_hap, maf,
        _p(out["allele_paths"], C.c_uint32), _p(out["filtered"], C.c_uint8), _p(out["n_used"], C.c_uint32),
        _p(out["p_value"], C.c_double), _p(out["r2"], C.c_double), _p(out["beta"], C.c_double),
        _p(out["se"], C.c_double), _p(out["code"], C.c_uint8), n_threads)
    out["seconds"] = secs
    return out


def binary_covar_batch(bits, n_cols, is_case, paths: Paths, covar=None, use_covariates=False, min_ind=3, min_hap=5,
                       maf=0.05, n_threads=0):
    """BinaryCovarSnarlAnalyzer (src/snarl_analyzer.cpp:392-428): quantitative table + logistic regression per snarl."""
    S, P = paths.n_snarls, paths.n_paths
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    is_case = np.ascontiguousarray(is_case, dtype=np.uint8)
    if covar is None or np.size(covar) == 0:
        cov_p, ncov = C.POINTER(C.c_double)(), 0
    else:
        covar = np.ascontiguousarray(covar, dtype=np.float64)
        cov_p, ncov = _p(covar, C.c_double), covar.shape[1]
    out = dict(allele_paths=np.zeros(P, np.uint32), filtered=np.zeros(S, np.uint8), n_used=np.zeros(S, np.uint32),
               p_value=np.zeros(S), beta=np.zeros(S), se=np.zeros(S), code=np.zeros(S, np.uint8))
    secs = lib().orc_binary_covar_batch(
        _p(bits, C.c_uint8), n_cols, _p(is_case, C.c_uint8), cov_p, ncov, int(bool(use_covariates)), S,
        _p(paths.snarl_path_off, C.c_uint64), _p(paths.path_edge_off, C.c_uint64), _p(paths.edge_row, C.c_uint32),
        min_ind, min_hap, maf,
        _p(out["allele_paths"], C.c_uint32), _p(out["filtered"], C.c_uint8), _p(out["n_used"], C.c_uint32),
        _p(out["p_value"], C.c_double), _p(out["beta"], C.c_double), _p(out["se"], C.c_double),
        _p(out["code"], C.c_uint8), n_threads)
    out["seconds"] = secs
    return out


def eqtl_batch(bits, n_cols, expression, paths: Paths, snarl_gene_off=None, gene_idx=None, covar=None,
               min_ind=3, min_hap=5, maf=0.05, n_threads=0):
    S, P = paths.n_snarls, paths.n_paths
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    expression = np.ascontiguousarray(expression, dtype=np.float64)
    G = expression.shape[0]
    if snarl_gene_off is None:
        off_p, idx_p, n_pairs = C.POINTER(C.c_uint64)(), C.POINTER(C.c_uint32)(), S * G
    else:
        snarl_gene_off = np.ascontiguousarray(snarl_gene_off, dtype=np.uint64)
        gene_idx = np.ascontiguousarray(gene_idx, dtype=np.uint32)
        off_p, idx_p, n_pairs = _p(snarl_gene_off, C.c_uint64), _p(gene_idx, C.c_uint32), int(snarl_gene_off[-1])
    if covar is None or np.size(covar) == 0:
        cov_p, ncov = C.POINTER(C.c_double)(), 0
    else:
        covar = np.ascontiguousarray(covar, dtype=np.float64)
        cov_p, ncov = _p(covar, C.c_double), covar.shape[1]
    out = _quant_out(S, P, n_pairs)
    secs = lib().orc_eqtl_batch(
        _p(bits, C.c_uint8), n_cols, _p(expression, C.c_double), G, off_p, idx_p, cov_p, ncov, S,
        _p(paths.snarl_path_off, C.c_uint64), _p(paths.path_edge_off, C.c_uint64), _p(paths.edge_row, C.c_uint32),
        min_ind, min_hap, maf,
        _p(out["allele_paths"], C.c_uint32), _p(out["filtered"], C.c_uint8), _p(out["n_used"], C.c_uint32),
        _p(out["p_value"], C.c_double), _p(out["r2"], C.c_double), _p(out["beta"], C.c_double),
        _p(out["se"], C.c_double), _p(out["code"], C.c_uint8), n_threads)
    out["seconds"] = secs
    return out


def quantitative_table(bits, n_cols, paths: Paths, snarl: int, min_ind=3, min_hap=5, maf=0.05):
    """Regression table of one snarl after all column surgery -> (df, used, allele_paths, filtered)."""
    bits = np.ascontiguousarray(bits, dtype=np.uint8)
    p0, p1 = int(paths.snarl_path_off[snarl]), int(paths.snarl_path_off[snarl + 1])
    n = n_cols // 2
    ap = np.zeros(max(p1 - p0, 1), np.uint32)
    used = np.zeros(max(n, 1), np.uint64)
    df = np.zeros(max(n * max(p1 - p0, 1), 1), np.float64)
    nc = C.c_uint64()
    filt = C.c_uint8()
    nu = lib().orc_quantitative_table(
        _p(bits, C.c_uint8), n_cols, _p(paths.path_edge_off, C.c_uint64), _p(paths.edge_row, C.c_uint32),
        p0, p1 - p0, min_ind, min_hap, maf, _p(ap, C.c_uint32), _p(used, C.c_uint64), _p(df, C.c_double),
        C.byref(nc), C.byref(filt))
    return (df[: nu * nc.value].reshape(nu, nc.value).copy(), used[:nu].copy(), ap[: p1 - p0].copy(), bool(filt.value))


def num_threads() -> int:
    return lib().orc_num_threads()


def bh_adjust(p) -> np.ndarray:
    """adjust_pvalues_with_BH (src/post_processing.cpp:7-38), restated in numpy: sort by raw p (:12-14), adjusted[i] =
    p * n / (i + 1) in that order of operations (:19-22), running minimum from the largest p down (:26-28), clamp to 1
    (:31-33), back to the original order (:36-38).  Ties get equal values whatever order the sort leaves them in."""
    a = np.asarray(p, dtype=np.float64)
    n = a.size
    if n == 0:
        return a.copy()
    order = np.argsort(a, kind="stable")
    adj = (a[order] * np.float64(n)) / np.arange(1, n + 1, dtype=np.float64)
    adj = np.minimum.accumulate(adj[::-1])[::-1]
    out = np.empty(n, dtype=np.float64)
    out[order] = np.minimum(1.0, adj)
    return out
