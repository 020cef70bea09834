"""NumPy front end of the C ABI (include/stoat_gpu.h): one `Context` per GPU.

Every call goes through libstoat_gpu.so; nothing here computes statistics.  Arrays follow the ABI:
the presence matrix is the reference's `EdgeBySampleMatrix` byte layout (src/matrix.cpp:67-84), snarl
paths are the CSR of include/stoat_gpu.h.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L


class StoatGpuError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"stoat_gpu error {code}: {msg}")
        self.code = code


@dataclass
class SnarlPaths:
    """Flattened snarl paths of one batch (`stoat_gpu_paths`)."""
    snarl_path_off: np.ndarray  # uint64 [S+1]
    path_edge_off: np.ndarray   # uint64 [P+1]
    edge_row: np.ndarray        # uint32 [sum R]

    def __post_init__(self):
        self.snarl_path_off = np.ascontiguousarray(self.snarl_path_off, dtype=np.uint64)
        self.path_edge_off = np.ascontiguousarray(self.path_edge_off, dtype=np.uint64)
        self.edge_row = np.ascontiguousarray(self.edge_row, dtype=np.uint32)

    @property
    def n_snarls(self) -> int:
        return len(self.snarl_path_off) - 1

    @property
    def n_paths(self) -> int:
        return int(self.snarl_path_off[-1]) if len(self.snarl_path_off) else 0

    def as_c(self) -> L.PathsC:
        er = self.edge_row if self.edge_row.size else np.zeros(1, np.uint32)
        self._keep = er
        return L.PathsC(self.n_snarls, self.snarl_path_off.ctypes.data_as(L.u64p),
                        self.path_edge_off.ctypes.data_as(L.u64p), er.ctypes.data_as(L.u32p))


def _ptr(a: np.ndarray):
    return C.c_void_p(a.ctypes.data)


class Context:
    """`stoat_gpu_ctx`: device state for one chromosome chunk on one GPU."""

    def __init__(self, device: int = 0):
        self._lib = L.load()
        self._ctx = L.CTX()
        rc = self._lib.stoat_gpu_create(C.byref(self._ctx), device)
        if rc != L.OK:
            msg = self._lib.stoat_gpu_last_error(None).decode()
            self._ctx = None
            raise StoatGpuError(rc, msg)
        self.device = device
        self.n_rows = self.n_cols = 0

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.stoat_gpu_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != L.OK:
            raise StoatGpuError(rc, self._lib.stoat_gpu_last_error(self._ctx).decode())

    # ---- state --------------------------------------------------------------------------------------
    def set_matrix(self, bits: np.ndarray, n_rows: int, n_cols: int):
        bits = np.ascontiguousarray(bits, dtype=np.uint8)
        need = (n_rows * n_cols + 7) // 8
        if bits.size < need:
            raise ValueError(f"matrix needs {need} bytes, got {bits.size}")
        self._check(self._lib.stoat_gpu_set_matrix(self._ctx, _ptr(bits), n_rows, n_cols))
        self.n_rows, self.n_cols = n_rows, n_cols

    def set_matrix_pairs(self, rows, cols, n_rows: int, n_cols: int):
        """The matrix from its set bits (row, column) — what make_edge_matrix emits per allele edge and carrier."""
        r = np.ascontiguousarray(rows, dtype=np.uint32)
        c = np.ascontiguousarray(cols, dtype=np.uint32)
        if r.size != c.size:
            raise ValueError("rows and cols differ in length")
        self._check(self._lib.stoat_gpu_set_matrix_pairs(self._ctx, _ptr(r), _ptr(c), r.size, n_rows, n_cols))
        self.n_rows, self.n_cols = n_rows, n_cols

    def set_matrix_ptr(self, host_ptr: int, n_rows: int, n_cols: int):
        """Same, from a raw host pointer (e.g. a pinned torch tensor's data_ptr())."""
        self._check(self._lib.stoat_gpu_set_matrix(self._ctx, C.c_void_p(host_ptr), n_rows, n_cols))
        self.n_rows, self.n_cols = n_rows, n_cols

    def set_matrix_dev(self, dev_ptr: int, n_rows: int, n_cols: int):
        self._check(self._lib.stoat_gpu_set_matrix_dev(self._ctx, C.c_void_p(dev_ptr), n_rows, n_cols))
        self.n_rows, self.n_cols = n_rows, n_cols

    def set_matrix_pitched_dev(self, dev_ptr: int, n_rows: int, n_cols: int, src_pitch: int):
        """rows already in the device layout (src_pitch bytes each, zero beyond n_cols): no repack"""
        self._check(self._lib.stoat_gpu_set_matrix_pitched_dev(self._ctx, C.c_void_p(dev_ptr), n_rows, n_cols, src_pitch))

    def gather_rows_dev(self, rows) -> tuple[int, int]:
        """(device pointer, pitch) of a context-owned buffer holding rows[i] of the local matrix, in that order"""
        r = np.ascontiguousarray(rows, dtype=np.uint32)
        ptr, pitch = C.c_void_p(), C.c_uint64()
        self._check(self._lib.stoat_gpu_gather_rows_dev(self._ctx, _ptr(r), r.size, C.byref(ptr), C.byref(pitch)))
        return int(ptr.value or 0), int(pitch.value)

    def set_binary_phenotype(self, is_case: np.ndarray):
        a = np.ascontiguousarray(is_case, dtype=np.uint8)
        self._check(self._lib.stoat_gpu_set_binary_phenotype(self._ctx, _ptr(a), a.size))

    def set_quantitative_phenotype(self, y: np.ndarray):
        a = np.ascontiguousarray(y, dtype=np.float64)
        self._check(self._lib.stoat_gpu_set_quantitative_phenotype(self._ctx, _ptr(a), a.size))

    def set_covariates(self, covar):
        if covar is None or np.size(covar) == 0:
            self._check(self._lib.stoat_gpu_set_covariates(self._ctx, None, 0, 0))
            return
        a = np.ascontiguousarray(covar, dtype=np.float64)
        self._check(self._lib.stoat_gpu_set_covariates(self._ctx, _ptr(a), a.shape[0], a.shape[1]))

    def bh_adjust(self, p) -> np.ndarray:
        """Benjamini-Hochberg adjusted p-values, adjust_pvalues_with_BH (src/post_processing.cpp:7-38)."""
        a = np.ascontiguousarray(p, dtype=np.float64)
        out = np.empty_like(a)
        self._check(self._lib.stoat_gpu_bh_adjust(self._ctx, _ptr(a), a.size, _ptr(out)))
        return out

    BH_STATIC_MAX_N = 1 << 20  # STOAT_GPU_BH_STATIC_MAX_N of include/stoat_gpu.h

    def bh_adjust_static(self, p) -> np.ndarray:
        """As bh_adjust through stoat_gpu_bh_adjust_static: scratch in a static device array of the library, no allocation
        on the device; at most BH_STATIC_MAX_N p-values.  Calls that overlap on the GPU race on the shared scratch."""
        a = np.ascontiguousarray(p, dtype=np.float64)
        out = np.empty_like(a)
        self._check(self._lib.stoat_gpu_bh_adjust_static(self._ctx, _ptr(a), a.size, _ptr(out)))
        return out

    def bh_adjust_static_dev(self, p_dev: int, n: int, out_dev: int) -> None:
        """The same with device pointers of the caller (nothing crosses PCIe but the validity flag)."""
        self._check(self._lib.stoat_gpu_bh_adjust_static(self._ctx, C.c_void_p(int(p_dev)), int(n), C.c_void_p(int(out_dev))))

    def debug_umma_probe(self, a, b) -> np.ndarray:
        """int32 C [128][rows of b] = a [128][k] @ b[rows][k].T through the tcgen05 kernel of the dense eQTL scan (tests)"""
        a = np.ascontiguousarray(a, dtype=np.int8)
        b = np.ascontiguousarray(b, dtype=np.int8)
        c = np.zeros((128, b.shape[0]), np.int32)
        self._check(self._lib.stoat_gpu_debug_umma_probe(self._ctx, _ptr(a), _ptr(b), b.shape[0], a.shape[1], _ptr(c)))
        return c

    def format_values(self, values, codes=None) -> list[str]:
        """The value column as the reference prints it (set_precision + the literal strings of the codes), formatted on the
        device (include/stoat_gpu.h: stoat_gpu_format_values)."""
        v = np.ascontiguousarray(values, dtype=np.float64)
        c = None if codes is None else np.ascontiguousarray(codes, dtype=np.uint8)
        slots = np.zeros((v.size, 16), np.uint8)
        lens = np.zeros(v.size, np.uint8)
        self._check(self._lib.stoat_gpu_format_values(self._ctx, _ptr(v), None if c is None else _ptr(c), v.size, _ptr(slots),
                                                      _ptr(lens)))
        raw = slots.tobytes()
        return [raw[16 * i:16 * i + int(lens[i])].decode() for i in range(v.size)]

    @property
    def row_pitch(self) -> int:
        return int(self._lib.stoat_gpu_row_pitch(self._ctx))

    @property
    def stream(self) -> int:
        return int(self._lib.stoat_gpu_stream(self._ctx))

    def read_row(self, row: int) -> np.ndarray:
        out = np.zeros(self.row_pitch, np.uint8)
        self._check(self._lib.stoat_gpu_read_row(self._ctx, row, _ptr(out), out.size))
        return out

    def timing(self) -> dict:
        t = L.Timing()
        self._check(self._lib.stoat_gpu_last_timing(self._ctx, C.byref(t)))
        return {k: getattr(t, k) for k, _ in L.Timing._fields_}

    def mean_timing(self, n_runs: int) -> dict:
        """table_ms / stats_ms averaged over the last `n_runs` staged runs (the runs themselves are asynchronous)."""
        t = L.Timing()
        self._check(self._lib.stoat_gpu_mean_timing(self._ctx, int(n_runs), C.byref(t)))
        return {k: getattr(t, k) for k, _ in L.Timing._fields_}

    # ---- output holders -----------------------------------------------------------------------------
    @staticmethod
    def _binary_arrays(S: int, P: int) -> dict:
        return dict(g0=np.zeros(P, np.uint32), g1=np.zeros(P, np.uint32), total_sum=np.zeros(S, np.uint64),
                    filtered=np.zeros(S, np.uint8), p_chi2=np.zeros(S), p_fisher=np.zeros(S),
                    chi2_code=np.zeros(S, np.uint8), fisher_code=np.zeros(S, np.uint8))

    @staticmethod
    def _binary_c(o: dict) -> L.BinaryOut:
        return L.BinaryOut(o["g0"].ctypes.data_as(L.u32p), o["g1"].ctypes.data_as(L.u32p),
                           o["total_sum"].ctypes.data_as(L.u64p), o["filtered"].ctypes.data_as(L.u8p),
                           o["p_chi2"].ctypes.data_as(L.f64p), o["p_fisher"].ctypes.data_as(L.f64p),
                           o["chi2_code"].ctypes.data_as(L.u8p), o["fisher_code"].ctypes.data_as(L.u8p))

    @staticmethod
    def _quant_arrays(S: int, P: int) -> dict:
        return dict(allele_paths=np.zeros(P, np.uint32), filtered=np.zeros(S, np.uint8),
                    n_used=np.zeros(S, np.uint32), p_value=np.zeros(S), r2=np.zeros(S), beta=np.zeros(S),
                    se=np.zeros(S), code=np.zeros(S, np.uint8))

    @staticmethod
    def _quant_c(o: dict) -> L.QuantOut:
        return L.QuantOut(o["allele_paths"].ctypes.data_as(L.u32p), o["filtered"].ctypes.data_as(L.u8p),
                          o["n_used"].ctypes.data_as(L.u32p), o["p_value"].ctypes.data_as(L.f64p),
                          o["r2"].ctypes.data_as(L.f64p), o["beta"].ctypes.data_as(L.f64p),
                          o["se"].ctypes.data_as(L.f64p), o["code"].ctypes.data_as(L.u8p))

    @staticmethod
    def _thr(min_ind, min_hap, maf) -> L.Thresholds:
        return L.Thresholds(int(min_ind), int(min_hap), float(maf))

    # ---- the drop-in calls: host buffers in, host buffers out -----------------------------------------
    def binary_batch(self, paths: SnarlPaths, min_ind=3, min_hap=5, maf=0.05, out: dict | None = None) -> dict:
        o = out if out is not None else self._binary_arrays(paths.n_snarls, paths.n_paths)
        pc, thr, oc = paths.as_c(), self._thr(min_ind, min_hap, maf), self._binary_c(o)
        self._check(self._lib.stoat_gpu_binary_batch(self._ctx, C.byref(pc), C.byref(thr), C.byref(oc)))
        return o

    def quantitative_batch(self, paths: SnarlPaths, min_ind=3, min_hap=5, maf=0.05, out: dict | None = None) -> dict:
        o = out if out is not None else self._quant_arrays(paths.n_snarls, paths.n_paths)
        pc, thr, oc = paths.as_c(), self._thr(min_ind, min_hap, maf), self._quant_c(o)
        self._check(self._lib.stoat_gpu_quantitative_batch(self._ctx, C.byref(pc), C.byref(thr), C.byref(oc)))
        return o

    # ---- binary trait with covariates: logistic regression (BinaryCovarSnarlAnalyzer, src/snarl_analyzer.cpp:392-428) ---
    @staticmethod
    def _covar_arrays(S: int, P: int) -> dict:
        return dict(allele_paths=np.zeros(P, np.uint32), filtered=np.zeros(S, np.uint8),
                    n_used=np.zeros(S, np.uint32), p_value=np.zeros(S), beta=np.zeros(S), se=np.zeros(S),
                    code=np.zeros(S, np.uint8))

    @staticmethod
    def _covar_c(o: dict) -> L.BinaryCovarOut:
        return L.BinaryCovarOut(o["allele_paths"].ctypes.data_as(L.u32p), o["filtered"].ctypes.data_as(L.u8p),
                                o["n_used"].ctypes.data_as(L.u32p), o["p_value"].ctypes.data_as(L.f64p),
                                o["beta"].ctypes.data_as(L.f64p), o["se"].ctypes.data_as(L.f64p),
                                o["code"].ctypes.data_as(L.u8p))

    def binary_covar_batch(self, paths: SnarlPaths, use_covariates=False, min_ind=3, min_hap=5, maf=0.05,
                           out: dict | None = None) -> dict:
        """use_covariates=False is the reference as written (its covariates never reach the design, stats_test.cpp:59-62)."""
        o = out if out is not None else self._covar_arrays(paths.n_snarls, paths.n_paths)
        pc, thr, oc = paths.as_c(), self._thr(min_ind, min_hap, maf), self._covar_c(o)
        self._check(self._lib.stoat_gpu_binary_covar_batch(self._ctx, C.byref(pc), C.byref(thr), int(bool(use_covariates)),
                                                           C.byref(oc)))
        return o

    def binary_covar_run_staged(self, use_covariates=False, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_binary_covar_run_staged(self._ctx, C.byref(thr), int(bool(use_covariates))))

    def binary_covar_fetch(self, out: dict | None = None) -> dict:
        S, P = self._staged
        o = out if out is not None else self._covar_arrays(S, P)
        oc = self._covar_c(o)
        self._check(self._lib.stoat_gpu_binary_covar_fetch(self._ctx, C.byref(oc)))
        return o

    # ---- eQTL: one OLS per (snarl, gene of its window) ----------------------------------------------
    def _genes_c(self, expression, snarl_gene_off, gene_idx):
        expr = np.ascontiguousarray(expression, dtype=np.float64)
        keep = [expr]
        if snarl_gene_off is None:
            off_p, idx_p = None, None
        else:
            off = np.ascontiguousarray(snarl_gene_off, dtype=np.uint64)
            idx = np.ascontiguousarray(gene_idx, dtype=np.uint32)
            if idx.size == 0:
                idx = np.zeros(1, np.uint32)
            keep += [off, idx]
            off_p, idx_p = off.ctypes.data_as(L.u64p), idx.ctypes.data_as(L.u32p)
        g = L.EqtlGenes(expr.shape[0], expr.ctypes.data_as(L.f64p), off_p, idx_p)
        g._keep = keep
        return g

    def eqtl_batch(self, paths: SnarlPaths, expression, snarl_gene_off=None, gene_idx=None, min_ind=3, min_hap=5,
                   maf=0.05) -> dict:
        """expression [n_genes, n_samples]; gene windows as CSR (None = every gene for every snarl).  Per-pair
        outputs in CSR order (dense: pair = snarl * n_genes + gene)."""
        S, P = paths.n_snarls, paths.n_paths
        G = np.shape(expression)[0]
        n_pairs = S * G if snarl_gene_off is None else int(np.asarray(snarl_gene_off)[-1])
        o = dict(allele_paths=np.zeros(P, np.uint32), filtered=np.zeros(S, np.uint8), n_used=np.zeros(S, np.uint32),
                 p_value=np.zeros(n_pairs), r2=np.zeros(n_pairs), beta=np.zeros(n_pairs), se=np.zeros(n_pairs),
                 code=np.zeros(n_pairs, np.uint8))
        oc = L.EqtlOut(o["allele_paths"].ctypes.data_as(L.u32p), o["filtered"].ctypes.data_as(L.u8p),
                       o["n_used"].ctypes.data_as(L.u32p), o["p_value"].ctypes.data_as(L.f64p),
                       o["r2"].ctypes.data_as(L.f64p), o["beta"].ctypes.data_as(L.f64p),
                       o["se"].ctypes.data_as(L.f64p), o["code"].ctypes.data_as(L.u8p))
        pc, thr, gc = paths.as_c(), self._thr(min_ind, min_hap, maf), self._genes_c(expression, snarl_gene_off, gene_idx)
        self._check(self._lib.stoat_gpu_eqtl_batch(self._ctx, C.byref(pc), C.byref(gc), C.byref(thr), C.byref(oc)))
        return o

    def eqtl_dense_summary(self, paths: SnarlPaths, expression, p_threshold=1e-5, min_ind=3, min_hap=5,
                           maf=0.05) -> dict:
        """Dense (snarl x gene) scan reduced on the device: per snarl min p, its gene, and #genes with p < threshold."""
        S = paths.n_snarls
        o = dict(min_p=np.zeros(S), argmin_gene=np.zeros(S, np.uint32), n_below=np.zeros(S, np.uint32))
        oc = L.EqtlSummary(o["min_p"].ctypes.data_as(L.f64p), o["argmin_gene"].ctypes.data_as(L.u32p),
                           o["n_below"].ctypes.data_as(L.u32p))
        pc, thr, gc = paths.as_c(), self._thr(min_ind, min_hap, maf), self._genes_c(expression, None, None)
        self._check(self._lib.stoat_gpu_eqtl_dense_summary(self._ctx, C.byref(pc), C.byref(gc), C.byref(thr),
                                                           C.c_double(p_threshold), C.byref(oc)))
        return o

    # ---- resident mode ------------------------------------------------------------------------------
    def set_expression(self, expression):
        """Upload + centre the expression matrix [n_genes, n_samples] for dense scans (resident)."""
        a = np.ascontiguousarray(expression, dtype=np.float64)
        self._check(self._lib.stoat_gpu_set_expression(self._ctx, _ptr(a), a.shape[0]))

    def set_expression_dev(self, dev_ptr: int, n_genes: int):
        self._check(self._lib.stoat_gpu_set_expression_dev(self._ctx, C.c_void_p(dev_ptr), n_genes))

    def eqtl_dense_run_staged(self, p_threshold=1e-5, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_eqtl_dense_run_staged(self._ctx, C.byref(thr), C.c_double(p_threshold)))

    def eqtl_summary_fetch(self) -> dict:
        S, _ = self._staged
        o = dict(min_p=np.zeros(S), argmin_gene=np.zeros(S, np.uint32), n_below=np.zeros(S, np.uint32))
        oc = L.EqtlSummary(o["min_p"].ctypes.data_as(L.f64p), o["argmin_gene"].ctypes.data_as(L.u32p),
                           o["n_below"].ctypes.data_as(L.u32p))
        self._check(self._lib.stoat_gpu_eqtl_summary_fetch(self._ctx, C.byref(oc)))
        return o

    def stage_paths(self, paths: SnarlPaths):
        pc = paths.as_c()
        self._check(self._lib.stoat_gpu_stage_paths(self._ctx, C.byref(pc)))
        self._staged = (paths.n_snarls, paths.n_paths)

    def binary_run_staged(self, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_binary_run_staged(self._ctx, C.byref(thr)))

    def binary_counts_staged(self):
        self._check(self._lib.stoat_gpu_binary_counts_staged(self._ctx))

    def binary_counts_dev(self) -> tuple[int, int]:
        p, n = C.c_void_p(), C.c_uint64()
        self._check(self._lib.stoat_gpu_binary_counts_dev(self._ctx, C.byref(p), C.byref(n)))
        return int(p.value), int(n.value)

    def binary_stats_staged(self, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_binary_stats_staged(self._ctx, C.byref(thr)))

    def binary_fetch(self, out: dict | None = None) -> dict:
        S, P = self._staged
        o = out if out is not None else self._binary_arrays(S, P)
        oc = self._binary_c(o)
        self._check(self._lib.stoat_gpu_binary_fetch(self._ctx, C.byref(oc)))
        return o

    def quantitative_run_staged(self, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_quantitative_run_staged(self._ctx, C.byref(thr)))

    # ---- sample-sharded quantitative run (include/stoat_gpu.h): moments | all-reduce | solve ----------
    def set_quantitative_phenotype_shard(self, y_local, mean: float, sum_centred: float, sum_centred_sq: float,
                                         max_abs_centred: float):
        a = np.ascontiguousarray(y_local, dtype=np.float64)
        self._keep_y = a
        self._check(self._lib.stoat_gpu_set_quantitative_phenotype_shard(
            self._ctx, _ptr(a), a.size, float(mean), float(sum_centred), float(sum_centred_sq), float(max_abs_centred)))

    def quantitative_moments_staged(self, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_quantitative_moments_staged(self._ctx, C.byref(thr)))

    def quantitative_partials_dev(self) -> tuple[int, int, int, int]:
        """(device pointer, length) of the uint32 class counts and of the float64 sums to all-reduce in place"""
        pi, ni, pd, nd = C.c_void_p(), C.c_uint64(), C.c_void_p(), C.c_uint64()
        self._check(self._lib.stoat_gpu_quantitative_partials_dev(self._ctx, C.byref(pi), C.byref(ni), C.byref(pd),
                                                                 C.byref(nd)))
        return int(pi.value), int(ni.value), int(pd.value), int(nd.value)

    def quantitative_solve_staged(self, min_ind=3, min_hap=5, maf=0.05):
        thr = self._thr(min_ind, min_hap, maf)
        self._check(self._lib.stoat_gpu_quantitative_solve_staged(self._ctx, C.byref(thr)))

    def quantitative_fetch(self, out: dict | None = None) -> dict:
        S, P = self._staged
        o = out if out is not None else self._quant_arrays(S, P)
        oc = self._quant_c(o)
        self._check(self._lib.stoat_gpu_quantitative_fetch(self._ctx, C.byref(oc)))
        return o


def format_value(value: float, code: int) -> str:
    """The reference's output strings: set_precision (src/utils.cpp:5-15) + "NA"/"0"/"1"."""
    buf = C.create_string_buffer(64)
    n = L.load().stoat_gpu_format_value(float(value), int(code), buf, 64)
    if n < 0:
        raise ValueError("format_value: buffer too small")
    return buf.value.decode()
