"""ctypes binding of libstoat_gpu.so (include/stoat_gpu.h).  Fails loudly: no fallback of any kind."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libstoat_gpu.so")

OK, ERR_ARG, ERR_CUDA, ERR_STATE, ERR_OOM = 0, -1, -2, -3, -4
P_VALUE, P_NA, P_ZERO, P_ONE, P_DBLMAX, P_UNDEF, P_OVERLIMIT = range(7)
MISSING_EDGE = 0xFFFFFFFF

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
f64p = C.POINTER(C.c_double)


class Thresholds(C.Structure):
    _fields_ = [("min_individuals", C.c_uint64), ("min_haplotypes", C.c_uint64), ("maf_threshold", C.c_double)]


class PathsC(C.Structure):
    _fields_ = [("n_snarls", C.c_uint64), ("snarl_path_off", u64p), ("path_edge_off", u64p), ("edge_row", u32p)]


class BinaryOut(C.Structure):
    _fields_ = [("g0", u32p), ("g1", u32p), ("total_sum", u64p), ("filtered", u8p), ("p_chi2", f64p),
                ("p_fisher", f64p), ("chi2_code", u8p), ("fisher_code", u8p)]


class QuantOut(C.Structure):
    _fields_ = [("allele_paths", u32p), ("filtered", u8p), ("n_used", u32p), ("p_value", f64p), ("r2", f64p),
                ("beta", f64p), ("se", f64p), ("code", u8p)]


class BinaryCovarOut(C.Structure):
    _fields_ = [("allele_paths", u32p), ("filtered", u8p), ("n_used", u32p), ("p_value", f64p), ("beta", f64p),
                ("se", f64p), ("code", u8p)]


class EqtlGenes(C.Structure):
    _fields_ = [("n_genes", C.c_uint64), ("expression", f64p), ("snarl_gene_off", u64p), ("gene_idx", u32p)]


class EqtlOut(C.Structure):
    _fields_ = [("allele_paths", u32p), ("filtered", u8p), ("n_used", u32p), ("p_value", f64p), ("r2", f64p),
                ("beta", f64p), ("se", f64p), ("code", u8p)]


class EqtlSummary(C.Structure):
    _fields_ = [("min_p", f64p), ("argmin_gene", u32p), ("n_below", u32p)]


class Timing(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("repack_ms", C.c_float), ("table_ms", C.c_float), ("stats_ms", C.c_float),
                ("d2h_ms", C.c_float), ("launches", C.c_uint32), ("n_general", C.c_uint32), ("n_fisher", C.c_uint32)]


# every symbol include/stoat_gpu.h declares: name -> (restype, argtypes)
CTX = C.c_void_p
SIGNATURES = {
    "stoat_gpu_abi_version": (C.c_int, []),
    "stoat_gpu_create": (C.c_int, [C.POINTER(CTX), C.c_int]),
    "stoat_gpu_destroy": (None, [CTX]),
    "stoat_gpu_last_error": (C.c_char_p, [CTX]),
    "stoat_gpu_set_matrix": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_uint64]),
    "stoat_gpu_set_matrix_dev": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_uint64]),
    "stoat_gpu_set_binary_phenotype": (C.c_int, [CTX, C.c_void_p, C.c_uint64]),
    "stoat_gpu_set_quantitative_phenotype": (C.c_int, [CTX, C.c_void_p, C.c_uint64]),
    "stoat_gpu_set_covariates": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_uint64]),
    "stoat_gpu_binary_batch": (C.c_int, [CTX, C.POINTER(PathsC), C.POINTER(Thresholds), C.POINTER(BinaryOut)]),
    "stoat_gpu_quantitative_batch": (C.c_int, [CTX, C.POINTER(PathsC), C.POINTER(Thresholds), C.POINTER(QuantOut)]),
    "stoat_gpu_gather_rows_dev": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "stoat_gpu_set_matrix_pitched_dev": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]),
    "stoat_gpu_debug_umma_probe": (C.c_int, [CTX, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]),
    "stoat_gpu_format_values": (C.c_int, [CTX, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "stoat_gpu_binary_covar_batch": (C.c_int, [CTX, C.POINTER(PathsC), C.POINTER(Thresholds), C.c_int,
                                               C.POINTER(BinaryCovarOut)]),
    "stoat_gpu_binary_covar_run_staged": (C.c_int, [CTX, C.POINTER(Thresholds), C.c_int]),
    "stoat_gpu_binary_covar_fetch": (C.c_int, [CTX, C.POINTER(BinaryCovarOut)]),
    "stoat_gpu_eqtl_batch": (C.c_int, [CTX, C.POINTER(PathsC), C.POINTER(EqtlGenes), C.POINTER(Thresholds),
                                       C.POINTER(EqtlOut)]),
    "stoat_gpu_eqtl_dense_summary": (C.c_int, [CTX, C.POINTER(PathsC), C.POINTER(EqtlGenes), C.POINTER(Thresholds),
                                               C.c_double, C.POINTER(EqtlSummary)]),
    "stoat_gpu_set_expression": (C.c_int, [CTX, C.c_void_p, C.c_uint64]),
    "stoat_gpu_set_expression_dev": (C.c_int, [CTX, C.c_void_p, C.c_uint64]),
    "stoat_gpu_eqtl_dense_run_staged": (C.c_int, [CTX, C.POINTER(Thresholds), C.c_double]),
    "stoat_gpu_eqtl_summary_fetch": (C.c_int, [CTX, C.POINTER(EqtlSummary)]),
    "stoat_gpu_stage_paths": (C.c_int, [CTX, C.POINTER(PathsC)]),
    "stoat_gpu_binary_run_staged": (C.c_int, [CTX, C.POINTER(Thresholds)]),
    "stoat_gpu_quantitative_run_staged": (C.c_int, [CTX, C.POINTER(Thresholds)]),
    "stoat_gpu_binary_counts_staged": (C.c_int, [CTX]),
    "stoat_gpu_binary_counts_dev": (C.c_int, [CTX, C.POINTER(C.c_void_p), u64p]),
    "stoat_gpu_binary_stats_staged": (C.c_int, [CTX, C.POINTER(Thresholds)]),
    "stoat_gpu_binary_fetch": (C.c_int, [CTX, C.POINTER(BinaryOut)]),
    "stoat_gpu_quantitative_fetch": (C.c_int, [CTX, C.POINTER(QuantOut)]),
    "stoat_gpu_set_quantitative_phenotype_shard": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_double, C.c_double,
                                                             C.c_double, C.c_double]),
    "stoat_gpu_quantitative_moments_staged": (C.c_int, [CTX, C.POINTER(Thresholds)]),
    "stoat_gpu_quantitative_partials_dev": (C.c_int, [CTX, C.POINTER(C.c_void_p), u64p, C.POINTER(C.c_void_p), u64p]),
    "stoat_gpu_quantitative_solve_staged": (C.c_int, [CTX, C.POINTER(Thresholds)]),
    "stoat_gpu_last_timing": (C.c_int, [CTX, C.POINTER(Timing)]),
    "stoat_gpu_mean_timing": (C.c_int, [CTX, C.c_uint32, C.POINTER(Timing)]),
    "stoat_gpu_stream": (C.c_uint64, [CTX]),
    "stoat_gpu_set_matrix_pairs": (C.c_int, [CTX, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]),
    "stoat_gpu_bh_adjust": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_void_p]),
    "stoat_gpu_bh_adjust_static": (C.c_int, [CTX, C.c_void_p, C.c_uint64, C.c_void_p]),
    "stoat_gpu_row_pitch": (C.c_uint64, [CTX]),
    "stoat_gpu_read_row": (C.c_int, [CTX, C.c_uint64, C.c_void_p, C.c_uint64]),
    "stoat_gpu_format_value": (C.c_int, [C.c_double, C.c_uint8, C.c_char_p, C.c_size_t]),
}

_lib = None


def load() -> C.CDLL:
    """Load the CUDA library.  Raises if it has not been built — there is no Python / CPU stand-in."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m stoat_b200.build` (nvcc, sm_100a). "
            "stoat_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
