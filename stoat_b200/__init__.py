"""stoat_b200 — B200 (sm_100a) implementation of STOAT's per-snarl association-testing hot path.

The product is `libstoat_gpu.so` (hand-written CUDA kernels behind the C ABI of include/stoat_gpu.h);
this package is its thin NumPy front end plus synthetic-data helpers.  There is no CPU fallback:
importing works anywhere, but every computation requires the built library and a B200.
"""
from .api import Context, SnarlPaths, StoatGpuError, format_value  # noqa: F401

__all__ = ["Context", "SnarlPaths", "StoatGpuError", "format_value"]
