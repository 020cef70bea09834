"""The C-ABI shared library loads and exports every symbol include/stoat_gpu.h declares (no compute: CPU box)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "stoat_gpu.h")) as f:
        src = f.read()
    return sorted(set(re.findall(r"\b(stoat_gpu_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def built_lib():
    from stoat_b200 import build

    return build.build()


def test_every_declared_symbol_is_exported(built_lib):
    names = declared_symbols()
    assert len(names) >= 20
    lib = ctypes.CDLL(built_lib)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/stoat_gpu.h but not exported"
    from stoat_b200 import _lib

    assert sorted(_lib.SIGNATURES) == names  # the Python binding covers exactly the header
    assert _lib.load().stoat_gpu_abi_version() == 1


def test_format_value_matches_reference_strings(built_lib):
    # host-only helper of the ABI: set_precision (src/utils.cpp:5-15), tests/unittest/utils_unit.cpp:9-23
    from stoat_b200 import format_value

    assert format_value(0.00001234, 0) == "1.2340e-05"
    assert format_value(0.123456, 0) == "0.1235"
    assert format_value(0.333333333, 0) == "0.3333"
    assert format_value(1.0, 0) == "1"
    assert format_value(0.0, 0) == "0"
    assert format_value(0.0, 1) == "NA"
    assert format_value(5.0, 2) == "0"
    assert format_value(5.0, 3) == "1"
    assert format_value(0.0, 4) == "1.798e+308"


def test_no_cpu_fallback(built_lib):
    """Without a GPU the product must fail loudly instead of computing anything on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import stoat_b200

    with pytest.raises(stoat_b200.StoatGpuError) as e:
        stoat_b200.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under stoat_b200/ may import, link or load it."""
    pkg = os.path.join(ROOT, "stoat_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert "import oracle" not in src and "from oracle" not in src, fn
                assert "stoat_oracle" not in src and "orc_" not in src, fn
