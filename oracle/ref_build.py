#!/usr/bin/env python3
"""Build oracle/_ref/libstoat_ref.so: the REFERENCE'S OWN hot-path sources + oracle/ref_harness.cpp.  TEST INFRASTRUCTURE.

Recipe (no cmake, none of the reference's build system): g++ -O3 (the reference's OPTIMIZE build, CMakeLists.txt:31-35)
on the few source files the path needs, compiled where they lie under /root/reference/src — nothing is copied into this
repository — against
  * the reference's vendored Eigen (deps/eigen) and libhandlegraph headers (deps/libhandlegraph/src/include),
  * the stand-in headers of oracle/ref_stubs/ for the system packages this image lacks (Boost.Math/Multiprecision,
    htslib, libbdsg, libvgio; see oracle/ref_stubs/README.md).
Outputs go to oracle/_ref/ only (git-ignored, travels to the GPU box with the snapshot).

Runs only where /root/reference exists (this container).  `python oracle/ref_build.py [--force]`.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("STOAT_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT, "libstoat_ref.so")
# the path's translation units + what they link against inside the reference
REF_SOURCES = ["snarl_analyzer.cpp", "binary_table.cpp", "quantitative_table.cpp", "stats_test.cpp", "matrix.cpp",
               "utils.cpp", "writer.cpp", "snarl_data_t.cpp", "log.cpp", "arg_parser.cpp"]
HG_SOURCES = ["handle_graph.cpp", "path_handle_graph.cpp", "path_metadata.cpp", "path_position_handle_graph.cpp", "types.cpp"]
FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-w"]


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "src"))


def build(force: bool = False) -> str | None:
    if not available():
        return LIB if os.path.exists(LIB) else None
    os.makedirs(OUT, exist_ok=True)
    inc = ["-I" + os.path.join(HERE, "ref_stubs"), "-I" + os.path.join(REF, "src"), "-I" + os.path.join(REF, "deps", "eigen"),
           "-I" + os.path.join(REF, "deps", "libhandlegraph", "src", "include")]
    units = [(os.path.join(REF, "src", s), "ref_" + s) for s in REF_SOURCES] + \
            [(os.path.join(REF, "deps", "libhandlegraph", "src", s), "hg_" + s) for s in HG_SOURCES] + \
            [(os.path.join(HERE, "ref_harness.cpp"), "harness.cpp")]
    stubs = [os.path.join(dp, f) for dp, _, fs in os.walk(os.path.join(HERE, "ref_stubs")) for f in fs]
    stubs += [os.path.join(HERE, "special_functions.hpp"), os.path.abspath(__file__)]
    newest_dep = max(os.path.getmtime(p) for p in stubs)
    objs, procs = [], []
    for src, name in units:
        obj = os.path.join(OUT, name.replace(".cpp", ".o"))
        objs.append(obj)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(src), newest_dep):
            continue
        cmd = ["g++"] + FLAGS + inc + ["-c", src, "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode:
            sys.stderr.write(out)
            raise RuntimeError("oracle/_ref: compile failed: " + " ".join(cmd))
    if procs or not os.path.exists(LIB):
        subprocess.check_call(["g++", "-shared", "-fopenmp", "-Wl,--no-undefined", "-o", LIB] + objs)
    return LIB


# ---------------------------------------------------------------------------------------------------
# The C++ host shim compiled the way a maintainer would compile it INSIDE the reference tree:
# -DSTOAT_GPU_SHIM_USE_REFERENCE_TYPES, i.e. against stoat::Snarl_data_t / Edge_t / Path_traversal_t and
# stoat_vcf::EdgeBySampleMatrix themselves, linked with the reference's own matrix.cpp / snarl_data_t.cpp objects
# (libstoat_ref.so) and libstoat_gpu.so.  The only change to the reference is §2(a) of INTEGRATION.md — three accessors in
# class EdgeBySampleMatrix — applied here to a COPY of src/matrix.hpp written to oracle/_ref/patched/ (a build output,
# git-ignored; the shim finds it first on the include path, the reference's own sources keep including theirs).
# ---------------------------------------------------------------------------------------------------
ROOT = os.path.dirname(HERE)
SHIM_REFTYPES = os.path.join(OUT, "shim_reftypes_selftest")
ACCESSORS = """
    // --- INTEGRATION.md section 2(a): what the GPU shim reads -------------------------------------
    const std::vector<uint8_t>& data() const { return matrix_1D; }
    size_t cols() const { return cols_; }
    size_t rows() const { return row_header.size(); }
"""


def build_shim_reftypes(force: bool = False) -> str | None:
    gpu_lib = os.path.join(ROOT, "stoat_b200", "libstoat_gpu.so")
    if not available() or not os.path.exists(LIB) or not os.path.exists(gpu_lib):
        return SHIM_REFTYPES if os.path.exists(SHIM_REFTYPES) else None
    srcs = [os.path.join(ROOT, "host", f) for f in ("shim_selftest.cpp", "stoat_gpu_shim.cpp")]
    deps = srcs + [os.path.join(ROOT, "host", "stoat_gpu_shim.hpp"), LIB, gpu_lib, os.path.abspath(__file__)]
    if not force and os.path.exists(SHIM_REFTYPES) and all(os.path.getmtime(d) <= os.path.getmtime(SHIM_REFTYPES) for d in deps):
        return SHIM_REFTYPES
    patched = os.path.join(OUT, "patched")
    os.makedirs(patched, exist_ok=True)
    text = open(os.path.join(REF, "src", "matrix.hpp")).read()
    anchor = "protected:"
    if text.count(anchor) != 1:
        raise RuntimeError("oracle/_ref: src/matrix.hpp no longer has exactly one `protected:` — revisit INTEGRATION.md 2(a)")
    with open(os.path.join(patched, "matrix.hpp"), "w") as f:
        f.write(text.replace(anchor, ACCESSORS + "\n" + anchor))
    inc = ["-I" + patched, "-I" + os.path.join(HERE, "ref_stubs"), "-I" + os.path.join(REF, "src"),
           "-I" + os.path.join(REF, "deps", "eigen"), "-I" + os.path.join(REF, "deps", "libhandlegraph", "src", "include")]
    cmd = ["g++", "-std=c++17", "-O2", "-w", "-fopenmp", "-DSTOAT_GPU_SHIM_USE_REFERENCE_TYPES"] + inc + srcs + \
          ["-o", SHIM_REFTYPES, "-L" + OUT, "-lstoat_ref", "-L" + os.path.dirname(gpu_lib), "-lstoat_gpu", "-pthread",
           "-Wl,-rpath,$ORIGIN", "-Wl,-rpath,$ORIGIN/../../stoat_b200"]
    subprocess.check_call(cmd)
    return SHIM_REFTYPES


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(build_shim_reftypes(force="--force" in sys.argv))
