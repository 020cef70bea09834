"""Build libstoat_gpu.so (CUDA kernels + C ABI) in-tree for sm_100a with nvcc.

The shared object lands next to this file so that it travels to the GPU box with the repository
snapshot; it is git-ignored.  `python -m stoat_b200.build` or `__graft_entry__.build()`.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libstoat_gpu.so")
SOURCES = ["abi.cu", "kernels_matrix.cu", "kernels_binary.cu", "kernels_quant.cu", "kernels_eqtl.cu", "kernels_post.cu", "kernels_logit.cu", "kernels_eqtl_umma.cu"]
HEADERS = ["ctx.hpp", "ptx.cuh", "device_math.cuh", "ols.cuh", "stage_pipeline.cuh", "quant_general.cuh", "quant_lean.cuh", "quant_reg.cuh", "quant_mma.cuh", "dosage_masks.cuh", "eqtl_dense.hpp", "eqtl_pair.cuh", os.path.join("..", "..", "include", "stoat_gpu.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
] + os.environ.get("STOAT_NVCC_EXTRA", "").split()


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(HERE, "build", src.replace(".cu", ".o"))
        cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    link = [nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart", "-lpthread"]
    subprocess.check_call(link)
    return LIB


HOST = os.path.join(os.path.dirname(HERE), "host")
SHIM_TEST = os.path.join(HOST, "shim_selftest")


def build_host(force: bool = False) -> str:
    """The C++ host shim (host/stoat_gpu_shim.cpp) + its self-test driver, linked against libstoat_gpu.so."""
    srcs = [os.path.join(HOST, f) for f in ("shim_selftest.cpp", "stoat_gpu_shim.cpp")]
    deps = srcs + [os.path.join(HOST, "stoat_gpu_shim.hpp"), LIB]
    if not force and os.path.exists(SHIM_TEST) and all(os.path.getmtime(d) <= os.path.getmtime(SHIM_TEST) for d in deps):
        return SHIM_TEST
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra"] + srcs + ["-o", SHIM_TEST, "-L" + HERE, "-lstoat_gpu",
                                                                      "-Wl,-rpath," + HERE, "-Wl,-rpath,$ORIGIN/../stoat_b200", "-pthread"]
    subprocess.check_call(cmd)
    return SHIM_TEST


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
