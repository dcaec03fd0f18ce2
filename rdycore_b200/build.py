"""In-tree build of librdyb200.so (host C + CUDA sm_100a) and of the CPU oracles.

nvcc cross-compiles for sm_100a without a GPU.  The shared library lands next to
this file so that it travels to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "librdyb200.so")
OBJDIR = os.path.join(HERE, "build")

C_SOURCES = ["rdyb_error.c", "host_mesh.c", "device_layout.c", "petsc_operator_b200.c", "rdy_driver.c", "io_formats.c"]
CU_SOURCES = ["swe_device.cu", "swe_noalloc.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v", "-Wno-deprecated-gpu-targets",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA path cannot be built")


def _newer(src: list[str], dst: str) -> bool:
    if not os.path.exists(dst):
        return True
    t = os.path.getmtime(dst)
    return any(os.path.getmtime(s) > t for s in src)


def _run(cmd: list[str], log: list[str]) -> None:
    p = subprocess.run(cmd, capture_output=True, text=True)
    log.append("$ " + " ".join(cmd) + "\n" + p.stdout + p.stderr)
    if p.returncode != 0:
        sys.stderr.write(log[-1])
        raise RuntimeError(f"build step failed: {' '.join(cmd)}")


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/ into librdyb200.so; returns the path."""
    os.makedirs(OBJDIR, exist_ok=True)
    headers = [os.path.join(ROOT, "include", h) for h in os.listdir(os.path.join(ROOT, "include"))]
    headers += [os.path.join(CSRC, h) for h in os.listdir(CSRC) if h.endswith((".h", ".cuh"))]
    inc = ["-I" + os.path.join(ROOT, "include"), "-I" + CSRC]
    log: list[str] = []
    objs = []
    for s in C_SOURCES:
        src = os.path.join(CSRC, s)
        if not os.path.exists(src):
            continue
        obj = os.path.join(OBJDIR, s + ".o")
        if force or _newer([src] + headers, obj):
            _run(["gcc", "-std=gnu11", "-O2", "-fPIC", "-Wall", "-Wextra", *inc, "-c", src, "-o", obj], log)
        objs.append(obj)
    nvcc = _nvcc()
    for s in CU_SOURCES:
        src = os.path.join(CSRC, s)
        if not os.path.exists(src):
            continue
        obj = os.path.join(OBJDIR, s + ".o")
        if force or _newer([src] + headers, obj):
            _run([nvcc, *NVCC_FLAGS, *inc, "-c", src, "-o", obj], log)
        objs.append(obj)
    if force or _newer(objs, LIB):
        _run([nvcc, "-shared", "-Wno-deprecated-gpu-targets", "-o", LIB, *objs, "-lm", "-ldl"], log)
    if verbose:
        print("\n".join(log))
    with open(os.path.join(OBJDIR, "build.log"), "a") as f:
        f.write("\n".join(log))
    return LIB


def _all_sources() -> list[str]:
    out = [os.path.join(ROOT, "include", h) for h in os.listdir(os.path.join(ROOT, "include"))]
    out += [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    return out


def stale() -> bool:
    """True when a source file is newer than the built library."""
    return _newer(_all_sources(), LIB)


def build_oracles() -> None:
    """Build the CPU oracles (test infrastructure): the port always, oracle/_ref where the reference tree exists."""
    # `adapter`: rdycore_adapter.c (factories with the reference's signatures) against the reference's headers + librdyb200.so - the
    # object under test of tests/test_gpu_rdycore_adapter.py, built where the reference tree exists, after the library
    p = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port", "ref", "adapter", "bs2002"], capture_output=True, text=True)
    if p.returncode != 0:
        sys.stderr.write(p.stdout + p.stderr)
        raise RuntimeError("building the CPU oracles failed")


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
    build_oracles()
