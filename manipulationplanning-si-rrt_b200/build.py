"""In-tree build of libsirrt_b200.so (nvcc, sm_100a only) and of the C++ host programs.

The built library stays next to this file (git-ignored, but it travels to the GPU box with the
gpurun snapshot).  There is no JIT cache and no fallback: if the library is missing the package
import fails loudly.
"""
from __future__ import annotations

import os
import shutil
import subprocess

_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_DIR, "csrc")
HOST = os.path.join(_DIR, "host")
LIB = os.path.join(_DIR, "libsirrt_b200.so")
INCLUDE = os.path.join(os.path.dirname(_DIR), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-pthread", "-shared", "--cudart=static",
]


def _nvcc() -> str:
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: libsirrt_b200 needs the CUDA toolkit to build")


def _newest(paths) -> float:
    return max(os.path.getmtime(p) for p in paths)


def lib_sources():
    out = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".hpp", ".inl"))]
    out += [os.path.join(INCLUDE, f) for f in sorted(os.listdir(INCLUDE))]
    return out


def build_lib(force: bool = False, verbose: bool = False) -> str:
    srcs = lib_sources()
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _newest(srcs):
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB, os.path.join(CSRC, "sirrt_api.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd, cwd=CSRC)
    return LIB


def build_host(force: bool = False):
    """C++ host programs above the C ABI (the CollisionManager shim smoke test, the SI-RRT planner)."""
    built = []
    if not os.path.isdir(HOST):
        return built
    build_lib()
    for f in sorted(os.listdir(HOST)):
        if not f.endswith("_main.cpp"):
            continue
        src = os.path.join(HOST, f)
        exe = os.path.join(HOST, f[: -len("_main.cpp")])
        deps = [src] + [os.path.join(HOST, h) for h in os.listdir(HOST) if h.endswith((".hpp", ".h"))]
        if force or not os.path.exists(exe) or os.path.getmtime(exe) < _newest(deps + [LIB]):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", "-I", INCLUDE, "-I", HOST, "-o", exe, src,
                                   "-L", _DIR, "-lsirrt_b200", f"-Wl,-rpath,{_DIR}", "-Wl,-rpath,$ORIGIN/.."])
        built.append(exe)
    return built


if __name__ == "__main__":
    import sys

    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
