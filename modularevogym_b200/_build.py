"""In-tree builds of the native code (no JIT cache: the built .so files travel with the repo).

* ``libmevo_b200.so``  -- CUDA engine + C ABI (``include/mevo_b200.h``), sm_100a only.
* ``simulator_cpp*.so`` -- pybind11 module with the reference's class names (``Sim``, ``Viewer``,
  ``Camera``) plus ``BatchSim``; links against ``libmevo_b200.so`` via ``$ORIGIN`` rpath.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig
from typing import List

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libmevo_b200.so")
EXT_SUFFIX = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
PYMOD_PATH = os.path.join(PKG_DIR, "simulator_cpp" + EXT_SUFFIX)

# --fmad=false: every product and sum is rounded separately, like the reference's x86-64 build
# (FP64 parity contract).  -lineinfo: ncu source pages map to these files.
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "--fmad=false", "-std=c++17",
    "-Xcompiler", "-fPIC", "-diag-suppress", "177",
]

LIB_SOURCES = ["engine.cu", "task.cu", "render.cu", "c_api.cpp", "topology.cpp"]
LIB_HEADERS = ["engine.h", "kernels.cuh", "wkernel.cuh", "gkernel.cuh", "device_layout.h", "topology.h", os.path.join("..", "..", "include", "mevo_b200.h")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA engine cannot be built (there is no CPU fallback)")


def _stale(target: str, deps: List[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build_lib(force: bool = False, verbose: bool = False) -> str:
    deps = [os.path.join(CSRC, s) for s in LIB_SOURCES + LIB_HEADERS]
    if force or _stale(LIB_PATH, deps):
        extra = os.environ.get("MEVO_NVCC_EXTRA", "").split()          # development only (e.g. -DMEVO_PROF)
        # --exclude-libs,ALL: symbols of static archives (libstdc++.a where the toolchain links it statically, cudart_static) stay
        # private to this library instead of mixing with another libstdc++ loaded in the process (numpy, torch)
        cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-shared", "-Xlinker", "--exclude-libs,ALL", "-o", LIB_PATH] + \
              [os.path.join(CSRC, s) for s in LIB_SOURCES]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return LIB_PATH


def build_pymodule(force: bool = False, verbose: bool = False) -> str:
    import pybind11

    build_lib(force=False, verbose=verbose)
    src = os.path.join(CSRC, "simulator_cpp.cpp")
    deps = [src, os.path.join(PKG_DIR, "..", "include", "mevo_b200.h"), LIB_PATH]
    if force or _stale(PYMOD_PATH, deps):
        cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
        cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
               "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"],
               "-I", os.path.join(PKG_DIR, "..", "include"),
               src, "-o", PYMOD_PATH, "-L", PKG_DIR, "-lmevo_b200", "-Wl,-rpath,$ORIGIN", "-Wl,--exclude-libs,ALL"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return PYMOD_PATH


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_lib(force, verbose)
    if os.path.exists(os.path.join(CSRC, "simulator_cpp.cpp")):
        build_pymodule(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
