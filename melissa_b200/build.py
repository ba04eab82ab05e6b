"""Build the CUDA shared library in-tree (nvcc, sm_100a only)."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
LIB_PATH = os.path.join(PKG_DIR, "libmelissa_b200.so")
SOURCES = [os.path.join(PKG_DIR, "csrc", "melissa.cu")]
HEADERS = sorted(glob.glob(os.path.join(PKG_DIR, "csrc", "*.cuh")) + glob.glob(os.path.join(PKG_DIR, "csrc", "*.h")) +
                 glob.glob(os.path.join(ROOT, "include", "*.h")))


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA library cannot be built")


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(f) > t for f in SOURCES + HEADERS)


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile melissa_b200/csrc/*.cu into melissa_b200/libmelissa_b200.so."""
    if not force and not is_stale():
        return LIB_PATH
    cmd = [nvcc_path(), "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
           "-Xcompiler", "-fPIC", "-shared", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(PKG_DIR, "csrc"), "-o", LIB_PATH] + SOURCES
    cmd += os.environ.get("MELISSA_NVCC_FLAGS", "").split()      # e.g. -DSWEEP_XU=2 for tuning experiments
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    print(build_library(force=True, verbose="-v" in sys.argv))
