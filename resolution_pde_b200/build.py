"""Builds librpde_b200.so in-tree with nvcc for sm_100a.  `python resolution_pde_b200/build.py [--force]`."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "librpde_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-shared", "-Xcompiler", "-fPIC",
         "-lcuda"]


def sources():
    return sorted(os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cu", ".cuh"))) + \
        [os.path.join(HERE, "..", "include", "rpde_b200.h")]


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    return all(os.path.getmtime(s) <= t for s in sources())


def build(force=False, verbose=False):
    if not force and up_to_date():
        return OUT
    defs = ["-DRPDE_WITH_TC"] if os.path.exists(os.path.join(SRC, "gemm_tc.cuh")) else []
    defs += os.environ.get("RPDE_NVCC_DEFS", "").split()
    cmd = [NVCC] + FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, os.path.join(SRC, "rpde.cu")]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
