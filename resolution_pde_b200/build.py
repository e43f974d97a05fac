"""Builds librpde_b200.so in-tree with nvcc for sm_100a.  `python resolution_pde_b200/build.py [--force] [-v]`.

Every `csrc/*.cu` is one translation unit, compiled in parallel into `csrc/_obj/` and linked into one shared library.
A SHA-256 over the sources, the public header and the flags is stored beside the library
(`librpde_b200.so.srchash`); `build()` rebuilds whenever the stored hash does not match the tree, so a prebuilt library
that travels with the repository (gpurun snapshot) is only reused when it really was built from these sources.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(SRC, "_obj")
OUT = os.path.join(HERE, "librpde_b200.so")
HASH = OUT + ".srchash"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CFLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xcompiler", "-fPIC"]
LFLAGS = ["-shared", "-Xcompiler", "-fPIC", "-gencode", "arch=compute_100a,code=sm_100a", "-lcuda"]


def sources():
    return sorted(os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cu", ".cuh"))) + \
        [os.path.join(HERE, "..", "include", "rpde_b200.h")]


def defines():
    defs = ["-DRPDE_WITH_TC"] if os.path.exists(os.path.join(SRC, "gemm_tc.cuh")) else []
    return defs + os.environ.get("RPDE_NVCC_DEFS", "").split()


def source_hash():
    h = hashlib.sha256()
    for s in sources():
        h.update(os.path.basename(s).encode())
        with open(s, "rb") as f:
            h.update(f.read())
    h.update(" ".join(CFLAGS + LFLAGS + defines()).encode())
    return h.hexdigest()


def up_to_date():
    if not (os.path.exists(OUT) and os.path.exists(HASH)):
        return False
    with open(HASH) as f:
        return f.read().strip() == source_hash()


def _deps(path, seen=None):
    """the file and every header it includes with #include "..." (recursively, relative to the including file)"""
    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith("#include \""):
                _deps(os.path.join(os.path.dirname(path), line.split('"')[1]), seen)
    return seen


def _unit_hash(unit):
    """hash of one translation unit: its own text + the headers it includes, so editing one kernel file recompiles
    only the units that use it"""
    h = hashlib.sha256()
    for s in sorted(_deps(os.path.join(SRC, unit))):
        with open(s, "rb") as f:
            h.update(f.read())
    h.update(" ".join(CFLAGS + defines()).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    if not force and up_to_date():
        return OUT
    os.makedirs(OBJ, exist_ok=True)
    units = sorted(f for f in os.listdir(SRC) if f.endswith(".cu"))
    extra = ["-Xptxas", "-v"] if verbose else []

    def compile_one(u):
        obj = os.path.join(OBJ, u[:-3] + ".o")
        stamp = obj + ".hash"
        uh = _unit_hash(u)
        if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read().strip() == uh:
            return obj
        subprocess.run([NVCC] + CFLAGS + defines() + extra + ["-c", "-o", obj, os.path.join(SRC, u)], check=True)
        with open(stamp, "w") as f:
            f.write(uh + "\n")
        return obj

    with ThreadPoolExecutor(max_workers=max(1, min(len(units), os.cpu_count() or 1))) as ex:
        objs = list(ex.map(compile_one, units))
    subprocess.run([NVCC] + LFLAGS + ["-o", OUT] + objs, check=True)
    with open(HASH, "w") as f:
        f.write(source_hash() + "\n")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
