"""Build ``sanm_b200/lib/libsanm_b200.so`` (CUDA, sm_100a) in-tree with nvcc.

    python -m sanm_b200.build [-v]

nvcc cross-compiles without a GPU.  Objects go to ``build/`` (git-ignored); the
shared library stays in-tree so it travels with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "sanm_b200", "csrc")
OBJ = os.path.join(ROOT, "build", "cuda")
LIBDIR = os.path.join(ROOT, "sanm_b200", "lib")
LIB = os.path.join(LIBDIR, "libsanm_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a",
         "-lineinfo", "-Xcompiler", "-fPIC,-fopenmp,-Wall,-Wno-format-zero-length,-Wno-unknown-pragmas",
         # host code at the ISA level the oracle is built for (oracle/Makefile
         # ARCH; the reference itself builds with -march=native): GCC then
         # contracts a*b+c into FMA in the scalar decision code (poly.cpp,
         # pade.cpp) exactly as it does in the reference's build -- without it
         # the Bairstow root finder differs from the oracle's at 1e-9
         # (tests/test_pade_parity.py)
         "-Xcompiler", "-march=x86-64-v3",
         "-x", "cu", "-Xptxas", "-v"]


def _nccl_flags():
    """NCCL headers/libs bundled with torch (2.28.x); optional"""
    try:
        import nvidia.nccl as m
        base = list(m.__path__)[0]
        inc = os.path.join(base, "include")
        lib = os.path.join(base, "lib")
        if os.path.exists(os.path.join(inc, "nccl.h")):
            return inc, lib
    except Exception:
        pass
    return None, None


def sources():
    return sorted(f for f in os.listdir(SRC) if f.endswith((".cu", ".cpp")))


def build(verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    hdrs = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith(".h")]
    hdrs.append(os.path.join(ROOT, "include", "sanm_b200.h"))
    hdrs.append(os.path.abspath(__file__))
    newest_hdr = max(os.path.getmtime(h) for h in hdrs)
    nccl_inc, nccl_lib = _nccl_flags()
    extra = []
    if nccl_inc:
        extra = ["-DSANM_WITH_NCCL", "-I", nccl_inc]
    objs, procs = [], []
    for f in sources():
        src = os.path.join(SRC, f)
        obj = os.path.join(OBJ, os.path.splitext(f)[0] + ".o")
        log = obj + ".log"
        objs.append(obj)
        if (os.path.exists(obj) and os.path.getmtime(obj) >
                max(os.path.getmtime(src), newest_hdr)):
            continue
        cmd = [NVCC] + FLAGS + extra + ["-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd), flush=True)
        procs.append((f, log, subprocess.Popen(
            cmd, stdout=open(log, "w"), stderr=subprocess.STDOUT)))
    failed = []
    for f, log, p in procs:
        if p.wait() != 0:
            failed.append((f, open(log).read()))
    if failed:
        for f, out in failed:
            sys.stderr.write(f"==== nvcc failed on {f}\n{out}\n")
        raise RuntimeError("CUDA build failed: " + ", ".join(f for f, _ in failed))
    if procs or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-Xcompiler", "-fopenmp",
                                                     "-lgomp"]
        if nccl_lib:
            # resolved at run time against the NCCL that torch has loaded
            cmd += ["-L", nccl_lib, "-l:libnccl.so.2",
                    "-Xlinker", "-rpath," + nccl_lib]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
