"""Build the HOST-EMULATION library used by the CPU-only tests.

Compiles the product sources under ``sanm_b200/csrc`` with ``-DSANM_HOSTEMU``
(kernels become plain loops, see ``sanm_b200/csrc/rt.h``) into
``tests/_hostemu/libsanm_b200_hostemu.so``.  Test infrastructure only.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "sanm_b200", "csrc")
OUT = os.path.join(ROOT, "tests", "_hostemu")


def sources():
    return sorted(f for f in os.listdir(SRC) if f.endswith((".cu", ".cpp")))


def build(verbose=False) -> str:
    os.makedirs(OUT, exist_ok=True)
    lib = os.path.join(OUT, "libsanm_b200_hostemu.so")
    hdrs = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith(".h")]
    hdrs.append(os.path.join(ROOT, "include", "sanm_b200.h"))
    newest_hdr = max(os.path.getmtime(h) for h in hdrs)
    objs = []
    procs = []
    for f in sources():
        src = os.path.join(SRC, f)
        obj = os.path.join(OUT, os.path.splitext(f)[0] + ".o")
        objs.append(obj)
        if (os.path.exists(obj) and os.path.getmtime(obj) >
                max(os.path.getmtime(src), newest_hdr)):
            continue
        cmd = ["g++", "-std=c++17", "-O2", "-g", "-fPIC", "-fopenmp",
               "-DSANM_HOSTEMU", "-march=x86-64-v3", "-x", "c++", "-c", src, "-o", obj, "-Wall",
               "-Wno-format-zero-length", "-Wno-unknown-pragmas",
               "-Wno-maybe-uninitialized"]
        if verbose:
            print(" ".join(cmd))
        procs.append((f, subprocess.Popen(cmd)))
    for f, p in procs:
        if p.wait() != 0:
            raise RuntimeError(f"host-emulation compile of {f} failed")
    if procs or not os.path.exists(lib):
        subprocess.check_call(["g++", "-shared", "-fopenmp", "-o", lib] + objs)
    return lib


if __name__ == "__main__":
    print(build(verbose=True))
