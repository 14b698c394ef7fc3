#!/usr/bin/env python
"""Run the reference application twice -- `oracle/_ref/ref_fea` (reference CPU
path) and `cxx/_build/gpu_fea` (the same fea/*.cpp over libsanm_b200.so) -- on
one config and diff what they write: the stats JSON fields of
fea/main.cpp:271-295,425-432,576-581 and every .obj vertex.

    python tools/fea_diff.py [--lib-dir DIR] sys.json bob.json
    (--lib-dir: directory holding the libsanm_b200.so to use, e.g. a symlink
     to the host-emulation build for CPU-only checks)
"""
import argparse
import glob
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = os.path.join(ROOT, "oracle", "_ref", "config")
REF_FEA = os.path.join(ROOT, "oracle", "_ref", "ref_fea")
GPU_FEA = os.path.join(ROOT, "cxx", "_build", "gpu_fea")


def read_obj_vertices(path):
    out = []
    with open(path) as f:
        for line in f:
            if line.startswith("v "):
                out.append([float(t) for t in line.split()[1:4]])
    return out


def run(binary, configs, workdir, env=None, timeout=3600):
    os.makedirs(workdir, exist_ok=True)
    paths = [c if os.path.isabs(c) else os.path.join(CFG, c) for c in configs]
    t0 = time.time()
    p = subprocess.run([binary] + paths, cwd=workdir, env=env, timeout=timeout,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True)
    return p.returncode, p.stdout, time.time() - t0


def solve_lines(stdout):
    """the `solving ...` progress lines with the per-iteration numbers"""
    return [l for l in stdout.splitlines() if l.startswith("solving ")]


def iters_of(stdout):
    return [int(m) for m in re.findall(r"iter=(\d+)", stdout)]


def force_rms_of(stdout):
    return [float(m) for m in re.findall(r"force rms[:=]\s*([0-9.eE+-]+)", stdout)]


def g_quantum(v):
    """one unit of the 6th significant digit of v (what printf %g resolves)"""
    import math
    return 10.0 ** (math.floor(math.log10(v)) - 5) if v > 0 else 1e-300


def flat(v):
    """leaves of a nested JSON value, in order"""
    if isinstance(v, (list, tuple)):
        for e in v:
            yield from flat(e)
    elif isinstance(v, dict):
        for k in sorted(v):
            yield from flat(v[k])
    else:
        yield v


def compare(ref_dir, gpu_dir, vtx_rtol=1e-7):
    """returns (report dict, ok)"""
    rep = {"json": {}, "obj": {}}
    ok = True
    for rj in sorted(glob.glob(os.path.join(ref_dir, "*.json"))):
        name = os.path.basename(rj)
        gj = os.path.join(gpu_dir, name)
        if not os.path.exists(gj):
            rep["json"][name] = "missing on gpu side"
            ok = False
            continue
        a, b = json.load(open(rj)), json.load(open(gj))
        d = {}
        for k in a:
            if k.startswith("time") or k in ("threads", "solver_threads"):
                continue
            d[k] = (a[k], b.get(k))
            if isinstance(a[k], int) or isinstance(a[k], str):
                ok &= a[k] == b.get(k)
        rep["json"][name] = d
    for ro in sorted(glob.glob(os.path.join(ref_dir, "*.obj"))):
        name = os.path.basename(ro)
        go = os.path.join(gpu_dir, name)
        if not os.path.exists(go):
            rep["obj"][name] = "missing on gpu side"
            ok = False
            continue
        va, vb = read_obj_vertices(ro), read_obj_vertices(go)
        if len(va) != len(vb):
            rep["obj"][name] = f"vertex count {len(va)} vs {len(vb)}"
            ok = False
            continue
        scale = max((abs(c) for v in va for c in v), default=1.0) or 1.0
        err = max((abs(x - y) for p, q in zip(va, vb) for x, y in zip(p, q)),
                  default=0.0)
        rep["obj"][name] = err / scale
        # the reference prints vertices with %g: differences in units of the
        # last printed digit (6 significant digits) of the larger operand
        rep.setdefault("obj_quanta", {})[name] = max(
            (abs(x - y) / g_quantum(max(abs(x), abs(y)))
             for p, q in zip(va, vb) for x, y in zip(p, q) if x != y),
            default=0.0)
        ok &= err / scale <= vtx_rtol
    return rep, ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib-dir")
    ap.add_argument("--keep", action="store_true")
    ap.add_argument("configs", nargs="+")
    a = ap.parse_args()
    env = dict(os.environ)
    if a.lib_dir:
        env["LD_LIBRARY_PATH"] = a.lib_dir + ":" + env.get("LD_LIBRARY_PATH", "")
    tmp = tempfile.mkdtemp(prefix="fea_diff_")
    rc_r, out_r, t_r = run(REF_FEA, a.configs, os.path.join(tmp, "ref"))
    rc_g, out_g, t_g = run(GPU_FEA, a.configs, os.path.join(tmp, "gpu"), env)
    print(f"ref_fea rc={rc_r} {t_r:.1f}s   gpu_fea rc={rc_g} {t_g:.1f}s   [{tmp}]")
    for l in solve_lines(out_r):
        print("  ref:", l)
    for l in solve_lines(out_g):
        print("  gpu:", l)
    if rc_r or rc_g:
        print(out_r[-2000:] if rc_r else out_g[-2000:])
        sys.exit(1)
    rep, ok = compare(os.path.join(tmp, "ref"), os.path.join(tmp, "gpu"))
    ok &= iters_of(out_r) == iters_of(out_g)
    print(json.dumps(rep, indent=1))
    print("iters ref", iters_of(out_r), "gpu", iters_of(out_g))
    print("MATCH" if ok else "MISMATCH")
    sys.exit(0 if ok else 2)


if __name__ == "__main__":
    main()
