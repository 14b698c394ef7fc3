#!/usr/bin/env python
"""Headline benchmark: element-Taylor throughput (tet*order steps/s) of the ANM
inner loop on a synthetic scaled tet cuboid, plus end-to-end solve times of the
reference's own `fea` configs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one full series expansion sweep of the element phase over the
whole tet batch: order-0 evaluation, per-tet Jacobian, and for every order
k = 1..20 gather (remap_in) -> Taylor propagation (bias phase + coefficient
phase) -> element-to-node assembly (remap_out); libsanm/anm.cpp:205-290 without
the global solves (SURVEY 8d metric (i)).  Metric: tets * order per second.

 value    inputs (nodal series x_0..x_K) resident in HBM, device time only
 e2e      same sweep through the C ABI with HOST buffers: H2D of the nodal
          series from pinned memory and D2H of the assembled right-hand sides
          inside the timed region
 roofline algorithmic bytes of the sweep (8*(S*(k+1)+24) B per
          tet*order, S = 20/22/27) / device time of the sweep's element
          kernels, against MEASURED_PEAKS.json hbm_gbs
 cpu_baseline  the unmodified reference (oracle/_ref, ParallelTaylorCoeffProp)
          on the host cores, bounded sample
 solves   end-to-end ANM solve times of the fea configs (GPU path) next to the
          reference CPU path on the same configs (N=1, rank 0 only)
"""
from __future__ import annotations

import os as _os
# all kernel images are loaded when the library is first used, not lazily at
# each kernel's first launch (must be set before the CUDA context exists)
_os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ORDER = 20
S_MIN = {"neohookean_i": 22, "neohookean_c": 20, "arap": 27}
ORACLE_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_capi.so")
CONFIG_DIR = os.path.join(ROOT, "oracle", "_ref", "config")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--energy", default="neohookean_c")
    # per-GPU cuboid (vertices per axis).  Defaults: N = 1 -> BASELINE config 5,
    # make_cuboid(401,101,51) = 10.0 M tets; N > 1 -> 201x51x41 = 2.0 M tets
    # per GPU (weak scaling: one cuboid N times longer in x)
    ap.add_argument("--nx", type=int, default=None)
    ap.add_argument("--ny", type=int, default=None)
    ap.add_argument("--nz", type=int, default=None)
    ap.add_argument("--exchange", default="halo", choices=["halo", "allreduce"],
                    help="N > 1: halo = nodal vectors stay distributed, only "
                         "interface rows are exchanged; allreduce = every rank "
                         "ends with the full right-hand sides")
    ap.add_argument("--no-solves", action="store_true",
                    help="skip the fea end-to-end solve section")
    ap.add_argument("--no-cpu", action="store_true",
                    help="skip the cpu_baseline leg")
    ap.add_argument("--sweep-x", action="store_true",
                    help="N = 1: also measure the SURVEY 8d size sweep "
                         "x in {51,101,201,401} (y, z as the workload)")
    ap.add_argument("--no-energies", action="store_true",
                    help="skip the NH-I / ARAP side measurements")
    ap.add_argument("--cpu-tets", type=int, default=1000000,
                    help="tets of the bounded CPU sample (an x-slab of the "
                         "workload's cuboid, same cross-section)")
    a = ap.parse_args()
    big = a.gpus == 1
    a.nx = a.nx or (401 if big else 201)
    a.ny = a.ny or (101 if big else 51)
    a.nz = a.nz or (51 if big else 41)
    return a


def algo_bytes_per_sweep(n_tets: int, energy: str, order: int = ORDER) -> float:
    S = S_MIN[energy]
    return float(n_tets) * sum(8.0 * (S * (k + 1) + 24)
                               for k in range(1, order + 1))


def make_series(x0: np.ndarray, order: int, seed: int = 92702102):
    """nodal series: x_0 = rest shape + small perturbation (keeps det F > 0),
    x_k ~ U(-1,1) * 0.01 * 0.5^k (as tests/pade.cpp:69-71 scales its series)"""
    rng = np.random.default_rng(seed)
    n = x0.size
    x = np.empty((order + 1, n))
    x[0] = x0 + rng.uniform(-1e-3, 1e-3, n)
    for k in range(1, order + 1):
        x[k] = rng.uniform(-1, 1, n) * (0.01 * 0.5 ** k)
    return x


def cuboid_model(binding, nx, ny, nz, energy, spacing=0.025,
                 cube_x_range=None):
    """the cuboid-twist body of fea/main.cpp:665-690 (both end faces fixed);
    cube_x_range: this rank's slab of hexahedra (all vertices are kept)"""
    from sanm_b200 import fea, material as mat
    mesh = fea.TetrahedralMesh.make_cuboid(nx, ny, nz, spacing,
                                           cube_x_range=cube_x_range)
    body = fea.DeformableBody(
        binding, mat.MaterialProperty.from_young_poisson(1e7, 0.45), mesh)
    vx = mesh.vertices[:, 0]
    body.coord_fixed_mask[(vx <= spacing / 2) |
                          (vx >= spacing * (nx - 1.5))] = True
    return mesh, body.make_forward(energy)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,"
         "clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self._stop_evt = threading.Event()
        self._init_nvml()

    def _init_nvml(self):
        """called from __init__: loading NVML takes ~100 ms, longer than a short
        timed region"""
        self._nv = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            self._h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self._mx = nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM)
            self._reasons = getattr(
                nv, "nvmlDeviceGetCurrentClocksEventReasons",
                None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            self._nv = nv
        except Exception:
            self._nv = None

    def _sample_nvml(self):
        bits = [(0x08, "hw_slowdown"), (0x40, "hw_thermal_slowdown"),
                (0x20, "sw_thermal_slowdown"), (0x04, "sw_power_cap")]
        try:
            mhz = self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM)
            r = int(self._reasons(self._h))
            self.samples.append(
                [str(mhz), str(self._mx), "0"] +
                ["Active" if r & bit else "Not Active" for bit, _ in bits])
        except Exception:
            pass

    def _run_nvml(self) -> bool:
        """NVML directly (5 ms period): nvidia-smi takes ~100 ms per call"""
        if self._nv is None:
            return False
        while not self._stop_evt.is_set():
            self._sample_nvml()
            self._stop_evt.wait(0.005)
        return True

    def run(self):
        if self._run_nvml():
            return
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", "-i", str(self.index),
                     "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                    capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([s.strip() for s in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self) -> dict:
        self._stop_evt.set()
        self.join(timeout=6)
        mhz, reasons, mx = [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        for s in self.samples:
            try:
                mhz.append(float(s[0]))
                mx = float(s[1])
            except Exception:
                continue
            for nm, v in zip(names, s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(mhz)) if mhz else None,
                "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(mhz)}


# ------------------------------------------------------------ reference arm
def load_oracle():
    from sanm_b200.capi import Binding
    if not os.path.exists(ORACLE_LIB):
        raise RuntimeError("oracle/_ref/libref_capi.so is missing "
                           "(make -C oracle)")
    return Binding(ctypes.CDLL(ORACLE_LIB), "ref_")


def cpu_sample_dims(target_tets: int, ny: int, nz: int, nx_max: int):
    """an x-slab of the workload's cuboid (same ny x nz cross-section, same
    spacing, material and boundary treatment) with about target_tets tets"""
    per_layer = 5 * (ny - 1) * (nz - 1)
    layers = min(max(2, round(target_tets / per_layer)), nx_max - 1)
    return layers + 1, ny, nz


class ReferenceSweep:
    """the reference's own element phase (ParallelTaylorCoeffProp +
    SparseLinearDesc::apply, oracle/ref_capi.cpp ref_sweep_run) on the host
    cores, on a small cuboid of the workload's family"""

    def __init__(self, ref, args, threads: int):
        ref.call("set_num_threads", ctypes.c_int(threads))
        self.ref = ref
        energy = args.energy
        self.dims = cpu_sample_dims(args.cpu_tets, args.ny, args.nz, args.nx)
        nx, ny, nz = self.dims
        self.mesh, self.model = cuboid_model(ref, nx, ny, nz, energy)
        self.x = make_series(self.model.lt_inp.x0, ORDER)
        self.b = np.empty((ORDER + 1, self.model.lt_inp.nr_unknown_vtx))
        self.n_tets = self.mesh.nr_tet

    def run(self) -> float:
        """one sweep; returns its wall seconds"""
        m = self.model
        dp = ctypes.POINTER(ctypes.c_double)
        sec = ctypes.c_double()
        self.ref.call("sweep_run", m.graph.h, ctypes.c_int(m.y.id),
                      m.lt_inp.desc.h, m.lt_out.h, ctypes.c_int(ORDER),
                      self.x.ctypes.data_as(dp), self.b.ctypes.data_as(dp),
                      ctypes.byref(sec))
        return sec.value


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    rs = ReferenceSweep(load_oracle(), args, threads)
    for _ in range(min(args.warmup, 1)):  # a CPU code needs one warm pass
        rs.run()
    times = [rs.run() for _ in range(args.steps)]
    ms = 1e3 * float(np.mean(times))
    ntet = rs.n_tets
    value = ntet * ORDER / (ms * 1e-3)
    sample = (f"each step = one sweep over an x-slab {rs.dims} = {ntet} tets of "
              f"the workload's cuboid (same cross-section, spacing, material, "
              f"order {ORDER}, {args.energy}); throughput is per tet, "
              f"the slab bounds the run time")
    full_tets = 5 * (args.nx - 1) * (args.ny - 1) * (args.nz - 1)
    line = {
        "impl": "reference", "metric": "tet_order_steps_per_s",
        "value": value, "unit": "tet*order/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, full_tets, args.gpus),
        "cpu_baseline": {"value": value, "unit": "tet*order/s",
                         "cores": threads, "kind": "reference",
                         "sample": sample},
        "e2e": {"value": value, "unit": "tet*order/s",
                "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": ("reference CPU path (unmodified libsanm sources, "
                 "ParallelTaylorCoeffProp with %d threads, Eigen kernels; MKL "
                 "unavailable on this box, not on the element path)" % threads),
    }
    emit_line(line)
    # (the reference's exit-time profiler report goes to fd 1, which main()
    # has pointed at stderr: the process can end normally)


def workload_config(args, tets_per_gpu, world=1):
    return {"workload": ("synthetic scaled tet cuboid (cuboid_twist family), "
                         "element phase of one order-%d series expansion: "
                         "gather + Taylor bias/coeff + Jacobian + assembly; "
                         "N GPUs: one cuboid %dx longer, tets sharded by "
                         "x-slab; %s" % (ORDER, world, (
                             "nodal vectors stay distributed, the interface "
                             "rows of the right-hand sides are summed between "
                             "neighbouring ranks every order (ncclSend/Recv)"
                             if args.exchange == "halo" else
                             "nodal right-hand sides summed by NCCL all-reduce "
                             "every order"))),
            "exchange": args.exchange if world > 1 else "none",
            "energy_model": args.energy, "order": ORDER,
            "cuboid_vertices_per_gpu": [args.nx, args.ny, args.nz],
            "cuboid_vertices_global": [(args.nx - 1) * world + 1, args.ny,
                                       args.nz],
            "tets_per_gpu": int(tets_per_gpu),
            "l2": "Taylor stacks per GPU exceed the 126 MB L2 (no flush needed)"}


# ------------------------------------------------------------------ our arm
def pinned_empty(shape):
    """numpy view of pinned host memory (torch owns the allocation)"""
    import torch
    t = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    return t, t.numpy()


def timed_task(binding, tpath, ov):
    from sanm_b200 import fea
    log = fea.RunLog(verbose=False)
    if binding.has("prof_reset"):
        binding.prof_reset()
    t0 = time.perf_counter()
    res = fea.run_task(binding, tpath, ov, save=False, log=log)
    wall = time.perf_counter() - t0
    st = res[1]
    if isinstance(st, list):  # cuboid twist: one dict per stage
        tsolve = sum(s["time"] for s in st)
        iters = [s["iter_tot"] for s in st]
        frms = st[-1]["force_rms_recomp"]
        es = [e for s in st for e in s.get("element_stats", [])]
    else:
        tsolve = st.get("time_solve", st.get("time"))
        iters = st.get("iter", st.get("iter_tot"))
        frms = st["force_rms_recomp"]
        es = st.get("element_stats", [])
        es = es if isinstance(es, list) else [es]
    return {"solve_s": tsolve, "wall_s": wall, "anm_iters": iters,
            "force_rms": frms}, es

def sparse_summary(es):
    """factor / back-solve figures of a run from the solver statistics"""
    es = [e for e in es if e and e.get("factor_entries")]
    if not es:
        return None
    # the profiler is process-wide: the last solver's record holds the totals
    e = es[-1]
    sp = {"n": e["n"], "nnz_K": e["nnz_a"],
          "factor_entries": e["factor_entries"],
          "factor_flops": e["factor_flops"],
          "levels": e["nr_levels"], "max_front": e["max_front"],
          "factor_ms_total": e["factor_ms"], "nr_factor": e["nr_factor"],
          "solve_ms_total": e["solve_ms"], "nr_solves": e["nr_solves"],
          "element_ms_total": e["element_ms"],
          # (counted per solver object: summed over the run's solvers)
          "pade_recovered": sum(x.get("pade_recovered") or 0 for x in es),
          "lu_perturbed_pivots": e.get("lu_perturbed_pivots"),
          "lu_refine_steps": e.get("lu_refine_steps")}
    if e["nr_solves"] and e["solve_ms"] > 0:
        # B_solve = 8 * nnz(L+U) + 16 n per back-solve
        bytes_solve = 8.0 * e["factor_entries"] + 16.0 * e["n"]
        sp["solve_gb_s"] = (bytes_solve * e["nr_solves"] /
                            (e["solve_ms"] * 1e-3) / 1e9)
    if e["nr_factor"] and e["factor_ms"] > 0:
        sp["factor_tflop_s"] = (e["factor_flops"] * e["nr_factor"] /
                                (e["factor_ms"] * 1e-3) / 1e12)
    return sp


BIG_SOLVE = ("cuboid61_twist4_neo_incomp",
             os.path.join(ROOT, "bench_configs", "cuboid61_twist_neo_incomp.json"),
             [os.path.join(ROOT, "bench_configs", "override_rotate4.json")])


def big_solve(b, world):
    """End-to-end solve at scale, every rank takes part (N2 of the round-1
    verdict; SURVEY 8c stand-in for BASELINE config 4, whose mesh is absent):
    the 60^3-cell cuboid = 1.08 M tets, 658,617 unknowns, neo-Hookean
    incompressible, a 4 degree twist (3 + 1 ANM iterations).  Tets are sharded
    over the ranks (element phase, K assembly), the factorisation and the
    back-solves are replicated (SURVEY 8e "replicas only" for the solver), so
    the Amdahl split is reported next to the time."""
    name, tpath, ov = BIG_SOLVE
    if not os.path.exists(tpath):
        return None
    row = {"config": name, "n_gpus": world,
           "workload": "make_cuboid(61,61,61): 1,080,000 tets, 658,617 "
                       "unknowns, neohookean_i, order 20, twist 4 degrees"}
    cold, _ = timed_task(b, tpath, ov)
    again, es = timed_task(b, tpath, ov)
    row["solve_s"] = cold["solve_s"]
    row["solve_s_cached_topology"] = again["solve_s"]
    row["anm_iters"] = again["anm_iters"]
    row["force_rms"] = again["force_rms"]
    sp = sparse_summary(es)
    if sp:
        row["sparse"] = sp
        tot = again["solve_s"] * 1e3
        row["amdahl_ms"] = {
            "total": tot, "factor (replicated)": sp["factor_ms_total"],
            "back-solves (replicated)": sp["solve_ms_total"],
            "element phase (sharded)": sp["element_ms_total"],
            "rest (K assembly, Pade, host glue, exchanges)":
                tot - sp["factor_ms_total"] - sp["solve_ms_total"] -
                sp["element_ms_total"]}
    return row


def fea_solves(b, ref, cpu_threads):
    """end-to-end solve times of the reference's fea configs: our GPU path and
    the reference CPU path through the same front end (sanm_b200/fea.py)"""
    from sanm_b200 import fea
    cases = [("bob", "bob.json", []),
             ("armadillo_small_bend", "armadillo_small.json",
              ["armadillo_small_bend_override.json"]),
             ("cuboid_twist_order20", "cuboid_twist.json",
              ["override_order20.json"]),
             # stand-in for BASELINE config 4 (the full Armadillo mesh is
             # absent from the reference checkout)
             ("armadillo_small_neo_incomp", "armadillo_small.json",
              ["override_neo_incomp.json"])]
    out = []
    # The in-process legs go through the Python front end (sanm_b200/fea.py, a
    # test harness that is known to take a slightly different continuation
    # path than the reference application on cuboid_twist).  Where a Pade jump
    # of that path fails the consistency check the reference behaviour is to
    # abort; these legs opt into the retry and REPORT how often it fired
    # (sparse.pade_recovered).  The reference application itself runs in the
    # gpu_fea / ref_fea columns, without any recovery.
    os.environ["SANM_PADE_RECOVER"] = "1"

    for name, task, ovs in cases:
        tpath = os.path.join(CONFIG_DIR, task)
        if not os.path.exists(tpath):
            continue
        ov = [os.path.join(CONFIG_DIR, o) for o in ovs]
        row = {"config": name}
        if b is not None:
            # first run: topology caches cold (ordering, symbolic analysis,
            # assembly lists are part of solve_s, as PARDISO's phase 11 is part
            # of the reference's); second run: the same topology again, as in
            # every later ANM iteration / load step of one simulation
            from sanm_b200.capi import Binding as _B
            _B.call_times.clear()
            row["gpu"], _ = timed_task(b, tpath, ov)
            if _B.call_times:  # SANM_TRACE_CALLS=1: where the cold run went
                sys.stderr.write("[trace] %s cold %.3f s\n%s\n" % (
                    name, row["gpu"]["solve_s"], b.prof_dump()))
                for k, (t, n) in sorted(_B.call_times.items(),
                                        key=lambda kv: -kv[1][0])[:8]:
                    sys.stderr.write("[trace]   %-34s %9.2f ms  n=%d\n" %
                                     (k, t * 1e3, n))
            again, es = timed_task(b, tpath, ov)
            row["gpu"]["solve_s_cached_topology"] = again["solve_s"]
            sp = sparse_summary(es)
            if sp:
                row["gpu"]["sparse"] = sp
        row["_paths"] = (tpath, ov)
        out.append(row)
    os.environ.pop("SANM_PADE_RECOVER", None)
    # second pass: the reference's CPU solves, after ALL of ours -- its worker
    # threads keep cores busy for a while after a run and would be timed as
    # part of the next configuration's cold (host-side) analysis
    for row in out:
        tpath, ov = row.pop("_paths")
        if ref is not None:
            ref.call("set_num_threads", ctypes.c_int(1))
            row["cpu_reference"], _ = timed_task(ref, tpath, ov)
        if "gpu" in row and "cpu_reference" in row:
            row["iters_match"] = (row["gpu"]["anm_iters"] ==
                                  row["cpu_reference"]["anm_iters"])
    return out


FEA_CASES = [("bob", ["bob.json"]),
             ("armadillo_small_bend", ["armadillo_small.json",
                                       "armadillo_small_bend_override.json"]),
             ("cuboid_twist_order20", ["cuboid_twist.json",
                                       "override_order20.json"]),
             ("armadillo_small_neo_incomp", ["armadillo_small.json",
                                             "override_neo_incomp.json"])]


def fea_binary_solves(with_cpu: bool, mt_threads: int):
    """The reference APPLICATION itself on both paths: cxx/_build/gpu_fea =
    the reference's fea/*.cpp compiled unmodified against the libsanm-compatible
    headers over libsanm_b200.so, oracle/_ref/ref_fea = the same sources on the
    reference's libsanm (CPU; Eigen::SparseLU in place of MKL PARDISO, which is
    not on this box).  Solve time = what the application itself reports:
    `time_solve` of the stats JSON (fea/main.cpp:425) or the sum of the
    `timing(sec): ... solve=` lines of a staged run (fea/main.cpp:544).  Every
    gpu_fea run is one process, so its topology caches start cold."""
    import re
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import fea_diff
    if not os.path.exists(fea_diff.GPU_FEA):
        return None

    def one(binary, cfgs, sys_json="sys.json"):
        d = tempfile.mkdtemp(prefix="bench_fea_")
        rc, out, wall = fea_diff.run(binary, [sys_json] + cfgs, d, timeout=1500)
        row = {"rc": rc, "wall_s": wall,
               "anm_iters": fea_diff.iters_of(out)}
        stage = [float(m) for m in re.findall(
            r"timing\(sec\): prep=[0-9.eE+-]+ solve=([0-9.eE+-]+)", out)]
        js = {}
        for f in sorted(os.listdir(d)):
            if f.endswith(".json"):
                js = json.load(open(os.path.join(d, f)))
        if "time_solve" in js:
            row["solve_s"] = js["time_solve"]
        elif stage:
            row["solve_s"] = sum(stage)
            row["stage_solve_s"] = stage
        for k in ("displacement", "force_rms_recomp", "nr_inverted"):
            if k in js:
                row[k] = js[k]
        if rc != 0:
            row["tail"] = out[-400:]
        import shutil
        shutil.rmtree(d, ignore_errors=True)
        return row

    # largest sys-mtN.json of the reference's config directory that the host
    # can run (N = min(32, cores))
    mt = max([n for n in (1, 2, 4, 8, 16, 32) if n <= mt_threads] or [1])
    rows = []
    for name, cfgs in FEA_CASES:
        row = {"config": name, "gpu_fea": one(fea_diff.GPU_FEA, cfgs)}
        rows.append(row)
    if with_cpu and os.path.exists(fea_diff.REF_FEA):
        for row, (name, cfgs) in zip(rows, FEA_CASES):
            row["ref_fea"] = one(fea_diff.REF_FEA, cfgs)
            row["ref_fea"]["threads"] = 1
            if mt > 1:
                # the same application at sys-mtN.json, the shared
                # constant pools pre-sized (oracle/shim/presize_pool.cpp)
                os.environ["SANM_REF_PRESIZE_ELEMS"] = "8000000"
                try:
                    row["ref_fea_mt"] = one(fea_diff.REF_FEA, cfgs,
                                            "sys-mt%d.json" % mt)
                finally:
                    del os.environ["SANM_REF_PRESIZE_ELEMS"]
                row["ref_fea_mt"]["threads"] = mt
            row["iters_match"] = (row["gpu_fea"]["anm_iters"] ==
                                  row["ref_fea"]["anm_iters"])
            g, r = row["gpu_fea"], row["ref_fea"]
            if "displacement" in g and "displacement" in r:
                row["displacement_rel_diff"] = abs(
                    g["displacement"] - r["displacement"]) / max(
                        abs(r["displacement"]), 1e-300)
    return rows


def fp64_peak_tflops(torch):
    """yardstick for the factorisation: cuBLAS DGEMM 8192^3 on this box (used
    only as the roofline denominator)"""
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    c = torch.empty_like(a)
    torch.matmul(a, a, out=c)
    best = 1e30
    for _ in range(4):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        torch.matmul(a, a, out=c)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, c
    torch.cuda.empty_cache()
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def side_energy(b, torch, stream, energy, peak, steps=5, warmup=3,
                dims=(201, 51, 41)):
    """the same sweep for another energy model on the 2.0 M-tet cuboid:
    device-resident value and the step kernel's roofline fraction"""
    from sanm_b200.capi import ElementSweep
    mesh, model = cuboid_model(b, *dims, energy)
    n_tets = mesh.nr_tet
    sweep = ElementSweep(model.y, model.lt_inp.desc, model.lt_out, ORDER)
    sweep.set_x(make_series(model.lt_inp.x0, ORDER))
    for _ in range(warmup):
        sweep.run()
    torch.cuda.synchronize()
    b.prof_reset()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(stream)
    for _ in range(steps):
        sweep.run()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    step_ms, step_launches = b.prof_get("k_fused_step")
    S = S_MIN[energy]
    step_algo = float(n_tets) * sum(8.0 * (S * (k + 1) + 24)
                                    for k in range(2, ORDER + 1))
    out = {"energy_model": energy, "tets": int(n_tets),
           "cuboid_vertices": list(dims),
           "value": n_tets * ORDER / (ms * 1e-3), "unit": "tet*order/s",
           "ms_per_step": ms, "steps": steps, "warmup": warmup}
    if step_launches:
        ach = step_algo * steps / (step_ms * 1e-3) / 1e9
        out["roofline"] = {"kernel": "fused.cu StepKernel", "bound": "hbm",
                           "achieved": ach, "peak": peak, "unit": "GB/s",
                           "frac": ach / peak,
                           "launch_ms_avg": step_ms / step_launches}
    del sweep
    return out


def bind_to_gpu_numa_node(torch, local_rank):
    """Run this rank on the cores of the NUMA node its GPU hangs off, BEFORE
    the pinned host buffers are allocated (first touch puts them there): with
    8 ranks on one host the copies of the e2e path otherwise cross sockets.
    Returns a short description for the JSON line."""
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return "numa node unknown"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"numa node {node}: no allowed cores"
        os.sched_setaffinity(0, cpus)
        return f"numa node {node}, {len(cpus)} cores"
    except Exception as e:  # plumbing only: never fail the run over it
        return f"not bound ({type(e).__name__})"


def run_ours(args):
    import torch
    import torch.distributed as dist
    import sanm_b200
    from sanm_b200.capi import ElementSweep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(
            "cuda", local_rank))
    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(torch, local_rank) if world > 1 else "single rank"
    b = sanm_b200.load()
    b.call("set_device", ctypes.c_int(local_rank))
    # a dedicated (non-default) stream shared by torch's events and the
    # library's kernels: handle 0 (the legacy default stream) means "library's
    # own stream" to sanm_set_stream, and events recorded there would not see
    # the kernels
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    b.call("set_stream", ctypes.c_void_p(stream.cuda_stream))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- workload (weak scaling): ONE cuboid, `world` times longer in x; each
    # rank owns the tets of one x-slab (a contiguous range of the element
    # list), all ranks hold the full nodal vectors and the assembled
    # right-hand side of every order is summed over the ranks (csrc/comm.h)
    if world > 1:
        from sanm_b200 import dist as sdist
        sdist.init_comm(b)
    nxg = (args.nx - 1) * world + 1
    mesh, model = cuboid_model(
        b, nxg, args.ny, args.nz, args.energy,
        cube_x_range=(rank * (args.nx - 1), (rank + 1) * (args.nx - 1)))
    n_tets = mesh.nr_tet
    n = model.lt_inp.nr_unknown_vtx
    sweep = ElementSweep(model.y, model.lt_inp.desc, model.lt_out, ORDER)
    x_t, x_host = pinned_empty((ORDER + 1, n))
    x_host[:] = make_series(model.lt_inp.x0, ORDER, seed=92702102)
    b_t, b_host = pinned_empty((ORDER + 1, n))
    sweep.set_x(x_host)
    st = sweep.stats()
    halo = world > 1 and args.exchange == "halo"
    if halo:
        sweep.set_exchange(1)
    xlo, xhi, blo, bhi = sweep.local_ranges() if halo else (0, n, 0, n)

    # ---- value: device-resident
    for _ in range(args.warmup):
        sweep.run()
    barrier()
    b.prof_reset()
    launches0 = b.launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(True), torch.cuda.Event(True)
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        sweep.run()
    ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    launches = b.launch_count() - launches0
    prof = {t: b.prof_get(t) for t in
            ("taylor_next_order", "push_xi", "remap_out", "k_order_bias",
             "k_order_coef", "k_order_eval", "k_jacobian", "k_fused_eval",
             "k_fused_step", "k_fused_jacobian", "k_assemble_final",
             "comm_allreduce", "comm_halo")}
    clocks = sampler.stop()

    # ---- e2e: host buffers through the C ABI, copies inside the timed region
    # (sanm_sweep_run_host: x_k uploaded / b_k downloaded on two copy streams
    # while other orders are being propagated)
    b2_t, b2_host = pinned_empty((ORDER + 1, n))
    b_host[:] = 0
    sweep.run_host(x_host, b2_host)  # warms the second set of device buffers
    sweep.run_host(x_host, b_host)
    b_check = sweep.get_b()
    assert np.array_equal(b_check[:, blo:bhi], b_host[:, blo:bhi]), \
        "pipelined e2e result differs"
    del b_check
    barrier()
    ev2, ev3 = torch.cuda.Event(True), torch.cuda.Event(True)
    ev2.record(stream)
    # every step uploads its series and downloads its right-hand sides; the
    # asynchronous form lets the copies of one step overlap the kernels of the
    # next (two sets of device buffers), results land in alternating host buffers
    for i in range(args.steps):
        sweep.submit_host(x_host, b_host if i % 2 == 0 else b2_host)
    sweep.wait_host()
    ev3.record(stream)
    barrier()
    ms_e2e = ev2.elapsed_time(ev3)
    if args.steps > 1:
        assert np.array_equal(b_host[:, blo:bhi], b2_host[:, blo:bhi])
    checksum = float(np.abs(b_host[:, blo:bhi]).sum())
    assert np.isfinite(checksum) and checksum > 0

    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device="cuda")
    # bytes this rank moves per step (halo mode: its local ranges only)
    nbytes = torch.tensor([(xhi - xlo) * 8.0 * (ORDER + 1),
                           (bhi - blo) * 8.0 * (ORDER + 1)],
                          dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(nbytes, op=dist.ReduceOp.SUM)
    ms_total, ms_e2e = (float(v) for v in t.cpu())
    h2d_bytes, d2h_bytes = (int(v) for v in nbytes.cpu())
    # N > 1: the same sweep WITHOUT the exchange step = what one GPU does alone
    # on its 2.0 M tets, measured in this run on this box (the N = 1 line of a
    # default run is on the 10 M-tet contract config, a different per-GPU size)
    ms_local = None
    if world > 1:
        sweep.set_exchange(2)
        for _ in range(2):
            sweep.run()
        barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(stream)
        for _ in range(args.steps):
            sweep.run()
        e1.record(stream)
        barrier()
        tl = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tl, op=dist.ReduceOp.MAX)
        ms_local = float(tl.cpu()[0])
    big_row = None
    if not args.no_solves:
        # all ranks: the drivers exchange right-hand sides and K values
        del sweep
        b.call("set_stream", ctypes.c_void_p(None))
        barrier()
        big_row = big_solve(b, world)
        barrier()
        b.call("set_stream", ctypes.c_void_p(stream.cuda_stream))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    steps_done = float(n_tets) * ORDER * args.steps * world
    value = steps_done / (ms_total * 1e-3)
    e2e = steps_done / (ms_e2e * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # whole element phase of a sweep (eval, Jacobian, steps, assembly, comm)
    elem_ms = sum(prof[t][0] for t in ("taylor_next_order", "push_xi",
                                       "remap_out")) / args.steps
    algo = algo_bytes_per_sweep(n_tets, args.energy)
    # dominant kernel: the fused order step (coefficient phase k + bias phase
    # k+1 + assembly), launched for kb = 2..K; algorithmic bytes of launch kb
    # are B_elem(kb) * tets
    S = S_MIN[args.energy]
    step_ms, step_launches = prof["k_fused_step"]
    if step_launches:
        step_algo = float(n_tets) * sum(8.0 * (S * (k + 1) + 24)
                                        for k in range(2, ORDER + 1))
        kernel_name = "fused.cu StepKernel (order step kb = 2..%d)" % ORDER
        achieved = step_algo * args.steps / (step_ms * 1e-3) / 1e9
        bytes_per_launch = step_algo / (ORDER - 1)
    else:  # generic operator-list path
        kernel_name = "element kernels of one sweep (generic path)"
        achieved = algo / (elem_ms * 1e-3) / 1e9
        bytes_per_launch = None
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles",
                                         "step_kernel_traffic.json")))
        if tr.get("energy_model") == args.energy and \
                tr.get("tets") == int(n_tets):
            traffic = tr
    except Exception:
        pass
    line = {
        "metric": "tet_order_steps_per_s", "value": value,
        "unit": "tet*order/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, n_tets, world),
        "e2e": {"value": e2e, "unit": "tet*order/s",
                "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": d2h_bytes},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak,
            "traffic": traffic["dram_bytes"] if traffic else None,
            "traffic_note": (traffic["note"] if traffic else
                             "no ncu capture for this configuration"),
            "peak_source": ("MEASURED_PEAKS.json hbm_gbs (measured copy "
                            "bandwidth)" if peaks else "fallback 6650 GB/s"),
            "kernel": kernel_name,
            "algorithmic_bytes_per_launch_avg": bytes_per_launch,
            "launch_ms_avg": (step_ms / step_launches if step_launches
                              else None),
            "sweep": {"algorithmic_bytes": algo, "element_ms": elem_ms,
                      "achieved": algo / (elem_ms * 1e-3) / 1e9,
                      "frac": algo / (elem_ms * 1e-3) / 1e9 / peak},
            "state_doubles_impl": st["state_doubles"],
            "state_doubles_min": S_MIN[args.energy],
            "arena_bytes": st["arena_bytes"],
            "kernel_ms": {k: {"ms_per_sweep": v[0] / args.steps,
                              "launches_per_sweep": v[1] / args.steps}
                          for k, v in prof.items() if v[1]},
        },
    }
    ref = None
    if os.path.exists(ORACLE_LIB):
        ref = load_oracle()
    if ms_local is not None:
        per_gpu = float(n_tets) * ORDER * args.steps / (ms_local * 1e-3)
        line["weak_scaling"] = {
            "per_gpu_value_without_exchange": per_gpu,
            "efficiency_vs_same_size_single_gpu": value / (world * per_gpu),
            "note": "same sweep, same 2.0 M tets per GPU, exchange step "
                    "switched off (slowest rank); the N = 1 line of a default "
                    "run uses the 10 M-tet contract config instead",
            "host_placement": numa}
    if big_row is not None:
        line["solve_at_scale"] = big_row
    if world == 1 and args.sweep_x:
        # element/assembly scaling sweep over the cuboid length
        line["size_sweep"] = [
            side_energy(b, torch, stream, args.energy, peak,
                        dims=(nx, args.ny, args.nz))
            for nx in (51, 101, 201, 401) if nx != args.nx]
    if world == 1 and not args.no_energies:
        line["energies"] = [side_energy(b, torch, stream, e, peak)
                            for e in ("neohookean_i", "arap")
                            if e != args.energy]
    if world == 1 and not args.no_solves:
        fp64_peak = fp64_peak_tflops(torch)
        # before the CPU sweep below, for the reason given in fea_solves
        b.call("set_stream", ctypes.c_void_p(None))
        # (the CPU side of every config is the reference application itself,
        # ref_fea, below -- not the Python front end on the oracle binding)
        line["solves"] = fea_solves(b, None, host_threads())
        # solver rooflines: back-solve bytes against the HBM peak,
        # factorisation flops against the DGEMM yardstick measured just now
        rl = []
        rows = list(line["solves"])
        if big_row is not None and big_row.get("sparse"):
            rows.append({"config": big_row["config"],
                         "gpu": {"sparse": big_row["sparse"]}})
        for row in rows:
            sp = row.get("gpu", {}).get("sparse")
            if not sp:
                continue
            if "solve_gb_s" in sp:
                rl.append({"config": row["config"],
                           "kernel": "sparse_lu.cu Forward/Backward (one "
                                     "back-solve)", "bound": "hbm",
                           "achieved": sp["solve_gb_s"], "peak": peak,
                           "unit": "GB/s", "frac": sp["solve_gb_s"] / peak})
            if "factor_tflop_s" in sp:
                rl.append({"config": row["config"],
                           "kernel": "sparse_lu.cu numeric factorisation",
                           "bound": "fp64", "achieved": sp["factor_tflop_s"],
                           "peak": fp64_peak, "unit": "TFLOP/s",
                           "frac": sp["factor_tflop_s"] / fp64_peak,
                           "peak_source": "cuBLAS DGEMM 8192^3 measured in "
                                          "this run (yardstick only)"})
        line["roofline_solver"] = rl
        # the reference application itself (unmodified fea/*.cpp) on both paths
        fb = fea_binary_solves(not args.no_cpu, host_threads())
        for row in fb or []:
            dst = next((r for r in line["solves"]
                        if r["config"] == row["config"]), None)
            if dst is None:
                line["solves"].append(row)
            else:
                dst.update({k: v for k, v in row.items() if k != "config"})
    if world == 1 and not args.no_cpu and ref is not None:
        threads = host_threads()
        rs = ReferenceSweep(ref, args, threads)
        rs.run()
        sec = min(rs.run() for _ in range(2))
        ntet = rs.n_tets
        line["cpu_baseline"] = {
            "value": ntet * ORDER / sec, "unit": "tet*order/s",
            "cores": threads, "kind": "reference",
            "sample": (f"x-slab {rs.dims} = {ntet} tets of the workload's "
                       f"cuboid, order {ORDER}, {args.energy}; unmodified "
                       f"reference ParallelTaylorCoeffProp, {threads} threads "
                       f"(sys-mt{threads}), {sec:.2f} s per sweep")}
    emit_line(line)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit_line(line: dict):
    """the ONE JSON line, on the process's real stdout"""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    # native libraries (NCCL's version banner, the reference's profiler dump)
    # write to fd 1: keep stdout for the JSON line alone
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
