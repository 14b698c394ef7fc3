#!/usr/bin/env python
"""Regenerate the golden fixtures from the oracle (the unmodified reference,
oracle/_ref/libref_capi.so; build it with `make -C oracle`).

    python tests/golden/make_golden.py

Writes tests/golden/*.npz.  The inputs are seeded, so the script is
reproducible; tests/test_golden.py checks the implementation under test (and,
when present, the oracle itself) against these files.
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from sanm_b200 import material as mat  # noqa: E402
from sanm_b200.capi import Binding, reference_sweep  # noqa: E402
import test_fea_parity as F  # noqa: E402
import test_taylor_parity as T  # noqa: E402

ENERGIES = (mat.NEOHOOKEAN_C, mat.NEOHOOKEAN_I, mat.ARAP)
SWEEP_DIMS, SWEEP_ORDER = (5, 4, 3), 10
SOLVE_DIMS, SOLVE_ORDER = (7, 3, 3), 10
PROP_N, PROP_ORDER = 5, 12


def prop_inputs(energy):
    rng = np.random.default_rng(20211)
    xs = T.make_F_coeffs(rng, PROP_N, PROP_ORDER)
    dminv = np.eye(3)[None] + rng.uniform(-0.1, 0.1, (PROP_N, 3, 3))
    bias = rng.uniform(-0.05, 0.05, (PROP_N, 3, 3))

    def build(g):
        Ds = g.placeholder() + g.constant(bias)
        F_ = Ds.batched_matmul(g.constant(dminv))
        return mat.pk1(energy, T.MATERIAL, F_)
    return xs, build


def main():
    ref = Binding(ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref",
                                           "libref_capi.so")), "ref_")
    out = {}
    for e in ENERGIES:
        # Taylor propagation of the stress graph (check_taylor_prop,
        # tests/symbolic.cpp:89-137)
        xs, build = prop_inputs(e)
        y0, biases, ys, jac = T.run_prop(ref, build, xs, PROP_ORDER)
        out[f"prop_y0_{e}"] = y0
        out[f"prop_bias_{e}"] = np.array(biases)
        out[f"prop_y_{e}"] = np.array(ys)
        out[f"prop_jac_{e}"] = jac
        # element sweep on a cuboid (anm.cpp:205-290 without the solves)
        _, _, model = F.cuboid_body(ref, SWEEP_DIMS, e)
        x = F.series(model, SWEEP_ORDER)
        out[f"sweep_b_{e}"] = reference_sweep(model.y, model.lt_inp.desc,
                                              model.lt_out, x)
        # full solve f(x) + load = 0
        it, xsol, rms = F.solve_eqn(ref, SOLVE_DIMS, e, SOLVE_ORDER)
        out[f"solve_it_{e}"] = it
        out[f"solve_x_{e}"] = xsol
        out[f"solve_rms0_{e}"] = rms[0]
    np.savez_compressed(os.path.join(HERE, "fea_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "fea_golden.npz"),
          sum(v.nbytes for v in map(np.asarray, out.values())), "bytes raw")
    os._exit(0)


if __name__ == "__main__":
    main()
