"""sanm_b200 -- B200-native (sm_100a, FP64) implementation of SANM's ANM inner
loop behind the libsanm API surface.

The compute path is the CUDA library ``sanm_b200/lib/libsanm_b200.so`` (built by
``sanm_b200/build.py`` / ``__graft_entry__.build()``), reached through the C ABI
declared in ``include/sanm_b200.h``.  There is no CPU fallback: :func:`load`
raises if the library is missing and every entry point fails with
``SANMCudaError`` when no CUDA device is present.
"""
from __future__ import annotations

import ctypes
import os

from . import capi
from .capi import (ANMEqnSolver, ANMImplicitSolver, ANMSolverVecScale, Binding,
                   ComputingGraph, Hyper, SANMAssertionError, SANMCudaError,
                   SANMError, SANMNumericalError, SparseLinearDesc,
                   SparseSolver, SymbolVar, TaylorCoeffProp)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libsanm_b200.so")
_binding = None


def load() -> Binding:
    """Load the CUDA library (once) and return its binding."""
    global _binding
    if _binding is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m "
                "sanm_b200.build` (needs nvcc); sanm_b200 has no CPU fallback")
        # load every kernel image when the library is first used instead of at
        # each kernel's first launch: a solve then never pays module loading
        # (has to be set before the CUDA context exists to take effect)
        os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
        _binding = Binding(ctypes.CDLL(LIB_PATH), "sanm_")
    return _binding
