"""The drop-in boundary: the CUDA shared library loads without a GPU and
exports every symbol that include/sanm_b200.h declares; without a device every
compute entry point fails loudly (there is no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import HAS_CUDA, ROOT


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "sanm_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(sanm_[a-z0-9_]+)\s*\(", hdr)))


@pytest.fixture(scope="module")
def cuda_lib():
    from sanm_b200 import build
    path = build.LIB
    if not os.path.exists(path):
        path = build.build()
    return ctypes.CDLL(path)


def test_header_symbols_exported(cuda_lib, hostemu):
    names = declared_symbols()
    assert len(names) > 60
    for n in names:
        assert hasattr(cuda_lib, n), f"{n} missing from libsanm_b200.so"
        assert hasattr(hostemu.lib, n), f"{n} missing from the host emulation"


@pytest.mark.skipif(HAS_CUDA, reason="needs a machine without a CUDA device")
def test_no_cpu_fallback(cuda_lib):
    """without a device the product library refuses to compute"""
    cuda_lib.sanm_last_error.restype = ctypes.c_char_p
    n = ctypes.c_uint64(4)
    h = ctypes.c_void_p()
    st = cuda_lib.sanm_sparse_create(n, ctypes.byref(h))
    if st == 0:
        x = (ctypes.c_double * 4)(1, 2, 3, 4)
        idx = (ctypes.c_uint64 * 4)(0, 1, 2, 3)
        st = cuda_lib.sanm_sparse_add_constraints(h, n, idx, idx, x)
        if st == 0:
            st = cuda_lib.sanm_sparse_prepare(h, ctypes.c_double(0))
    assert st == 4, cuda_lib.sanm_last_error()   # SANM_ERR_CUDA
    assert b"CUDA device" in cuda_lib.sanm_last_error()
