"""pytest configuration.

Markers
    gpu   needs a CUDA device (run on the B200 box: ``pytest -m gpu``)

Fixtures
    impl  the implementation under test.  Under ``-m gpu`` this is the product
          CUDA library (``sanm_b200.load()``).  In the CPU-only container it is
          the HOST-EMULATION build of the same kernel sources
          (``tests/_hostemu``, compiled with ``-DSANM_HOSTEMU`` by
          ``tests/hostemu_build.py``): test infrastructure that checks kernel and
          host logic against the oracle without a GPU.  It is never shipped or
          loaded by the package.
    ref   the oracle: the unmodified reference built into ``oracle/_ref``
"""
import ctypes
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


def _has_cuda():
    try:
        lib = ctypes.CDLL("libcudart.so")
    except OSError:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:
            return False
    n = ctypes.c_int(0)
    return lib.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0


HAS_CUDA = _has_cuda()


@pytest.fixture(scope="session")
def ref():
    from sanm_b200.capi import Binding
    path = os.path.join(ROOT, "oracle", "_ref", "libref_capi.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libref_capi.so not built (run make -C oracle)")
    return Binding(ctypes.CDLL(path), "ref_")


@pytest.fixture(scope="session")
def hostemu():
    """kernel-logic emulation on the host (CPU container only)"""
    from sanm_b200.capi import Binding
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import hostemu_build
    path = hostemu_build.build()
    return Binding(ctypes.CDLL(path), "sanm_")


@pytest.fixture(scope="session")
def cuda_impl():
    import sanm_b200
    if not HAS_CUDA:
        pytest.skip("no CUDA device")
    return sanm_b200.load()


@pytest.fixture(scope="session",
                params=["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)])
def impl(request):
    """every parity test runs twice: `hostemu` (CPU container, kernel logic) and
    `cuda` (marked gpu: the product library on the B200, through the C ABI)"""
    if request.param == "cuda":
        return request.getfixturevalue("cuda_impl")
    return request.getfixturevalue("hostemu")
