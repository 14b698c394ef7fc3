"""sanm_remap_apply_static: the allocation-free variant of sanm_remap_apply
(device staging in a static device array, include/sanm_b200.h).  No
counterpart in the reference.

Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests: it only needs the
library, builds its own small maps, and never looks at earlier state of the
scratch array (the second call below deliberately follows a different map so
that stale contents would show).
"""
import ctypes

import numpy as np
import pytest

from sanm_b200.capi import SANMError, SparseLinearDesc

MAX_DOUBLES = 2097152  # SANM_REMAP_STATIC_MAX_DOUBLES


def random_map(b, nout, nin, rng, empty_every=7):
    """CSR gather with ragged rows, some of them empty"""
    counts = rng.integers(1, 6, nout)
    counts[::empty_every] = 0
    rowptr = np.zeros(nout + 1, dtype=np.uint64)
    rowptr[1:] = np.cumsum(counts)
    nnz = int(rowptr[-1])
    idx = rng.integers(0, nin, nnz).astype(np.uint64)
    coeff = rng.uniform(-2, 2, nnz)
    desc = SparseLinearDesc(b, (nout,), (nin,), rowptr, idx, coeff)
    return desc, (rowptr, idx, coeff)


def check(desc, csr, x):
    want = desc.apply(x)            # existing entry point (pool-allocated staging)
    got = desc.apply_static(x)      # static scratch
    # same kernel, same operation order: bitwise equal (the bar of the other
    # remap tests is 1e-10 relative)
    assert np.array_equal(got, want)
    rowptr, idx, coeff = csr
    ref = np.zeros(len(rowptr) - 1)
    for i in range(len(ref)):
        p, q = int(rowptr[i]), int(rowptr[i + 1])
        s = 0.0
        for k in range(p, q):
            s += x[int(idx[k])] * coeff[k]
        ref[i] = s
    scale = max(np.abs(ref).max(), 1e-300)
    assert np.abs(got - ref).max() <= 1e-10 * scale


def run_two_calls(b):
    rng = np.random.default_rng(2024)
    big, big_csr = random_map(b, 3000, 1777, rng)
    small, small_csr = random_map(b, 513, 2100, rng, empty_every=3)
    # two calls in a row on one map, then another shape over the same scratch
    # (rows without entries must read zero, not the previous call's output)
    check(big, big_csr, rng.uniform(-1, 1, 1777))
    check(big, big_csr, rng.uniform(-1, 1, 1777))
    check(small, small_csr, rng.uniform(-1, 1, 2100))
    check(big, big_csr, rng.uniform(-1, 1, 1777))


def run_limit(b):
    # round_up(nin, 32) + nout one past the limit: refused with a clear error
    nin = 64
    nout = MAX_DOUBLES - nin + 1
    ident = SparseLinearDesc(
        b, (nout,), (nin,), np.zeros(nout + 1, dtype=np.uint64),
        np.zeros(0, dtype=np.uint64), np.zeros(0))
    with pytest.raises(SANMError) as ei:
        ident.apply_static(np.zeros(nin))
    assert "static scratch" in str(ei.value)
    # exactly at the limit it runs
    nout -= 1
    fits = SparseLinearDesc(
        b, (nout,), (nin,), np.zeros(nout + 1, dtype=np.uint64),
        np.zeros(0, dtype=np.uint64), np.zeros(0))
    y = fits.apply_static(np.ones(nin))
    assert y.shape == (nout,) and not y.any()


@pytest.mark.gpu
def test_remap_apply_static_gpu(cuda_impl):
    launches0 = cuda_impl.launch_count()
    run_two_calls(cuda_impl)
    run_limit(cuda_impl)
    assert cuda_impl.launch_count() > launches0


def test_remap_apply_static_hostemu(hostemu):
    """host emulation of the same entry point (kernel logic, CPU container)"""
    run_two_calls(hostemu)
    run_limit(hostemu)
