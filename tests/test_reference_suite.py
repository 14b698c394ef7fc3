"""The reference's OWN Catch2 suite (tests/symbolic.cpp, tests/pade.cpp,
tests/tensor.cpp of the reference checkout, compiled unmodified by cxx/Makefile
against the libsanm-compatible headers) on this library: every ComputingGraph /
ANM solver / Pade / TensorND case of the reference runs through the C ABI.

  hostemu  (CPU container) linked against the host-emulation build of the kernels
  gpu      cxx/_build/gpu_tests on the B200 (built in the container by
           __graft_entry__.build(); the binary travels with the snapshot)

All 36 test cases of the three files run.  `Tensor.SVD` calls three host-side
SVD derivative routines of libsanm/tensor_svd.h whose counterparts on the GPU
path live on the device; plain host versions for the test are in
cxx/tests_support.cpp.  tests/ano.cpp is out of scope.
"""
import os
import subprocess
import tempfile

import pytest

from conftest import HAS_CUDA, ROOT

REF = "/root/reference"
#: test cases of tests/symbolic.cpp + pade.cpp + tensor.cpp
NR_CASES = 36


def run_suite(binary, timeout):
    # the suite writes its convergence logs (log_det_*.txt ...) into the cwd
    with tempfile.TemporaryDirectory(prefix="ref_suite_") as cwd:
        p = subprocess.run([binary], stdout=subprocess.PIPE, cwd=cwd,
                           stderr=subprocess.STDOUT, text=True, timeout=timeout)
    tail = "\n".join(p.stdout.splitlines()[-60:])
    assert p.returncode == 0, tail
    assert "All tests passed" in p.stdout, tail
    assert f"in {NR_CASES} test cases" in p.stdout, tail


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "tests")),
                    reason="needs the reference checkout (build container)")
def test_reference_catch2_hostemu(hostemu):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "cxx"),
                           "-j", "8", "_build/hostemu_tests"])
    run_suite(os.path.join(ROOT, "cxx", "_build", "hostemu_tests"), 900)


@pytest.mark.gpu
def test_reference_catch2_gpu(cuda_impl):
    binary = os.path.join(ROOT, "cxx", "_build", "gpu_tests")
    if not os.path.exists(binary):
        pytest.skip("cxx/_build/gpu_tests not built (needs the reference "
                    "checkout at build time)")
    run_suite(binary, 1200)
