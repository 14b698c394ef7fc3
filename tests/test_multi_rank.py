"""Multi-GPU path with world_size 2: the tets are sharded over the ranks and
the assembled nodal vectors / K values are summed across them (csrc/comm.h).

CPU: two processes over gloo with the host-emulation library (callback
transport).  ``-m gpu``: two processes over NCCL on the GPUs of the box (both
ranks share a GPU when only one is present).

Checks, for every rank: the sharded solve reaches the same solution as a
single-rank run (1e-7 relative, ANM iterations within one), the sharded
element sweep reproduces the single-rank right-hand sides to 1e-12 relative,
and all ranks hold bitwise identical results (they take their scalar
decisions independently, so this is what keeps them in lock step)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from sanm_b200 import fea, material as mat
from sanm_b200.capi import ElementSweep

import test_fea_parity as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORLD = 2


def launch(tmp_path, impl_name, port):
    out = str(tmp_path / "res")
    procs = []
    for r in range(WORLD):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(WORLD),
                   LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OUT=out, SANM_TEST_IMPL=impl_name,
                   SANM_NO_FUSED="0", SANM_NO_FUSED_IO="0")
        procs.append(subprocess.Popen(
            [sys.executable, os.path.join(ROOT, "tests",
                                          "multi_rank_worker.py")],
            env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    logs = [p.communicate(timeout=600)[0].decode() for p in procs]
    for p, log in zip(procs, logs):
        assert p.returncode == 0, log[-3000:]
    return [np.load(out + f".rank{r}.npz") for r in range(WORLD)]


def check(impl, res):
    dims, order = (9, 4, 4), 8
    for energy in (mat.NEOHOOKEAN_C, mat.ARAP):
        it, x, _ = T.solve_eqn(impl, dims, energy, order)
        _, _, model = T.cuboid_body(impl, dims, energy)
        xs = T.series(model, order)
        sw = ElementSweep(model.y, model.lt_inp.desc, model.lt_out, order)
        sw.set_x(xs)
        sw.run()
        b1 = sw.get_b()
        for r in range(WORLD):
            # the summation order of the shards differs from the single-rank
            # run at rounding level, which may move a Pade range estimate
            # (see test_fea_parity.test_eqn_solve_cuboid): same fixed point,
            # iteration count within one
            assert abs(int(res[r][f"solve_it_{energy}"]) - it) <= 1
            xr = res[r][f"solve_x_{energy}"]
            assert np.abs(xr - x).max() <= 1e-7 * np.abs(x).max()
            br = res[r][f"sweep_b_{energy}"]
            for k in range(order + 1):
                scale = max(np.abs(b1[k]).max(), 1e-300)
                assert np.abs(br[k] - b1[k]).max() <= 1e-12 * scale, (energy, k)
            # lock step: every rank has exactly the same numbers
            assert np.array_equal(xr, res[0][f"solve_x_{energy}"])
            assert np.array_equal(br, res[0][f"sweep_b_{energy}"])


def test_two_ranks_gloo(hostemu, tmp_path):
    res = launch(tmp_path, "hostemu", 29731)
    check(hostemu, res)


@pytest.mark.gpu
def test_two_ranks_nccl(cuda_impl, tmp_path):
    import torch
    if torch.cuda.device_count() < WORLD:
        pytest.skip("NCCL needs one GPU per rank (run under gpurun --gpus 2)")
    res = launch(tmp_path, "cuda", 29741)
    check(cuda_impl, res)
