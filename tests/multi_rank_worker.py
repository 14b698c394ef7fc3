"""Worker of tests/test_multi_rank.py: one of WORLD_SIZE processes (gloo on
CPU with the host-emulation library; NCCL with the CUDA library under
``-m gpu``).  Writes its results to $OUT.rank<r>.npz."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    from sanm_b200 import dist as sdist, fea, material as mat
    from sanm_b200.capi import ANMEqnSolver, Binding, ElementSweep, Hyper
    import test_fea_parity as T

    use_cuda = os.environ.get("SANM_TEST_IMPL") == "cuda"
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    if use_cuda:
        import sanm_b200
        torch.cuda.set_device(rank % torch.cuda.device_count())
        dist.init_process_group("nccl", device_id=torch.device(
            "cuda", rank % torch.cuda.device_count()))
        b = sanm_b200.load()
        b.call("set_device", ctypes.c_int(rank % torch.cuda.device_count()))
    else:
        dist.init_process_group("gloo")
        import hostemu_build
        b = Binding(ctypes.CDLL(hostemu_build.build()), "sanm_")
    sdist.init_comm(b)

    dims, order = (9, 4, 4), 8
    out = {}
    for energy in (mat.NEOHOOKEAN_C, mat.ARAP):
        # (B) full solve, global descriptors, sharded inside the library
        it, x, rms = T.solve_eqn(b, dims, energy, order)
        out[f"solve_it_{energy}"] = it
        out[f"solve_x_{energy}"] = x
        # (A) element sweep over pre-sharded descriptors
        nx, ny, nz = dims
        c0 = (nx - 1) * rank // world
        c1 = (nx - 1) * (rank + 1) // world
        mesh = fea.TetrahedralMesh.make_cuboid(nx, ny, nz, 0.025,
                                               cube_x_range=(c0, c1))
        body = fea.DeformableBody(
            b, mat.MaterialProperty.from_young_poisson(1e7, 0.45), mesh)
        body.coord_fixed_mask[mesh.vertices[:, 0] <= 0.0125] = True
        model = body.make_forward(energy)
        x = T.series(model, order)
        sw = ElementSweep(model.y, model.lt_inp.desc, model.lt_out, order)
        sw.set_x(x)
        sw.run()
        out[f"sweep_b_{energy}"] = sw.get_b()
        # distributed nodal vectors: halo sums instead of the all-reduce; the
        # local range must carry the same bits (also through run_host, which
        # moves the local ranges only)
        full = out[f"sweep_b_{energy}"]
        sw.set_exchange(1)
        xlo, xhi, blo, bhi = sw.local_ranges()
        assert 0 <= xlo < xhi <= x.shape[1] and 0 <= blo < bhi <= full.shape[1]
        sw.run()
        assert np.array_equal(sw.get_b()[:, blo:bhi], full[:, blo:bhi]), energy
        got = sw.run_host(x, np.zeros_like(full))
        assert np.array_equal(got[:, blo:bhi], full[:, blo:bhi]), energy
        out[f"sweep_range_{energy}"] = np.array([xlo, xhi, blo, bhi])
    np.savez(os.environ["OUT"] + f".rank{rank}.npz", **out)
    sdist.finalize_comm(b)
    dist.destroy_process_group()
    os._exit(0)


if __name__ == "__main__":
    main()
