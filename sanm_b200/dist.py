"""Multi-GPU plumbing: one process per GPU, ``torch.distributed`` for
rendezvous, the library's own NCCL communicator for the data path
(``csrc/comm.h``).

    dist.init_process_group("nccl", ...)      # by the caller (torchrun)
    sanm_b200.dist.init_comm(binding)          # after sanm_set_device()

After ``init_comm`` every solver created by this process propagates only its
contiguous shard of the tets (the split of ``libsanm/symbolic.cpp:525-529``)
and sums assembled nodal vectors over the ranks; an ``ElementSweep`` expects
descriptors that are already restricted to the local tets
(``TetrahedralMesh.make_cuboid(cube_x_range=...)``).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

_keepalive = []


def init_comm(binding, transport: str = "auto") -> None:
    """transport: "nccl" (device buffers over NVLink), "callback" (host
    buffers through torch.distributed, used by the CPU-only tests) or "auto"
    (nccl when the process group's backend is nccl)"""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        return
    if transport == "auto":
        transport = "nccl" if dist.get_backend() == "nccl" else "callback"
    if transport == "nccl":
        ids = [None]
        if rank == 0:
            buf = C.create_string_buffer(128)
            binding.call("comm_unique_id", buf)
            ids = [buf.raw]
        dist.broadcast_object_list(ids, src=0)
        binding.call("comm_init_nccl", C.c_char_p(ids[0]), C.c_int(rank),
                     C.c_int(world))
        return

    fn_t = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_uint64)

    def allreduce(buf, n):
        arr = np.ctypeslib.as_array(buf, shape=(int(n),))
        t = torch.from_numpy(arr)
        dist.all_reduce(t)

    cb = fn_t(allreduce)
    _keepalive.append(cb)
    binding.call("comm_init_callback", cb, C.c_int(rank), C.c_int(world))


def finalize_comm(binding) -> None:
    binding.call("comm_finalize")
