"""Batch sharding over one-process-per-GPU ranks.  Columns (samples) are split in
contiguous blocks; theta is replicated.  The only data-path exchange is the scalar error-norm
exchange every attempted step (whole-batch RMS norm of the matrix-valued ODE state, reference
src/base_icnf.jl:280-285; inside the controller kernel over NVLink peer memory, NCCL as the fallback) plus the loss/gradient allreduce in training; both happen inside
libcnfb200 on its NCCL communicator.  torch.distributed is only used to hand the NCCL unique id
from rank 0 to the other ranks."""
import ctypes as C

import numpy as np

from . import _cabi


def shard_range(B, rank, world):
    """contiguous column block [lo, hi) of rank `rank` out of `world` (first B % world ranks get one more)"""
    base, rem = divmod(B, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def broadcast_unique_id(make_id, rank, group=None):
    """rank 0 calls make_id() -> 128 bytes; every rank returns the same bytes (any backend)."""
    import torch
    import torch.distributed as dist
    buf = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        buf = torch.frombuffer(bytearray(make_id()), dtype=torch.uint8).clone()
    if dist.get_backend(group) == "nccl":
        buf = buf.cuda()
    dist.broadcast(buf, src=0, group=group)
    return bytes(buf.cpu().numpy().tobytes())


def init_comm(icnf, rank=None, world=None):
    """give `icnf`'s handle an NCCL communicator over all ranks of the default process group"""
    import torch.distributed as dist
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    if world == 1:
        return
    lib = _cabi.load()

    def make_id():
        b = C.create_string_buffer(128)
        rc = lib.cnfb200_comm_unique_id(b)
        if rc != 0:
            raise _cabi.CnfError(f"cnfb200_comm_unique_id failed (rc={rc})")
        return b.raw

    uid = broadcast_unique_id(make_id, rank)
    h = icnf.handle()
    icnf._check(lib.cnfb200_comm_init(h, world, rank, uid), "cnfb200_comm_init")


def errnorm_exchange(icnf):
    """how the per-attempt error-norm partials travel: 'peer-memory' (inside the controller kernel, P2P stores
    over NVLink) or 'nccl' (an ncclAllReduce launch per attempt)"""
    return "peer-memory" if _cabi.load().cnfb200_comm_p2p(icnf.handle()) else "nccl"
