"""world_size-2 gloo test (CPU) of the host-side multi-rank logic: the NCCL unique id hand-off and
K9 shard invariance of the whole-batch error norm (the per-step scalar allreduce libcnfb200 does
with NCCL is emulated here with gloo on the oracle's partial sums)."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch
    from helpers import O, cnf, oracle_cfg, wl
    uid = cnf.parallel.broadcast_unique_id(lambda: bytes(range(128)), rank)
    # shard the columns; each rank evaluates its partial sum of squares, allreduce, compare
    cid, B = "C4", 96
    inp = wl.make_inputs(cid, B=B)
    lo, hi = cnf.parallel.shard_range(B, rank, world)
    cfg = oracle_cfg(cid)
    D = cfg.D
    u = np.zeros((B, D + 3), np.float32)
    u[:, :wl.CONFIGS[cid]["nvars"]] = inp["xs"]
    du = O.augmented_f(cfg, inp["theta"], 1, u[lo:hi], inp["ys"][lo:hi], inp["eps"][lo:hi])
    part = torch.tensor([float((du.astype(np.float64) ** 2).sum()), float(hi - lo)], dtype=torch.float64)
    dist.all_reduce(part)
    full = O.augmented_f(cfg, inp["theta"], 1, u, inp["ys"], inp["eps"])
    q.put((rank, uid == bytes(range(128)), float(part[0]), float((full.astype(np.float64) ** 2).sum()),
           int(part[1])))
    dist.destroy_process_group()


def test_two_rank_gloo_sharding():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, uid_ok, s_sharded, s_full, n in res:
        assert uid_ok and n == 96
        assert abs(s_sharded - s_full) <= 1e-12 * abs(s_full)
