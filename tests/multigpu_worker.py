"""torchrun worker for tests/test_multigpu.py: K9 shard invariance.  Every rank solves its column
block with the communicator on (whole-batch error norm through the per-step allreduce); rank 0
also solves the full batch alone; results must agree (steps equal or +-1, values to 1e-4)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from helpers import cnf, make_icnf, wl
    out = {}
    for cid, B, train in (("C2", 1 << 14, False), ("C4", 1 << 12, True)):
        inp = wl.make_inputs(cid, B=B)
        lo, hi = cnf.parallel.shard_range(B, rank, world)
        icnf = make_icnf(cid)
        icnf.resource.device = local
        cnf.parallel.init_comm(icnf, rank, world)
        tail = lambda s: ((inp["ys"][s].T,) if inp["ys"] is not None else ()) + (inp["theta"], {})
        sl = slice(lo, hi)
        if train:
            val, g = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"][sl].T, *tail(sl), eps=inp["eps"][sl].T)
            st = cnf.last_stats(icnf)
            res = dict(loss=val, grad=g, naccept=st["naccept"])
        else:
            lp, _ = cnf.inference(icnf, cnf.TestMode(), inp["xs"][sl].T, *tail(sl))
            st = cnf.last_stats(icnf)
            res = dict(lp=lp, naccept=st["naccept"])
        if rank == 0:
            solo = make_icnf(cid)
            full = slice(0, B)
            if train:
                v1, g1 = cnf.loss_and_grad(solo, cnf.TrainMode(), inp["xs"].T, *tail(full), eps=inp["eps"].T)
                s1 = cnf.last_stats(solo)
                out[cid] = dict(dsteps=abs(s1["naccept"] - res["naccept"]),
                                loss_rel=abs(v1 - res["loss"]) / abs(v1),
                                grad_rel=float(np.linalg.norm(g1 - res["grad"]) / np.linalg.norm(g1)))
            else:
                lp1, _ = cnf.inference(solo, cnf.TestMode(), inp["xs"].T, *tail(full))
                s1 = cnf.last_stats(solo)
                out[cid] = dict(dsteps=abs(s1["naccept"] - res["naccept"]),
                                lp_rel=float(np.max(np.abs(lp1[lo:hi] - res["lp"]) / np.abs(lp1[lo:hi]))))
        dist.barrier()
    # a rank that fails on the host mid-solve (simulated out-of-memory state slot at its 3rd attempt) must take its
    # peers out of the solve with it: they return promptly with retcode Unstable instead of spinning on its mailbox
    import time
    inp = wl.make_inputs("C2", B=1 << 14)
    lo, hi = cnf.parallel.shard_range(1 << 14, rank, world)
    icnf = make_icnf("C2")
    icnf.resource.device = local
    cnf.parallel.init_comm(icnf, rank, world)
    os.environ["CNFB200_TEST_SLOT_FAIL"] = f"{world - 1}:2"
    t0 = time.perf_counter()
    try:
        cnf.inference(icnf, cnf.TestMode(), inp["xs"][lo:hi].T, inp["theta"], {})
        outcome = ("ret", cnf.last_stats(icnf)["retcode"])
    except cnf._cabi.CnfError as e:
        outcome = ("err", "rc=-4" in str(e))
    took = time.perf_counter() - t0
    del os.environ["CNFB200_TEST_SLOT_FAIL"]
    dist.barrier()
    lp, _ = cnf.inference(icnf, cnf.TestMode(), inp["xs"][lo:hi].T, inp["theta"], {})   # the handles still work
    st = cnf.last_stats(icnf)
    gathered = [None] * world
    dist.all_gather_object(gathered, dict(outcome=outcome, took=took, after=(st["retcode"], st["naccept"], bool(np.isfinite(lp).all()))))
    out["collective_failure"] = gathered
    if rank == 0:
        print("RESULT " + json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
