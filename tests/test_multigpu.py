"""K9 on real GPUs (needs >= 2 B200s; skipped otherwise): sharded solve with the NCCL per-step
error-norm allreduce == single-GPU solve of the whole batch."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["peer-memory", "nccl"])
def test_shard_invariance_two_ranks(exchange):
    """both forms of the per-attempt error-norm exchange: inside the controller kernel over NVLink peer memory (default)
    and as an ncclAllReduce launch (CNFB200_P2P=0)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29631" if exchange == "peer-memory" else "29632",
           os.path.join(ROOT, "tests", "multigpu_worker.py")]
    env = dict(os.environ)
    if exchange == "nccl":
        env["CNFB200_P2P"] = "0"
    else:
        env.pop("CNFB200_P2P", None)
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")][-1]
    res = json.loads(line[7:])
    assert res["C2"]["dsteps"] <= 1 and res["C4"]["dsteps"] <= 1
    if res["C2"]["dsteps"] == 0:
        assert res["C2"]["lp_rel"] < 1e-4
    if res["C4"]["dsteps"] == 0:
        assert res["C4"]["loss_rel"] < 1e-4 and res["C4"]["grad_rel"] < 1e-3
    cf = res["collective_failure"]
    assert cf[-1]["outcome"] == ["err", True]                         # the failing rank reports its own error
    assert all(r["outcome"] == ["ret", 3] for r in cf[:-1])             # its peers leave the solve: retcode Unstable
    assert all(r["took"] < 20.0 for r in cf)                            # promptly, not after the 60 s mailbox trap
    assert all(r["after"][0] == 0 and r["after"][2] for r in cf)        # and every handle still solves afterwards
    assert len({r["after"][1] for r in cf}) == 1
