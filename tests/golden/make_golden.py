"""Generates tests/golden/*.npz.  The reference (Julia) cannot be run here and ships no golden
vectors for this path, so these fixtures pin the *restatement*: inputs, outputs of the float64
torch restatement (oracle/ref_f64.py) replaying the FP32 oracle's accepted steps, and the FP32
oracle's own outputs and step counts at generation time.  They guard against silent drift of
either implementation.  Run from the repo root:  python tests/golden/make_golden.py
When a machine with Julia is available, oracle/dump_reference.jl produces the same file layout
from the real package and closes "parity unpinned"."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, oracle_cfg, wl  # noqa: E402
from oracle import ref_f64 as F  # noqa: E402

CASES = [("C0", 16, 1, 0), ("C1", 64, 1, 0), ("C2", 128, 0, 1), ("C3", 8, 1, 0), ("C4", 32, 1, 0),
         ("C4", 32, 0, 1)]

for cid, B, mode, tm in CASES:
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid, trace_mode=tm)
    lp, regs, z, st = O.logdensity(cfg, inp["theta"], mode, inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    dts = O.last_dts()
    r = F.run(cfg, inp["theta"], mode, inp["xs"], inp["ys"], inp["eps"], *c["tspan"],
              frozen_dts=[float(d) for d in dts], requires_grad=(mode == 1))
    out = dict(xs=inp["xs"], eps=inp["eps"], theta=inp["theta"], dts=dts, logp32=lp, regs32=regs, z32=z,
               naccept=st["naccept"], nreject=st["nreject"], logp64=r["logp"], regs64=r["regs"],
               z64=r["z"], loss64=r["loss"])
    if inp["ys"] is not None:
        out["ys"] = inp["ys"]
    if mode == 1:
        out["grad_theta64"], out["grad_x64"] = r["grad_theta"], r["grad_x"]
    name = os.path.join(ROOT, "tests", "golden", f"{cid}_B{B}_mode{mode}_tm{tm}.npz")
    np.savez_compressed(name, **out)
    print(name, st["naccept"], st["nreject"], os.path.getsize(name))
