"""shared helpers for the parity tests (tests only)"""
import numpy as np

import cnfb200_loader
from oracle import oracle as O

cnf = cnfb200_loader.load_package()
wl = cnf.workloads

TAGS = {"RNODE": cnf.RNODE, "ICNF": cnf.FFJORD, "CondRNODE": cnf.CondRNODE}


def oracle_cfg(cid, **over):
    c = dict(wl.CONFIGS[cid])
    c.update(over)
    return O.make_config(c["nvars"], c["naug"], c["dims"], C_=c["C"], trace_mode=c["trace_mode"],
                         lambda1=c["lambda1"], lambda2=c["lambda2"], lambda3=c["lambda3"],
                         reltol=c["reltol"], abstol=c["abstol"])


def make_icnf(cid, kernel_path=0, **over):
    c = dict(wl.CONFIGS[cid])
    c.update(over)
    dims = c["dims"]
    nn = cnf.Chain(*[cnf.Dense(dims[i], dims[i + 1], "tanh") for i in range(len(dims) - 1)])
    cm = cnf.DIVecJacMatrixMode() if c["trace_mode"] == 0 else cnf.DIJacVecMatrixMode()
    sk = {}
    if c["reltol"]:
        sk["reltol"] = c["reltol"]
    if c["abstol"]:
        sk["abstol"] = c["abstol"]
    return cnf.construct(TAGS[c["tag"]], nn, c["nvars"], c["naug"], compute_mode=cm,
                         tspan=c["tspan"], sol_kwargs=sk, lambda1=c["lambda1"],
                         lambda2=c["lambda2"], lambda3=c["lambda3"],
                         resource=cnf.B200Libs(0, kernel_path), rng=np.random.default_rng(1))


def args_of(inp):
    """(ys?, ps, st) tail of the reference call signature"""
    return ((inp["ys"].T,) if inp["ys"] is not None else ()) + (inp["theta"], {})


def relerr(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))


def relnorm(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
