"""Writes the inputs of every committed fixture (tests/golden/*.npz) as the .npy files oracle/dump_reference.jl reads:
<case>_xs.npy (nvars x B), <case>_eps.npy (D x B), <case>_theta.npy and, for conditional flows, <case>_ys.npy (C x B) —
Julia orientation (column = sample), so NPZ.npzread returns the matrices the reference expects.
    python tests/golden/export_inputs.py tests/golden/ref_inputs"""
import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def main(out):
    os.makedirs(out, exist_ok=True)
    for path in sorted(glob.glob(os.path.join(HERE, "C*_B*_mode*_tm*.npz"))):
        if path.endswith("_ref.npz"):
            continue
        name = os.path.basename(path)[:-4]
        g = np.load(path)
        np.save(os.path.join(out, f"{name}_xs.npy"), np.ascontiguousarray(g["xs"].T))
        np.save(os.path.join(out, f"{name}_eps.npy"), np.ascontiguousarray(g["eps"].T))
        np.save(os.path.join(out, f"{name}_theta.npy"), g["theta"])
        if "ys" in g:
            np.save(os.path.join(out, f"{name}_ys.npy"), np.ascontiguousarray(g["ys"].T))
        print(name, g["xs"].shape, "->", out)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "ref_inputs"))
