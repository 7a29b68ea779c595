"""N2 measurement: the reference's fit loop (README shape: RNODE nvars=1 naugs=1, Dense(2=>6)->Dense(6=>2),
Beta(2,4) n=1024, tspan (0,13), batch_size=32 default) through per-mini-batch C-ABI calls vs the
device-resident cnfb200_fit_epochs.  usage: bench_fit.py [n_epochs] [batch_size]"""
import sys, os, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cnfb200_loader
cnf = cnfb200_loader.load_package()

n_epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 3
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 32
rng = np.random.default_rng(3)
X = rng.beta(2.0, 4.0, size=(1024, 1)).astype(np.float32)


def mk():
    nn = cnf.Chain(cnf.Dense(2, 6, "tanh"), cnf.Dense(6, 2, "tanh"))
    return cnf.construct(cnf.RNODE, nn, 1, 1, tspan=(0.0, 13.0), resource=cnf.B200Libs(0), rng=np.random.default_rng(11))


out = {}
for name in ("per_batch_calls", "device_loop"):
    m = mk()
    cnf.seed_device_rng(m, 77)
    model = cnf.ICNFModel(m, n_epochs=n_epochs, batch_size=bs, optimizers=(cnf.Lion(),))
    if name == "per_batch_calls":
        model.optimizers[0].m = np.zeros(cnf.param_offsets(m.dims)[2], np.float32)
    cnf.fit(cnf.ICNFModel(mk(), n_epochs=1, batch_size=512), 0, X[:512])   # warm up the library
    t = time.perf_counter()
    _, _, rep = cnf.fit(model, 0, X)
    dt = time.perf_counter() - t
    steps = n_epochs * -(-1024 // bs)
    out[name] = {"seconds": dt, "minibatches": steps, "ms_per_minibatch": 1e3 * dt / steps, "loss_last": rep["loss_history"][-1]}
out["speedup"] = out["per_batch_calls"]["seconds"] / out["device_loop"]["seconds"]
print(json.dumps(out))
