"""Caller-side adapters, mirrored so the path can be exercised end to end from Python the way the
reference's users (and its regression test) drive it.  In a Julia deployment these stay the
reference's own code (src/exts/mlj_ext, src/exts/dist_ext) and only call the overloaded
inference / generate / loss.

  ICNFModel / CondICNFModel, fit, transform     src/exts/mlj_ext/core_icnf.jl:1-146, core_cond_icnf.jl
  ICNFDist / CondICNFDist: logpdf, pdf, rand     src/exts/dist_ext/core_icnf.jl:1-58, core_cond_icnf.jl
  Lion                                           Optimisers.Lion (default optimiser of ICNFModel)
"""
import time

import numpy as np

from . import api


class Lion:
    """Optimisers.Lion(eta = 0.001, beta = (0.9, 0.999)): x -= eta * sign(b1 m + (1-b1) g);
    m = b2 m + (1-b2) g"""

    def __init__(self, eta=1e-3, beta=(0.9, 0.999)):
        self.eta, self.beta, self.m = eta, beta, None

    def update(self, ps, g):
        if self.m is None:
            self.m = np.zeros_like(ps)
        b1, b2 = self.beta
        ps = ps - self.eta * np.sign(b1 * self.m + (1 - b1) * g)
        self.m = b2 * self.m + (1 - b2) * g
        return ps.astype(np.float32)


class ICNFModel:
    """mirror of `ICNFModel(m, loss; optimizers, n_epochs, adtype, use_batch, batch_size, sol_kwargs)`"""
    cond = False

    def __init__(self, m, loss=None, *, optimizers=None, n_epochs=300, adtype=None, use_batch=True,
                 batch_size=32, sol_kwargs=None):
        self.m, self.loss = m, loss
        self.optimizers = optimizers if optimizers is not None else (Lion(),)
        self.n_epochs, self.use_batch, self.batch_size = n_epochs, use_batch, batch_size
        self.adtype, self.sol_kwargs = adtype, dict(sol_kwargs or {})
        self.fitresult, self.report = None, None


class CondICNFModel(ICNFModel):
    cond = True


def fit(model, verbosity, X, Y=None):
    """MLJModelInterface.fit(model, verbosity, X[, Y]): X is samples x features (table orientation);
    DataLoader(batchsize, shuffle = true, partial = true); per optimiser `n_epochs` epochs of
    (loss, grad) -> update.  Returns (fitresult = (ps, st), cache = None, report)."""
    m = model.m
    x = np.ascontiguousarray(np.asarray(X, np.float32))            # (N, nvars) == nvars x N column-major
    y = None if Y is None else np.ascontiguousarray(np.asarray(Y, np.float32))
    ps, st = api.setup(m.rng, m)
    N = x.shape[0]
    bs = model.batch_size if model.use_batch else N
    t0 = time.perf_counter()
    history = []
    device_loop = getattr(m, "_device_rng", False) and m.steer_rate == 0 and all(isinstance(o, Lion) and o.m is None for o in model.optimizers)
    for opt in model.optimizers:
        if device_loop:   # N2: the whole epoch loop in one library call (steering is per call there)
            perm = np.stack([m.rng.permutation(N) for _ in range(model.n_epochs)])
            ps, hist = api.fit_epochs(m, x.T, y.T if model.cond else None, ps, batch_size=bs, n_epochs=model.n_epochs,
                                      perm=perm, eta=opt.eta, beta=opt.beta)
            history.extend(float(v) for v in hist)
            continue
        for epoch in range(model.n_epochs):
            perm = m.rng.permutation(N)
            tot, cnt = 0.0, 0
            for lo in range(0, N, bs):
                idx = perm[lo:lo + bs]
                args = ((y[idx].T,) if model.cond else ()) + (ps, st)
                val, g = api.loss_and_grad(m, api.TrainMode(), x[idx].T, *args)
                ps = opt.update(ps, g)
                tot, cnt = tot + val * len(idx), cnt + len(idx)
            history.append(tot / cnt)
            if verbosity > 1:
                print(f"epoch {epoch}: loss {history[-1]:.5f}")
    model.fitresult = (ps, st)
    model.report = {"stats": {"time": time.perf_counter() - t0}, "loss_history": history}
    return model.fitresult, None, model.report


def transform(model, fitresult, Xnew, Ynew=None):
    """MLJModelInterface.transform: DataFrame(px = exp.(logp̂x)) -> returns the px vector"""
    ps, st = fitresult
    x = np.ascontiguousarray(np.asarray(Xnew, np.float32))
    args = ((np.asarray(Ynew, np.float32).T,) if model.cond else ()) + (ps, st)
    logp, _ = api.inference(model.m, api.TestMode(), x.T, *args)
    return np.exp(np.asarray(logp, np.float64))


class ICNFDist:
    """ICNFDist(m, mode, ps, st) / ICNFDist(mach, mode): a ContinuousMultivariateDistribution"""

    def __init__(self, m, mode, ps=None, st=None):
        if isinstance(m, ICNFModel):
            ps, st = m.fitresult
            m = m.m
        self.m, self.mode, self.ps, self.st = m, mode, ps, st if st is not None else {}

    def _tail(self, n=None):
        return (self.ps, self.st)

    def logpdf(self, A):
        A = np.asarray(A, np.float32)
        if A.ndim == 1:
            A = A[:, None]
        return np.asarray(api.inference(self.m, self.mode, A, *self._tail(A.shape[1]))[0])

    def pdf(self, A):
        return np.exp(self.logpdf(A).astype(np.float64))

    def rand(self, n=None):
        out = np.asarray(api.generate(self.m, self.mode, *self._tail(n or 1), n or 1))
        return out[:, 0] if n is None else out


class CondICNFDist(ICNFDist):
    """CondICNFDist(m, mode, ys, ps, st); rand(n) uses ys[:, 1:n] (core_cond_icnf.jl:57)"""

    def __init__(self, m, mode, ys, ps=None, st=None):
        super().__init__(m, mode, ps, st)
        self.ys = np.asarray(ys, np.float32)

    def _tail(self, n=None):
        ys = self.ys if n is None else self.ys[:, :n]
        return (ys, self.ps, self.st)
