"""Independent float64 restatement of the path in torch (matrix form, autograd-capable).
TEST INFRASTRUCTURE ONLY (same rules as oracle.py).

Purpose: (i) a second, structurally different implementation (whole-batch
matrix algebra instead of per-tile loops) to cross-check cnf_oracle.c; (ii) the gradient oracle:
autograd through the discrete Tsit5 recursion with the FP32 run's accepted dt sequence frozen.
Cites: src/icnf.jl:148-164, 318-350, 384-420, 481-502; src/base_icnf.jl:167-189, 266-309.
"""
import math

import numpy as np
import torch

A = [
    [],
    [0.161],
    [-0.008480655492356989, 0.335480655492357],
    [2.8971530571054935, -6.359448489975075, 4.3622954328695815],
    [5.325864828439257, -11.748883564062828, 7.4955393428898365, -0.09249506636175525],
    [5.86145544294642, -12.92096931784711, 8.159367898576159, -0.071584973281401,
     -0.028269050394068383],
    [0.09646076681806523, 0.01, 0.4798896504144996, 1.379008574103742, -3.290069515436081,
     2.324710524099774],
]
BT = [-0.00178001105222577714, -0.0008164344596567469, 0.007880878010261995,
      -0.1447110071732629, 0.5823571654525552, -0.45808210592918697, 0.015151515151515152]


def unpack(theta, dims, planar=False):
    """flat Lux ComponentArray layout -> [(W (dout,din), b (dout,))]; PlanarLayer (u, w, b) ->
    [(w' (1,d0), b), (u (D,1), None)] with a linear, bias-free second product"""
    if planar:
        d0, D = dims[0], dims[2]
        return [(theta[D:D + d0].reshape(1, d0), theta[D + d0:D + d0 + 1]), (theta[:D].reshape(D, 1), None)]
    out, o = [], 0
    for l in range(len(dims) - 1):
        din, dout = dims[l], dims[l + 1]
        W = theta[o:o + din * dout].reshape(din, dout).T  # column-major dout x din
        o += din * dout
        b = theta[o:o + dout]
        o += dout
        out.append((W, b))
    return out


def mlp(layers, h):
    """h: (B, d0) -> (B, dL); every Dense layer has tanh (exact tanh here)"""
    for W, b in layers:
        h = torch.tanh(h @ W.T + b) if b is not None else h @ W.T
    return h


def aug_f(layers, D, mode, trace_mode, flags, z, ys, eps):
    """returns du columns (B, R) given z (B, D).  mode 0 = Test (exact trace), 1 = Train."""
    def f(zz):
        inp = zz if ys is None else torch.cat([zz, ys], dim=1)
        return mlp(layers, inp)

    if mode == 0:
        zz = z if z.requires_grad else z.detach().requires_grad_(True)
        with torch.enable_grad():
            y = f(zz)
            tr = 0
            for i in range(D):
                gi, = torch.autograd.grad(y[:, i].sum(), zz, create_graph=z.requires_grad,
                                          retain_graph=True)
                tr = tr + gi[:, i]
        return torch.cat([y, -tr[:, None]], dim=1)
    zz = z if z.requires_grad else z.detach().requires_grad_(True)
    with torch.enable_grad():
        if trace_mode == 0:  # VecJac: v = J^T eps
            y = f(zz)
            v, = torch.autograd.grad(y, zz, eps, create_graph=True)
        else:  # JacVec: v = J eps
            y, v = torch.autograd.functional.jvp(f, zz, eps, create_graph=True)
    ldot = -(v * eps).sum(1, keepdim=True)
    E = y.norm(dim=1, keepdim=True) if flags & 4 else torch.zeros_like(ldot)
    n = v.norm(dim=1, keepdim=True) if flags & 8 else torch.zeros_like(ldot)
    out = torch.cat([y, ldot, E, n], dim=1)
    return out if z.requires_grad else out.detach()


def solve(rhs, u, t0, t1, reltol=1e-3, abstol=1e-6, maxiters=100000, frozen_dts=None):
    """Tsit5 with the whole-matrix RMS norm + PI controller in float64,
    or a replay of `frozen_dts` (accepted steps only).  Returns u_final, stats."""
    tdir = 1.0 if t1 >= t0 else -1.0
    if frozen_dts is not None:
        for dt in frozen_dts:
            k = [rhs(u)]
            for i in range(1, 6):
                s = u + dt * sum(a * kk for a, kk in zip(A[i], k))
                k.append(rhs(s))
            u = u + dt * sum(a * kk for a, kk in zip(A[6], k))
        return u, dict(naccept=len(frozen_dts), nreject=0)
    N = u.numel()
    rms = lambda x: float(torch.sqrt((x.detach() * x.detach()).sum() / N))
    dtmax = abs(t1 - t0)
    k1 = rhs(u)
    sk = abstol + u.abs() * reltol
    d0, d1 = rms(u / sk), rms(k1 / sk)
    dt0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
    dt0 = min(dt0, dtmax)
    f1 = rhs(u + tdir * dt0 * k1)
    d2 = rms((f1 - k1) / sk) / dt0
    mx = max(d1, d2)
    dt1 = max(1e-6, dt0 * 1e-3) if mx <= 1e-15 else 10 ** (-(2 + math.log10(mx)) / 5)
    dt = tdir * min(100 * dt0, dt1, dtmax)
    t, qold, nacc, nrej, it, dts = t0, 1e-4, 0, 0, 0, []
    while t != t1:
        it += 1
        if it > maxiters:
            break
        dt = tdir * min(abs(dt), dtmax, abs(t1 - t))
        k = [k1]
        for i in range(1, 6):
            k.append(rhs(u + dt * sum(a * kk for a, kk in zip(A[i], k))))
        un = u + dt * sum(a * kk for a, kk in zip(A[6], k))
        k.append(rhs(un))
        ut = dt * sum(b * kk for b, kk in zip(BT, k))
        EEst = rms(ut / (abstol + torch.maximum(u.abs(), un.abs()) * reltol))
        if EEst == 0:
            q11, q = 0.0, 0.1
        else:
            q11 = EEst ** 0.14
            q = max(0.1, min(5.0, q11 / qold ** 0.08 / 0.9))
        if EEst <= 1:
            dts.append(dt)
            tn = t + dt
            t = t1 if abs(tn - t1) < 100 * 2.2e-16 * max(abs(t), abs(t1)) else tn
            u, k1 = un, k[6]
            qold = max(EEst, 1e-4)
            dt = dt / q
            nacc += 1
        else:
            dt = dt / min(5.0, q11 / 0.9)
            nrej += 1
    return u, dict(naccept=nacc, nreject=nrej, dts=dts)


def run(cfg, theta, mode, xs, ys=None, eps=None, t0=0.0, t1=1.0, frozen_dts=None,
        requires_grad=False, reltol=None, abstol=None):
    """float64 logdensity / loss.  Returns dict(logp, regs, z, loss, stats[, grad_theta, grad_x])."""
    dims = list(cfg.dims[:cfg.n_layers + 1])
    D, nv = cfg.D, cfg.nvars
    th = torch.tensor(np.asarray(theta, np.float64), requires_grad=requires_grad)
    x = torch.tensor(np.asarray(xs, np.float64), requires_grad=requires_grad)
    y = None if ys is None else torch.tensor(np.asarray(ys, np.float64))
    e = None if eps is None else torch.tensor(np.asarray(eps, np.float64))
    layers = unpack(th, dims, bool(cfg.flags & 32))
    B = x.shape[0]
    R = D + (3 if mode == 1 else 1)
    u0 = torch.cat([x, torch.zeros(B, R - nv, dtype=torch.float64)], dim=1)
    flags = cfg.flags

    def rhs(u):
        z = u[:, :D]
        if requires_grad and not z.requires_grad:
            z = z.requires_grad_(True)
        return aug_f(layers, D, mode, cfg.trace_mode, flags, z, y, e)

    rt = reltol if reltol is not None else (cfg.reltol if cfg.reltol > 0 else 1e-3)
    at = abstol if abstol is not None else (cfg.abstol if cfg.abstol > 0 else 1e-6)
    u, st = solve(rhs, u0, t0, t1, rt, at, frozen_dts=frozen_dts)
    z = u[:, :D]
    logpz = -0.5 * (z * z).sum(1) - 0.5 * D * math.log(2 * math.pi)
    logp = logpz - u[:, D]
    useA = (flags & 16) and (flags & 2)
    Aa = z[:, nv:].norm(dim=1) if useA else torch.zeros(B, dtype=torch.float64)
    out = dict(stats=st, z=z.detach().numpy(), logp=logp.detach().numpy())
    if mode == 1:
        regs = torch.stack([u[:, D + 1], u[:, D + 2], Aa], dim=1)
        loss = (-logp + cfg.lambda1 * regs[:, 0] + cfg.lambda2 * regs[:, 1]
                + cfg.lambda3 * regs[:, 2]).mean()
    else:
        regs = Aa[:, None]
        loss = (-logp).mean()
    out["regs"] = regs.detach().numpy()
    out["loss"] = float(loss.detach())
    if requires_grad:
        gth, gx = torch.autograd.grad(loss, [th, x])
        out["grad_theta"], out["grad_x"] = gth.numpy(), gx.numpy()
    return out
