#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native augmented-dynamics solve.

  python bench.py --gpus N --steps K --warmup W [--config C2] [--impl reference]

BASELINE.json's metric has two halves, and one JSON line (rank 0) carries both:

  top level   log-density samples/s: C2 = BASELINE.json configs[1] (ICNF exact trace, DIJacVecMatrixMode,
              MLP 2->64->64->2, B = 2^20 per GPU, weak scaling over N GPUs).  `value` = device-resident inputs,
              CUDA-event time of the calls (max over ranks); `e2e` = the same call with HOST buffers (pinned),
              host<->device copies inside the timed region; `roofline` = the RK stage kernel, ALGORITHMIC flops
               / its mean CUDA-event time / the measured bf16 tensor peak; `cpu_baseline` = the CPU
              oracle (port of the reference's CPU path) on this box's cores; `parity` = the results of the timed
              call checked against the oracle on the same inputs at the benchmarked size.
  "train"     RNODE train samples/s: C3 (configs[2], 32->256->256->32, B = 2^18 per GPU, weak scaling): one
              loss + gradient call per step (forward solve, discrete-adjoint reverse sweep, gradient allreduce),
              theta re-uploaded every step; own value / e2e / roofline / parity (fixed slice against the oracle).
  "train_strong"  C5 (configs[4], 128->1024->1024->128) at the GLOBAL batch 2^22 split over the N GPUs (strong
              scaling, gradient allreduce); "cond_strong": C4 (configs[3], CondRNODE) global batch 2^20 split N ways.

A "step" is one pass of the hot path over one batch of synthetic input.  The run exits non-zero when a parity
block is outside north_star's tolerances (the line is still printed).  --config X runs only that config as the
top-level workload (no extra blocks).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = {"test": ("logdensity_samples_per_s", "samples/s"),
          "train": ("rnode_train_samples_per_s", "samples/s")}
# north_star tolerances
TOL_LOGP, TOL_LOSS, TOL_GRAD = 1e-4, 1e-4, 1e-3
# accepted-step counts may differ by one (north_star); the SAME value tolerances apply then (measured: the two solves
# still agree to ~1e-6)
TOL_FLIP_LOGP, TOL_FLIP_LOSS, TOL_FLIP_GRAD = TOL_LOGP, TOL_LOSS, TOL_GRAD


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--config", default="", help="run only this config (C0..C5) as the top-level workload")
    p.add_argument("--batch", type=int, default=0, help="override the per-GPU batch (debug)")
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--kernel-path", type=int, default=0)
    p.add_argument("--compute-mode", default="matrix", choices=["matrix", "vector"],
                   help="vector = per-sample step control (row N4; TestMode configs on narrow shapes)")
    p.add_argument("--mode", default="", choices=["", "test", "hutchinson"],
                   help="override the config's headline mode: log-density inference in TestMode (exact trace) "
                        "or in TrainMode (Hutchinson estimate)")
    p.add_argument("--strong", action="store_true", help="with --config: the config's batch is the GLOBAL batch")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-parity", action="store_true")
    p.add_argument("--no-extra", action="store_true", help="default run without the train / strong-scaling blocks")
    p.add_argument("--extras", default="train,cond_strong,train_strong",
                   help="which extra blocks the default run carries (comma list; debugging aid)")
    p.add_argument("--cpu-sample", type=int, default=0)
    p.add_argument("--tc-dbg", type=int, default=0, help="debug / timing-experiment bits of the tensor-core kernels (option tc_dbg)")
    p.add_argument("--stage-sample", type=int, default=13, help="time every N-th stage-kernel launch with CUDA events (1 = all)")
    return p.parse_args()


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.nvml, self.t, self.stop_flag = None, None, False

    def start(self):
        # in-process NVML (one cheap library call per sample) when the binding is there: eight ranks that each run
        # `nvidia-smi -lms 50` put 160 driver queries per second beside the timed host<->device copies
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                hnd = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
            except Exception:
                hnd = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.nvml = (pynvml, hnd)
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _poll(self):
        nv, hnd = self.nvml
        bits = [("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)]
        try:
            mx = nv.nvmlDeviceGetMaxClockInfo(hnd, nv.NVML_CLOCK_SM)
        except Exception:
            mx = 0
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(hnd, nv.NVML_CLOCK_SM)
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(hnd)
                self.rows.append([str(sm), str(mx), "0"] + ["Active" if mask & b else "Not Active" for _, b in bits])
            except Exception:
                pass
            time.sleep(0.025)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.nvml:
            self.stop_flag = True
            self.t.join(timeout=2)
        elif not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            self.proc.terminate()
            self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                pass
        busy = [x for x in sm if x > 0.5 * max(mx)] if mx else sm
        return {"sm_mhz": float(np.median(busy or sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "source": "NVML in-process, 25 ms" if self.nvml else "nvidia-smi -lms 200"}


def make_model(cnf, cid, device, kernel_path, vector=False):
    c = cnf.workloads.CONFIGS[cid]
    dims = c["dims"]
    nn = cnf.Chain(*[cnf.Dense(dims[i], dims[i + 1], "tanh") for i in range(len(dims) - 1)])
    cm = cnf.DIJacVecMatrixMode() if c["trace_mode"] == 1 else cnf.DIVecJacMatrixMode()
    if vector:
        cm = cnf.DIJacVecVectorMode() if c["trace_mode"] == 1 else cnf.DIVecJacVectorMode()
    tag = {"RNODE": cnf.RNODE, "ICNF": cnf.FFJORD, "CondRNODE": cnf.CondRNODE}[c["tag"]]
    sk = {k: c[k] for k in ("reltol", "abstol") if c[k]}
    return cnf.construct(tag, nn, c["nvars"], c["naug"], compute_mode=cm, tspan=c["tspan"],
                         sol_kwargs=sk, lambda1=c["lambda1"], lambda2=c["lambda2"],
                         lambda3=c["lambda3"], resource=cnf.B200Libs(device, kernel_path),
                         rng=np.random.default_rng(0))


def oracle_config(O, c):
    return O.make_config(c["nvars"], c["naug"], c["dims"], C_=c["C"], trace_mode=c["trace_mode"],
                         lambda1=c["lambda1"], lambda2=c["lambda2"], lambda3=c["lambda3"],
                         reltol=c["reltol"], abstol=c["abstol"])


def oracle_step(O, cfg, c, inp, kind):
    """one pass of the CPU path: kind 'train' = loss + gradient, 'hutchinson' = TrainMode inference, else TestMode"""
    if kind == "train":
        return O.loss_grad(cfg, inp["theta"], inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    if kind == "hutchinson":
        return O.logdensity(cfg, inp["theta"], 1, inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    return O.logdensity(cfg, inp["theta"], 0, inp["xs"], inp["ys"], None, *c["tspan"])


def cpu_sample_size(cid, kind, override, cores):
    """bounded CPU sample: C0-C2 at the full batch, the wide training configs on a slice sized for
    ~10-20 s of CPU work on `cores` threads"""
    if override:
        return override
    scale = max(1, cores // 8)
    full = {"C0": 64, "C1": 1024, "C2": 1 << 20}
    if cid in full:
        return full[cid]
    if kind == "train":
        return {"C3": 1 << 10, "C4": 1 << 12, "C5": 32}[cid] * scale
    return {"C3": 1 << 12, "C4": 1 << 14, "C5": 128}[cid] * scale


def parity_slice(cid, kind):
    """samples of the in-run parity check where the oracle cannot afford the benchmarked size"""
    if kind == "train":
        return {"C0": 64, "C1": 1024, "C2": 1 << 12, "C3": 1 << 11, "C4": 1 << 12, "C5": 64}[cid]
    return {"C0": 64, "C1": 1024, "C2": 1 << 20, "C3": 1 << 12, "C4": 1 << 14, "C5": 128}[cid]


def relerr(a, b):
    a, b = np.asarray(a, np.float64).reshape(-1), np.asarray(b, np.float64).reshape(-1)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))


def relnorm(a, b):
    a, b = np.asarray(a, np.float64).reshape(-1), np.asarray(b, np.float64).reshape(-1)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return {}


def tensor_peak(peaks, clocks):
    """SURVEY.md 8.4 / the judge's rule: the burst figure when the sampled SM clock sat at max during the timed
    region, the sustained one otherwise"""
    burst = float(peaks.get("bf16_tflops", 1590.0))
    sust = float(peaks.get("bf16_tflops_sustained", 1400.0))
    src = "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"
    sm, mx = clocks.get("sm_mhz"), clocks.get("sm_max_mhz")
    if sm and mx and sm >= 0.98 * mx:
        return burst, f"{src} bf16_tflops (burst: SM clock {sm:.0f} of {mx:.0f} MHz during the timed region)"
    return sust, f"{src} bf16_tflops_sustained (SM clock {sm} of {mx} MHz during the timed region)"


# --------------------------------------------------------------------------------------------------- reference arm
def run_reference(args, cnf):
    """--impl reference: the reference's CPU implementation of the path.  Julia is absent, so this is the CPU
    oracle (a port), on ALL host threads (torchrun pins OMP_NUM_THREADS=1: reset explicitly), each step a bounded
    sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as O
    cores = O.set_num_threads()
    wl = cnf.workloads

    def leg(cid, kind, steps, warmup):
        c = wl.CONFIGS[cid]
        Bs = cpu_sample_size(cid, kind, args.cpu_sample, cores)
        inp = wl.make_inputs(cid, B=Bs)
        cfg = oracle_config(O, c)
        for _ in range(warmup):
            oracle_step(O, cfg, c, inp, kind)
        t0 = time.perf_counter()
        for _ in range(steps):
            st = oracle_step(O, cfg, c, inp, kind)[3]
        dt = time.perf_counter() - t0
        val = Bs * steps / dt
        metric, unit = METRIC["train" if kind == "train" else "test"]
        full = Bs == c["B"]
        return {
            "metric": metric, "value": val, "unit": unit, "steps": steps, "warmup": warmup,
            "ms_per_step": 1e3 * dt / steps,
            "config": {"workload": f"{cid}: {c['tag']} dims={c['dims']} B={c['B']} mode={kind}",
                       "sample_per_step": Bs, "naccept": st["naccept"], "nreject": st["nreject"]},
            "cpu_baseline": {"value": val, "unit": unit, "cores": cores, "kind": "port",
                             "sample": (f"the full {Bs}-sample batch per step" if full else
                                        f"{Bs} samples per step of the {c['B']}-sample workload (CPU throughput is "
                                        "linear in the batch)")},
            "e2e": {"value": val, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }

    cid = args.config or "C2"
    c = wl.CONFIGS[cid]
    kind = "hutchinson" if args.mode == "hutchinson" else ("test" if args.mode == "test" else c["mode"])
    top = leg(cid, kind, args.steps, args.warmup)
    line = {"impl": "reference", "n_gpus": args.gpus, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "gpu_launches": 0}
    line.update(top)
    if not args.config and not args.no_extra:
        line["train"] = leg("C3", "train", min(args.steps, 3), 1)
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------------- one workload
class Ctx:
    pass


def run_leg(ctx, cid, kind, B, steps, warmup, strong=False, want_cpu=True, want_parity=True):
    """times one workload on every rank.  kind: 'test' (TestMode inference), 'hutchinson' (TrainMode inference),
    'train' (loss + gradient).  B = per-GPU batch.  Returns the block (rank 0) or None."""
    import torch
    import torch.distributed as dist
    cnf, args, wl = ctx.cnf, ctx.args, ctx.cnf.workloads
    rank, world, local, dev = ctx.rank, ctx.world, ctx.local, ctx.dev
    c = wl.CONFIGS[cid]
    train = kind == "train"
    inp = wl.make_inputs(cid, B=B, offset=rank)
    icnf = make_model(cnf, cid, local, args.kernel_path, args.compute_mode == "vector")
    if world > 1:
        cnf.parallel.init_comm(icnf, rank, world)
    # every N-th stage launch of the timed region is bracketed by CUDA events on the library's stream (N = 13 walks
    # through all stage indices of the 6-stage attempts): timing every launch costs two event records per launch,
    # several per cent of a C2 step (and a bracketed launch loses its programmatic overlap with its neighbours)
    cnf.set_option(icnf, "time_stages", max(1, args.stage_sample))
    if args.tc_dbg:
        cnf.set_option(icnf, "tc_dbg", args.tc_dbg)
    mode = cnf.TrainMode() if kind in ("train", "hutchinson") else cnf.TestMode()
    xs_d = torch.from_numpy(inp["xs"]).to(dev).t()
    ys_d = torch.from_numpy(inp["ys"]).to(dev).t() if inp["ys"] is not None else None
    eps_d = torch.from_numpy(inp["eps"]).to(dev).t()
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy().T
    xs_h = pin(inp["xs"])
    ys_h = pin(inp["ys"]) if inp["ys"] is not None else None
    eps_h = pin(inp["eps"])
    theta = inp["theta"]
    # training: a real fit! step uploads a new theta every step (cnfb200_set_params: H2D, transposes, hi/lo planes);
    # two vectors that differ in the last bit of one weight defeat the host mirror's unchanged-theta shortcut
    theta_b = theta.copy()
    theta_b[0] = np.nextafter(theta_b[0], np.float32(1.0))
    thetas = (theta, theta_b)
    nstep = [0]

    def step(xs, ys, eps):
        th = thetas[nstep[0] & 1] if train else theta
        nstep[0] += 1
        tail = ((ys,) if ys is not None else ()) + (th, {})
        if train:
            out = cnf.loss_and_grad(icnf, mode, xs, *tail, eps=eps)
        else:
            out = cnf.inference(icnf, mode, xs, *tail, **({"eps": eps} if kind == "hutchinson" else {}))
        return out, cnf.last_stats(icnf)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    warm_done = 0
    for _ in range(warmup):
        _, st_w = step(xs_d, ys_d, eps_d)
        warm_done += 1
        # a step of several seconds (C5 at the full batch on one GPU) is warm after one call: every workspace, the
        # step tape and the weight planes exist; all ranks see the same device time class, so they stop together
        slow = torch.tensor([1.0 if st_w["ms_total"] > 5000.0 else 0.0], device=dev)
        if world > 1:
            dist.all_reduce(slow, op=dist.ReduceOp.MAX)
        if float(slow[0]) > 0.0:
            break
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ms_dev = ms_stage = ms_solve = 0.0
    n_stage = launches = 0
    t_wall = time.perf_counter()
    for _ in range(steps):
        ctx.flush.zero_()
        torch.cuda.synchronize()
        _, st = step(xs_d, ys_d, eps_d)
        ms_dev += st["ms_total"]
        ms_stage += st["ms_stage"]
        ms_solve += st["ms_solve"]
        n_stage += st["n_stage"]
        launches += st["gpu_launches"]
    barrier()
    wall = time.perf_counter() - t_wall
    # end-to-end: same call, HOST buffers; H2D of the inputs and D2H of the result inside the call
    # (already warm: this only primes the pinned output pool; one call when a step takes seconds)
    # (the result of the previous call stays referenced while the next one runs, exactly as in the timed loop below:
    # the pool of page-locked result buffers reaches its steady state here, not inside the timed region, where a
    # fresh cudaHostAlloc of 4 MB costs milliseconds)
    out_h = st_h = None
    for _ in range(1 if ms_dev / steps > 2000.0 else max(min(warmup, 3), 2)):
        out_h, st_h = step(xs_h, ys_h, eps_h)
    barrier()
    t_e2e = time.perf_counter()
    e2e_each, e2e_dev = [], 0.0
    for _ in range(steps):
        t_one = time.perf_counter()
        out_h, st_h = step(xs_h, ys_h, eps_h)
        e2e_each.append(time.perf_counter() - t_one)
        e2e_dev += st_h["ms_total"]
    barrier()
    e2e_s = time.perf_counter() - t_e2e
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms_dev, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, e2e_s = float(t[0]), float(t[1])

    # ---- parity of what was just timed (rank 0 checks; multi-rank test legs gather every shard's log-density)
    parity = None
    if want_parity and not args.no_parity:
        parity = check_parity(ctx, cid, kind, B, inp, out_h, st_h)
    if rank != 0:
        del icnf
        torch.cuda.empty_cache()
        return None

    metric, unit = METRIC["train" if train else "test"]
    total = B * world * steps
    value = total / (ms_dev * 1e-3)
    D = c["nvars"] + c["naug"]
    nb = lambda a: a.nbytes if a is not None else 0
    h2d = nb(inp["xs"]) + nb(inp["ys"]) + (nb(inp["eps"]) if kind != "test" else 0) + (theta.nbytes if train else 0)
    d2h = (theta.nbytes + 4) if train else (B * 4 + B * 4 * (3 if kind == "hutchinson" else 1))
    roofline = make_roofline(ctx, cid, kind, B, st, ms_stage, n_stage, ms_solve, ms_dev, steps, clocks)
    par = f"dp{world} (columns sharded)"
    if world > 1:
        par += f", error-norm exchange: {cnf.parallel.errnorm_exchange(icnf)}"
        if train:
            par += ", gradient + loss allreduce (NCCL) after the reverse sweep"
    block = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": steps, "warmup": warm_done,
        "ms_per_step": ms_dev / steps, "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{cid}: {c['tag']} dims={c['dims']} B={B}/GPU mode={kind} "
                               f"trace={'JVP' if c['trace_mode'] else 'VJP'} tspan={c['tspan']}",
                   "global_batch": B * world, "parallelism": par,
                   "l2": "256 MB flush between timed steps", "naccept": st["naccept"],
                   "nreject": st["nreject"], "nf": st["nf"], "kernel_path": st["kernel_path"],
                   "timing": "sum of per-call CUDA-event times on the library stream, max over ranks",
                   "wall_ms_per_step_incl_flush": 1e3 * wall / steps},
        "clocks": clocks,
        "e2e": {"value": total / e2e_s, "unit": unit, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * e2e_s / steps,
                "ms_per_step_min": 1e3 * min(e2e_each), "ms_per_step_max": 1e3 * max(e2e_each),
                "device_ms_per_step": e2e_dev / steps},
        "gpu_launches": launches,
        "roofline": roofline,
    }
    if train:
        block["config"]["theta_upload"] = "every step (cnfb200_set_params inside the timed call: a different host theta each step)"
    if parity is not None:
        block["parity"] = parity
    if want_cpu and not args.no_cpu_baseline and world == 1:
        block["cpu_baseline"] = cpu_baseline(ctx, cid, kind, B)
    del icnf, xs_d, ys_d, eps_d
    torch.cuda.empty_cache()
    return block


def cpu_baseline(ctx, cid, kind, B):
    from oracle import oracle as O
    wl = ctx.cnf.workloads
    c = wl.CONFIGS[cid]
    cores = O.set_num_threads()
    Bs = min(cpu_sample_size(cid, kind, ctx.args.cpu_sample, cores), B)
    sinp = wl.make_inputs(cid, B=Bs)
    cfg = oracle_config(O, c)
    oracle_step(O, cfg, c, sinp, kind)
    t0 = time.perf_counter()
    reps = 0
    while reps < 3 and time.perf_counter() - t0 < 20:
        oracle_step(O, cfg, c, sinp, kind)
        reps += 1
    dt = (time.perf_counter() - t0) / reps
    unit = METRIC["train" if kind == "train" else "test"][1]
    return {"value": Bs / dt, "unit": unit, "cores": cores, "kind": "port",
            "sample": f"{Bs} of {B} samples, {reps} repetitions, same synthetic stream"}


def check_parity(ctx, cid, kind, B, inp, out_h, st_h):
    """the results of the timed e2e call (host buffers) against the CPU oracle on the same inputs.
    Test legs: at the benchmarked size when the oracle affords it (every rank's shard, gathered: the oracle solves
    the GLOBAL batch, whose whole-matrix error norm couples the shards); otherwise, and for training legs, a fixed
    slice solved by both sides as its own batch (single GPU, rank 0)."""
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    cnf, wl = ctx.cnf, ctx.cnf.workloads
    rank, world = ctx.rank, ctx.world
    c = wl.CONFIGS[cid]
    Bp = parity_slice(cid, kind)
    full = (kind != "train") and Bp >= B and B * world <= (1 << 23)
    if full:
        lp = np.ascontiguousarray(np.asarray(out_h[0], dtype=np.float32).reshape(-1))
        if world > 1:
            t = torch.from_numpy(lp).to(ctx.dev)
            parts = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
            dist.gather(t, parts, dst=0)
            if rank != 0:
                return None
            lp = torch.cat(parts).cpu().numpy()
        cores = O.set_num_threads()
        shards = [inp] + [wl.make_inputs(cid, B=B, offset=r) for r in range(1, world)]
        cat = lambda k: None if shards[0][k] is None else np.concatenate([s[k] for s in shards])
        ginp = dict(xs=cat("xs"), ys=cat("ys"), eps=cat("eps"), theta=inp["theta"])
        t0 = time.perf_counter()
        rlp, _, _, rst = oracle_step(O, oracle_config(O, c), c, ginp, kind)
        dt = time.perf_counter() - t0
        ds = int(st_h["naccept"] - rst["naccept"])
        err = relerr(lp, rlp)
        tol = TOL_LOGP if ds == 0 else TOL_FLIP_LOGP
        return {"max_rel_logp": err, "dsteps": ds, "naccept": int(st_h["naccept"]), "naccept_oracle": int(rst["naccept"]),
                "B_checked": int(B * world), "what": "every sample of the timed call (all ranks) vs the oracle's solve of "
                "the global batch", "tol_logp": tol, "ok": bool(abs(ds) <= 1 and err < tol), "oracle_s": dt, "oracle_cores": cores}
    if rank != 0:
        return None
    cores = O.set_num_threads()
    Bp = min(Bp, B)
    sinp = wl.make_inputs(cid, B=Bp)
    cfg = oracle_config(O, c)
    one = make_model(cnf, cid, ctx.local, ctx.args.kernel_path, ctx.args.compute_mode == "vector")   # no communicator
    mode = cnf.TrainMode() if kind in ("train", "hutchinson") else cnf.TestMode()
    tail = ((sinp["ys"].T,) if sinp["ys"] is not None else ()) + (sinp["theta"], {})
    t0 = time.perf_counter()
    if kind == "train":
        val, g = cnf.loss_and_grad(one, mode, sinp["xs"].T, *tail, eps=sinp["eps"].T)
        st = cnf.last_stats(one)
        rval, rg, _, rst = oracle_step(O, cfg, c, sinp, kind)
        ds = int(st["naccept"] - rst["naccept"])
        le, ge = abs(val - rval) / abs(rval), relnorm(g, rg)
        tl, tg = (TOL_LOSS, TOL_GRAD) if ds == 0 else (TOL_FLIP_LOSS, TOL_FLIP_GRAD)
        out = {"loss_rel": le, "grad_rel": ge, "dsteps": ds, "naccept": int(st["naccept"]),
               "naccept_oracle": int(rst["naccept"]), "B_checked": int(Bp), "tol_loss": tl, "tol_grad": tg,
               "ok": bool(abs(ds) <= 1 and le < tl and ge < tg), "kernel_path": st["kernel_path"]}
    else:
        lp, _ = cnf.inference(one, mode, sinp["xs"].T, *tail, **({"eps": sinp["eps"].T} if kind == "hutchinson" else {}))
        st = cnf.last_stats(one)
        rlp, _, _, rst = oracle_step(O, cfg, c, sinp, kind)
        ds = int(st["naccept"] - rst["naccept"])
        err = relerr(lp, rlp)
        tol = TOL_LOGP if ds == 0 else TOL_FLIP_LOGP
        out = {"max_rel_logp": err, "dsteps": ds, "naccept": int(st["naccept"]), "naccept_oracle": int(rst["naccept"]),
               "B_checked": int(Bp), "tol_logp": tol, "ok": bool(abs(ds) <= 1 and err < tol), "kernel_path": st["kernel_path"]}
    out["what"] = (f"a fixed {Bp}-sample slice of the workload solved as its own batch by the library (1 GPU) and by "
                   "the oracle: the oracle cannot afford the benchmarked size")
    out["oracle_s"] = time.perf_counter() - t0
    out["oracle_cores"] = cores
    del one
    return out


def make_roofline(ctx, cid, kind, B, st, ms_stage, n_stage, ms_solve, ms_dev, steps, clocks):
    """achieved = ALGORITHMIC flops of one augmented_f evaluation of B samples (F_E or F_H per sample;
    operand splitting, recomputation and padding are overhead that lowers the fraction) / the mean CUDA-event time of
    the stage kernel(s) of one evaluation / the peak of the pipe the kernel runs on."""
    import ctypes
    cnf, wl = ctx.cnf, ctx.cnf.workloads
    c = wl.CONFIGS[cid]
    D = c["nvars"] + c["naug"]
    R = D + (1 if kind == "test" else 3)
    fmode = "test" if kind == "test" else "train"
    flops_eval = wl.flops_per_eval(cid, fmode) * B
    stage_ms = ms_stage / max(n_stage, 1)
    alg = flops_eval / (stage_ms * 1e-3) / 1e12 if stage_ms > 0 else 0.0
    peaks = load_peaks()
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{cid}:path{st['kernel_path']}")
    except (OSError, ValueError):
        pass
    path = st["kernel_path"]
    common = {"traffic": traffic, "traffic_source": "static: dram__bytes_read+write per launch from the committed ncu --set full "
              "capture of this kernel (profiles/traffic.json), not measured in this run" if traffic else None,
              "stage_ms": stage_ms, "stage_launches_timed": n_stage, "stage_sampling": f"every {max(1, ctx.args.stage_sample)}-th stage launch of the timed steps",
              "kernel_path": path,
              "stage_share_of_step": stage_ms * st["nf"] * steps / ms_dev if ms_dev else None,
              "algorithmic_flops_per_eval": flops_eval,
              "algorithmic_bytes_per_eval": (2 * R + D) * 4 * B,
              "algorithmic_gbs": (2 * R + D) * 4 * B / (stage_ms * 1e-3) / 1e9 if stage_ms > 0 else None}
    if path in (3, 6, 7):
        peak, src = tensor_peak(peaks, clocks)
        dims = c["dims"]
        if path == 6:   # only layer 2 of the activation and of the D tangents runs on the tensor pipe
            executed = 3 * 2 * dims[1] * dims[2] * (1 + D) * B / (stage_ms * 1e-3) / 1e12 if stage_ms > 0 else 0.0
            kern = ("k_stage_narrow_tc: RK stage kernel, layer 2 of the activation and the D tangents as tcgen05 MMAs on "
                    "operands the SIMT phase writes into swizzled shared memory")
        else:
            executed = 3 * alg
            kern = ("k_tc_fused: one augmented_f evaluation, all layer products of a 128-sample tile chained on chip "
                    "(tcgen05, A operand in TMEM)" if path == 7 else
                    "one augmented_f evaluation = 2L tcgen05 layer GEMMs (k_tc_gemm) + prep")
        rf = {"bound": "tensor", "achieved": alg, "peak": peak, "unit": "TFLOP/s", "frac": alg / peak if peak else None,
              "kernel": kern, "peak_source": src,
              "note": "achieved = ALGORITHMIC flops (SURVEY.md 8.4); the FP16 hi/lo operand split executes 3 f16 MMA flops "
                      "per algorithmic flop of a tensor-core product (hi*hi + lo*hi + hi*lo), which is overhead here: a "
                      "kernel that kept the tensor pipe 100 % busy would read frac = 1/3",
              "executed_f16_tflops": executed, "executed_frac": executed / peak if peak else None}
    else:
        pk = ctypes.c_double(0.0)
        cnf._cabi.load().cnfb200_measure_fp32_peak(ctx.local, ctypes.byref(pk))
        achieved = alg
        extra = {}
        if path == 4:   # vector mode: one launch = every sample's whole solve; a warp runs as long as its slowest lane
            solve_ms = ms_solve / steps
            achieved = flops_eval * st["nf"] / (solve_ms * 1e-3) / 1e12 if solve_ms > 0 else 0.0
            extra = {"note": "achieved counts lane-executed evaluations (nf of the slowest sample for every lane)",
                     "stage_ms": solve_ms, "stage_launches": 1}
        rf = {"bound": "fp32_fma", "achieved": achieved, "peak": pk.value, "unit": "TFLOP/s",
              "frac": achieved / pk.value if pk.value else None,
              "kernel": "RK stage kernel (augmented_f over the whole batch)",
              "peak_source": "measured in this run by cnfb200_measure_fp32_peak (dependent-FFMA microbenchmark); "
                             "MEASURED_PEAKS.json has no FP32 figure"}
        rf.update(extra)
    for k, v in common.items():
        rf.setdefault(k, v)
    if kind == "train":
        # the whole training call: (6 nattempt + 3) F_H + 6 naccept F_B per sample
        natt = st["naccept"] + st["nreject"]
        fl = ((6 * natt + 3) * wl.flops_per_eval(cid, "train") + 6 * st["naccept"] * wl.flops_backward_per_eval(cid)) * B
        peak, src = tensor_peak(peaks, clocks)
        t = ms_dev / steps * 1e-3
        rf["step"] = {"algorithmic_flops": fl, "achieved": fl / t / 1e12 if t > 0 else None, "unit": "TFLOP/s",
                      "peak": peak, "frac": fl / t / 1e12 / peak if t > 0 and peak else None,
                      "note": "whole loss+gradient call: forward solve + reverse sweep + gradient allreduce"}
    return rf


# --------------------------------------------------------------------------------------------------- main
def main():
    args = parse()
    import cnfb200_loader
    cnf = cnfb200_loader.load_package()
    if args.impl == "reference":
        return run_reference(args, cnf)
    import torch
    import torch.distributed as dist
    ctx = Ctx()
    ctx.cnf, ctx.args = cnf, args
    ctx.rank = int(os.environ.get("RANK", "0"))
    ctx.local = int(os.environ.get("LOCAL_RANK", "0"))
    ctx.world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(ctx.local)
    if ctx.world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", ctx.local))
    ctx.dev = torch.device("cuda", ctx.local)
    ctx.flush = torch.empty(256 << 20, dtype=torch.uint8, device=ctx.dev)  # > 126 MB L2
    wl = cnf.workloads
    world = ctx.world

    cid = args.config or "C2"
    c = wl.CONFIGS[cid]
    kind = "hutchinson" if args.mode == "hutchinson" else ("test" if args.mode == "test" else c["mode"])
    B = args.batch or c["B"]
    strong = bool(args.config and args.strong)
    if strong:
        B = max(B // world, 1)
    line = run_leg(ctx, cid, kind, B, args.steps, args.warmup, strong=strong)
    ok = True
    if not args.config and not args.no_extra:
        extras = [
            # name, config, kind, per-GPU batch, steps, warmup, strong
            ("train", "C3", "train", wl.CONFIGS["C3"]["B"], min(args.steps, 10), max(3, min(args.warmup, 3)), False),
            ("cond_strong", "C4", "train", max(wl.CONFIGS["C4"]["B"] // world, 1), min(args.steps, 5), 3, True),
            ("train_strong", "C5", "train", max(wl.CONFIGS["C5"]["B"] // world, 1), min(args.steps, 2 if world < 4 else 5), 3, True),
        ]
        for name, xc, xk, xB, xs, xw, xstrong in extras:
            if name not in args.extras.split(","):
                continue
            try:
                blk = run_leg(ctx, xc, xk, xB, xs, xw, strong=xstrong, want_cpu=(name == "train"),
                              want_parity=True)
            except Exception as e:   # an extra block must not take the headline record down with it
                import traceback
                traceback.print_exc(file=sys.stderr)
                blk = {"error": f"{type(e).__name__}: {e}"}
                ok = False
            if ctx.rank == 0:
                line[name] = blk
    if ctx.rank == 0:
        for blk in [line] + [line.get(k) for k in ("train", "cond_strong", "train_strong")]:
            if isinstance(blk, dict) and isinstance(blk.get("parity"), dict) and not blk["parity"].get("ok", True):
                ok = False
        line["parity_ok"] = ok
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0 if ok else 3


_REAL_STDOUT = None


def _own_stdout():
    """the contract is ONE JSON line on stdout: everything native libraries print there (e.g. NCCL's
    version banner under NCCL_DEBUG) is sent to stderr, the line itself goes to the saved descriptor"""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


if __name__ == "__main__":
    _own_stdout()
    sys.exit(main())
