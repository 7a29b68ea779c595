#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native augmented-dynamics solve.

  python bench.py --gpus N --steps K --warmup W [--config C2] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic input: one
`inference(icnf, TestMode(), xs, ps, st)` call (log-density configs) or one loss+gradient call
(training configs) on the config's batch.  N=1 workload: C2 = BASELINE.json configs[1]
(ICNF exact trace, DIJacVecMatrixMode, MLP 2->64->64->2, B = 2^20).  N>1: weak scaling, every
rank solves its own shard of B columns; the step-size controller sees the whole batch through a
per-step NCCL allreduce of the error-norm partials.

Prints ONE JSON line (rank 0).  `value` = device-resident inputs, CUDA-event time of the calls
(max over ranks); `e2e` = the same call with HOST buffers (pinned), host<->device copies inside the
timed region; `roofline` = the RK stage kernel against the measured FP32-FMA peak; `cpu_baseline`
= the CPU oracle (port of the reference's CPU path) on a bounded sample on this box's cores.
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


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--config", default="C2")
    p.add_argument("--batch", type=int, default=0, help="override the per-GPU batch (debug)")
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--kernel-path", type=int, default=0)
    p.add_argument("--compute-mode", default="matrix", choices=["matrix", "vector"],
                   help="vector = per-sample step control (row N4; TestMode configs on narrow shapes)")
    p.add_argument("--mode", default="", choices=["", "test", "hutchinson"],
                   help="override the config's headline mode: log-density inference in TestMode (exact trace) "
                        "or in TrainMode (Hutchinson estimate)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--cpu-sample", type=int, default=0)
    return p.parse_args()


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
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
                "samples": len(sm)}


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


def oracle_step(O, cfg, c, inp, train):
    if train:
        _, _, _, st = O.loss_grad(cfg, inp["theta"], inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    else:
        _, _, _, st = O.logdensity(cfg, inp["theta"], 0, inp["xs"], inp["ys"], None, *c["tspan"])
    return st


def oracle_config(O, c):
    return O.make_config(c["nvars"], c["naug"], c["dims"], C_=c["C"], trace_mode=c["trace_mode"],
                         lambda1=c["lambda1"], lambda2=c["lambda2"], lambda3=c["lambda3"],
                         reltol=c["reltol"], abstol=c["abstol"])


def cpu_sample_size(cid, override):
    if override:
        return override
    return {"C0": 64, "C1": 1024, "C2": 1 << 16, "C3": 1 << 11, "C4": 1 << 14, "C5": 1 << 7}[cid]


def run_reference(args, cnf):
    """--impl reference: the reference's CPU implementation of the path.  Julia is absent, so this
    is the CPU oracle (a port), all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    wl = cnf.workloads
    cid = args.config
    c = wl.CONFIGS[cid]
    train = c["mode"] == "train"
    Bs = cpu_sample_size(cid, args.cpu_sample)
    inp = wl.make_inputs(cid, B=Bs)
    cfg = oracle_config(O, c)
    for _ in range(args.warmup):
        oracle_step(O, cfg, c, inp, train)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st = oracle_step(O, cfg, c, inp, train)
    dt = time.perf_counter() - t0
    val = Bs * args.steps / dt
    metric, unit = METRIC[c["mode"]]
    line = {
        "impl": "reference", "metric": metric, "value": val, "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{cid}: {c['tag']} dims={c['dims']} B={c['B']} mode={c['mode']}",
                   "sample_per_step": Bs, "naccept": st["naccept"], "nreject": st["nreject"]},
        "cpu_baseline": {"value": val, "unit": unit, "cores": O.num_threads(), "kind": "port",
                         "sample": f"{Bs} samples per step of the {c['B']}-sample workload "
                                   "(CPU throughput is linear in the batch)"},
        "e2e": {"value": val, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def main():
    args = parse()
    import cnfb200_loader
    cnf = cnfb200_loader.load_package()
    if args.impl == "reference":
        return run_reference(args, cnf)
    import torch
    import torch.distributed as dist
    wl = cnf.workloads
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cid = args.config
    c = wl.CONFIGS[cid]
    if args.mode:   # inference on a training config (e.g. C4: "Test (inference) and Train")
        c = dict(c, mode="test")
    train = c["mode"] == "train"
    B = args.batch or c["B"]
    D = c["nvars"] + c["naug"]
    inp = wl.make_inputs(cid, B=B, offset=rank)
    icnf = make_model(cnf, cid, local, args.kernel_path, args.compute_mode == "vector")
    if world > 1:
        cnf.parallel.init_comm(icnf, rank, world)
    cnf.set_option(icnf, "time_stages", 1)
    mode = cnf.TrainMode() if (train or args.mode == "hutchinson") else cnf.TestMode()
    dev = torch.device("cuda", local)
    # device-resident inputs (value) and pinned host inputs (e2e); `rows x B` views, Julia-style
    xs_d = torch.from_numpy(inp["xs"]).to(dev).t()
    ys_d = torch.from_numpy(inp["ys"]).to(dev).t() if inp["ys"] is not None else None
    eps_d = torch.from_numpy(inp["eps"]).to(dev).t()
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy().T
    xs_h = pin(inp["xs"])
    ys_h = pin(inp["ys"]) if inp["ys"] is not None else None
    eps_h = pin(inp["eps"])
    theta = inp["theta"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step(xs, ys, eps):
        tail = ((ys,) if ys is not None else ()) + (theta, {})
        if train:
            out = cnf.loss_and_grad(icnf, mode, xs, *tail, eps=eps)
        else:
            out = cnf.inference(icnf, mode, xs, *tail, **({"eps": eps} if args.mode == "hutchinson" else {}))
        return out, cnf.last_stats(icnf)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(xs_d, ys_d, eps_d)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ms_dev = ms_stage = ms_solve = 0.0
    n_stage = launches = 0
    t_wall = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()
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
    for _ in range(min(args.warmup, 2)):
        step(xs_h, ys_h, eps_h)
    barrier()
    t_e2e = time.perf_counter()
    for _ in range(args.steps):
        out_h, st_h = step(xs_h, ys_h, eps_h)
    barrier()
    e2e_s = time.perf_counter() - t_e2e
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms_dev, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, e2e_s = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    metric, unit = METRIC[c["mode"]]
    total = B * world * args.steps
    value = total / (ms_dev * 1e-3)
    # (theta is the same host vector in every step: the host mirror uploads it once and skips the copy afterwards)
    h2d = inp["xs"].nbytes + (inp["ys"].nbytes if inp["ys"] is not None else 0) + \
        (inp["eps"].nbytes if train else 0)
    d2h = (theta.nbytes + 4) if train else (B * 4 + B * 4)
    # roofline of the dominant kernel: one RK stage kernel = one augmented_f evaluation of B samples
    flops_eval = wl.flops_per_eval(cid, "train" if args.mode == "hutchinson" else c["mode"]) * B
    lib = cnf._cabi.load()
    import ctypes
    pk = ctypes.c_double(0.0)
    lib.cnfb200_measure_fp32_peak(local, ctypes.byref(pk))
    stage_ms = ms_stage / max(n_stage, 1)
    achieved = flops_eval / (stage_ms * 1e-3) / 1e12 if stage_ms > 0 else 0.0
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{cid}:path{st['kernel_path']}")
    except (OSError, ValueError):
        pass
    if st["kernel_path"] == 4:
        # vector mode: the whole adaptive solve of every sample is ONE kernel; a warp runs as long as its slowest lane,
        # so the executed work is nf_max evaluations per lane
        solve_ms = ms_solve / args.steps
        achieved = flops_eval * st["nf"] / (solve_ms * 1e-3) / 1e12 if solve_ms > 0 else 0.0
        roofline = {"bound": "fp32_fma", "achieved": achieved, "peak": pk.value, "unit": "TFLOP/s",
                    "frac": achieved / pk.value if pk.value else None, "traffic": traffic,
                    "kernel": "k_solve_vector: one thread = one sample's whole Tsit5 solve in registers",
                    "peak_source": "measured in this run by cnfb200_measure_fp32_peak",
                    "note": "achieved counts lane-executed evaluations (nf of the slowest sample for every lane: lanes that "
                            "finish early idle inside the warp), so it is an upper bound on useful flops",
                    "stage_ms": solve_ms, "stage_launches": 1, "kernel_path": 4,
                    "stage_share_of_step": ms_solve / ms_dev if ms_dev else None}
    elif st["kernel_path"] == 6:
        # narrow exact-trace nets with layer 2 (94 % of the flops) on tcgen05: the MMAs execute 3 f16 flops per
        # algorithmic flop of that layer (FP16 hi/lo split); layers 1 and 3 and all elementwise work stay SIMT
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except (OSError, ValueError):
            pass
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        dims = c["dims"]
        D = c["nvars"] + c["naug"]
        mma_flops = 3 * 2 * dims[1] * dims[2] * (1 + D) * B      # executed by tcgen05.mma per evaluation
        executed = mma_flops / (stage_ms * 1e-3) / 1e12 if stage_ms > 0 else 0.0
        roofline = {"bound": "tensor", "achieved": executed, "peak": peak, "unit": "TFLOP/s",
                    "frac": executed / peak, "traffic": traffic,
                    "kernel": "k_stage_narrow_tc: RK stage kernel, layer 2 of the activation and the D tangents as "
                              "tcgen05 MMAs on operands the SIMT phase writes into swizzled shared memory",
                    "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1.4 PFLOP/s sustained",
                    "note": "achieved = f16 MMA flops executed (3 per algorithmic flop of layer 2: hi*hi + lo*hi + hi*lo); "
                            "the kernel is bound by the SIMT prologue/epilogue (tanh, hi/lo split, layers 1 and 3: "
                            "~2.2 k instructions per sample, issue slots 46 % and the MUFU pipe 43 % busy in ncu), not by the tensor pipe; "
                            "algorithmic_tflops / fp32_fma_peak shows how far it is above the FP32-FMA roofline that "
                            "bounded the register-resident kernel (kernel_path 2)",
                    "algorithmic_tflops": achieved, "fp32_fma_peak": pk.value,
                    "algorithmic_over_fp32_fma_peak": achieved / pk.value if pk.value else None,
                    "stage_ms": stage_ms, "stage_launches": n_stage, "kernel_path": 6,
                    "stage_share_of_step": ms_stage / ms_dev if ms_dev else None}
    elif st["kernel_path"] == 3:
        # tensor path: FP16 hi/lo operand splitting executes 3 f16 MMA flops per algorithmic flop
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except (OSError, ValueError):
            pass
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        hbm_peak = float(peaks.get("hbm_gbs", 6400.0))
        bytes_eval = (wl.tc_bytes_per_eval_exact(cid) if (c["mode"] == "test" and args.mode != "hutchinson") else wl.tc_bytes_per_eval(cid)) * B
        t_hbm = bytes_eval / (hbm_peak * 1e9)
        t_tensor = 3 * flops_eval / (peak * 1e12)
        if t_hbm > t_tensor:
            # narrow hidden widths: the layer GEMMs stream hi/lo planes through HBM faster than they fill the tensor pipe
            gbs = bytes_eval / (stage_ms * 1e-3) / 1e9 if stage_ms > 0 else 0.0
            roofline = {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                        "traffic": traffic,
                        "kernel": "one augmented_f evaluation = 2L tcgen05 layer GEMMs (k_tc_gemm) + prep",
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy bandwidth)" if peaks else "fallback 6.4 TB/s",
                        "note": "achieved = algorithmic bytes of the evaluation (workloads.tc_bytes_per_eval: operands and "
                                "results of every layer GEMM as hi/lo 16-bit planes) / mean evaluation time; the tensor-pipe "
                                "time of the same evaluation at the sustained bf16 peak would be "
                                f"{1e3 * t_tensor:.3f} ms against {1e3 * t_hbm:.3f} ms of HBM time",
                        "algorithmic_tflops": achieved, "executed_f16_tflops": 3 * achieved,
                        "stage_ms": stage_ms, "stage_launches": n_stage, "kernel_path": 3,
                        "stage_share_of_step": ms_stage / ms_dev if ms_dev else None}
        else:
          roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                    "frac": achieved / peak, "traffic": traffic,
                    "kernel": "one augmented_f evaluation = 2L tcgen05 layer GEMMs (k_tc_gemm) + prep",
                    "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (dense bf16, stage timed inside a "
                                   "long step)" if peaks else "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)",
                    "note": "achieved counts ALGORITHMIC flops (F_H per sample); the kernels execute 3x that in "
                            "f16 MMA flops (FP16 hi/lo split, hi*hi + lo*hi + hi*lo, for FP32-level accuracy)",
                    "executed_f16_tflops": 3 * achieved, "stage_ms": stage_ms, "stage_launches": n_stage,
                    "kernel_path": 3, "stage_share_of_step": ms_stage / ms_dev if ms_dev else None}
    else:
        roofline = {"bound": "fp32_fma", "achieved": achieved, "peak": pk.value, "unit": "TFLOP/s",
                    "frac": achieved / pk.value if pk.value else None, "traffic": traffic,
                    "kernel": "RK stage kernel (augmented_f over the whole batch)",
                    "peak_source": "measured in this run by cnfb200_measure_fp32_peak (dependent-FFMA "
                                   "microbenchmark); MEASURED_PEAKS.json has no FP32 figure",
                    "stage_ms": stage_ms, "stage_launches": n_stage, "kernel_path": st["kernel_path"],
                    "stage_share_of_step": ms_stage / ms_dev if ms_dev else None}
    line = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{cid}: {c['tag']} dims={c['dims']} B={B}/GPU mode={c['mode']} "
                               f"trace={'JVP' if c['trace_mode'] else 'VJP'} tspan={c['tspan']}",
                   "global_batch": B * world, "parallelism": f"dp{world} (columns sharded)" + (f", error-norm exchange: {cnf.parallel.errnorm_exchange(icnf)}" if world > 1 else ""),
                   "l2": "256 MB flush between timed steps", "naccept": st["naccept"],
                   "nreject": st["nreject"], "nf": st["nf"], "timing": "sum of per-call CUDA-event "
                   "times on the library stream, max over ranks", "wall_ms_per_step_incl_flush":
                   1e3 * wall / args.steps},
        "clocks": clocks,
        "e2e": {"value": total / e2e_s, "unit": unit, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * e2e_s / args.steps},
        "gpu_launches": launches,
        "roofline": roofline,
    }
    if not args.no_cpu_baseline and world == 1:
        from oracle import oracle as O
        Bs = cpu_sample_size(cid, args.cpu_sample)
        sinp = wl.make_inputs(cid, B=Bs)
        cfg = oracle_config(O, c)
        oracle_step(O, cfg, c, sinp, train)
        t0 = time.perf_counter()
        reps = 0
        while reps < 3 and time.perf_counter() - t0 < 20:
            oracle_step(O, cfg, c, sinp, train)
            reps += 1
        dt = (time.perf_counter() - t0) / reps
        line["cpu_baseline"] = {"value": Bs / dt, "unit": unit, "cores": O.num_threads(), "kind": "port",
                                "sample": f"{Bs} of {B} samples, {reps} repetitions, same synthetic stream"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


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
    main()
