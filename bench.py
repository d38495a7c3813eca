#!/usr/bin/env python
"""Headline benchmark: HES lookahead tree-node posterior evals/s (fwd+bwd) - BASELINE.json metric.

Workload (BASELINE.json configs[1], SURVEY.md 8d cfg 2b): Ackley 2-D, 50 GP training points,
HES-TS-AM with 3 lookahead steps, 64 restarts x 64 Thompson paths = 4096 chains per GPU, FP64,
hidden size 64, euclidean spotlight cost.  One "step" = one AdamW iteration of the reference's
``Actor.query`` loop (actor.py:558-656): rollout forward (+ the no-grad rollout with the frozen
network), objective + best-restart bookkeeping, reverse rollout, weight gradients, AdamW.
A node eval = one query point evaluated under the Thompson path, forward and backward:
4096 chains x (3 design levels + 1 action) = 16,384 per step and GPU (the no-grad rollout's
16,384 forward-only evals are NOT counted).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [--gpus N] ...                  # CPU oracle on the host cores

Multi-GPU (torchrun, one rank per GPU): every rank owns 64 restarts x 64 paths (weak scaling); the
policy gradient is summed and the snapshot decision MIN-reduced over the ranks once per iteration - by this repo's
own kernels over NVLink peer memory, fused with the AdamW update (--exchange peer, the default when the box maps
peer memory), or by two NCCL all-reduces (--exchange nccl).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch

ENV, N_TRAIN, D, H, L = "Ackley", 50, 2, 64, 3
N_RESTARTS, N_PATHS = 64, 64
METRIC = "hes_lookahead_node_posterior_evals_per_s_fwd_bwd"
UNIT = "node evals/s"
FP64_PEAK_FILE = os.path.join(ROOT, "profiles", "pipe_peaks_r01.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-selfcheck", action="store_true", help="N > 1: skip the sharded parity self-check")
    ap.add_argument("--cpu-config", type=int, default=None, choices=[1, 3, 4],
                    help="run ONLY the cpu_baseline leg of another BASELINE config (oracle port on the host cores): "
                         "1 = one Ackley HES-TS-AM-1 BO step's Actor.query loop, 3 = Hartmann HES-MC tree, "
                         "4 = Cosine8 HES-TS-AM (n = 200)")
    ap.add_argument("--sustained-s", type=float, default=2.0,
                    help="seconds of back-to-back graph replays for the sustained-clock number (0 = skip)")
    ap.add_argument("--restarts", type=int, default=N_RESTARTS)
    ap.add_argument("--paths", type=int, default=N_PATHS)
    ap.add_argument("--lookahead", type=int, default=L)
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl"],
                    help="N > 1: per-iteration gradient/decision exchange (own kernels over NVLink peer memory, or NCCL)")
    return ap.parse_args()


# ----------------------------------------------------------------------------- workload
def cost_hypers():
    return dict(radius=0.075, k=1.0, max_noise=1e-5, p_norm=2.0, discount=0.0, discount_threshold=-1.0)


def host_workload(n_restarts, n_paths, rank=0):
    """Host-side (CPU tensors) inputs of one rank: GP data + hypers, initial policy, previous point."""
    from nonmyopia_b200.synthetic import make_problem
    prob = make_problem(ENV, seed=2)
    g = torch.Generator().manual_seed(1000 + rank)
    hidden_noise = torch.randn(n_restarts, H, generator=g).to(torch.float64)
    N = n_restarts * n_paths
    prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
    prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
    prev_h = hidden_noise.repeat(n_paths, 1).contiguous()          # chain = path * n_restarts + restart
    return prob, prev_x, prev_y, prev_h


def init_policy(d, hidden):
    from nonmyopia_b200.amortized_network import AmortizedNetwork
    torch.manual_seed(2)
    return AmortizedNetwork(d + 1, d, hidden, 1, torch.tensor([[0.0, 1.0]] * d, dtype=torch.float64)).double()


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.kill()
        self.t.join(timeout=1)
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower() == "active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ----------------------------------------------------------------------------- CPU oracle leg
def cpu_oracle_rate(n_restarts, lookahead, target_s, threads):
    """Time the CPU oracle (oracle/hes_oracle.py: plain-PyTorch restatement of the reference's
    Actor.query loop - forward, no-grad forward, autograd backward, AdamW, bookkeeping) on a bounded
    sample of the workload: ``n_restarts`` chains (the same chain count as one B200 step when called with
    restarts x paths) of ONE Thompson path - the reference's path object holds a single sample, the
    per-node arithmetic is identical - for as many Adam iterations as fit in ``target_s`` seconds."""
    from oracle import gp_math, hes_oracle as ho
    torch.set_num_threads(threads)
    prob, prev_x, prev_y, prev_h = host_workload(n_restarts, 1)
    ls = torch.full((1, D), prob.lengthscale, dtype=torch.float64)
    gp = gp_math.DenseGP.fit_caches(prob.train_X, prob.train_Y, kind=prob.kernel, lengthscale=ls,
                                    outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
    torch.manual_seed(2)
    path = gp_math.draw_matheron_path(gp, prob.train_Y, 1)
    net = init_policy(D, H)
    spec = ho.HESSpec(lookahead, [1] * lookahead, 1, ho.CostSpec(**cost_hypers()), enable_ts=True)
    budget = 2 * math.sqrt(D)

    def run(n_iter):
        P = {k: v.detach().clone() for k, v in net.state_dict().items()}
        t0 = time.perf_counter()
        ho.query_loop(spec, prev_x, prev_y, prev_h, P, budget=budget, lr=1e-3, n_iter=n_iter, past_initial=False,
                      path=path)
        return time.perf_counter() - t0

    run(2)                                   # warm-up (thread pools, allocator)
    t_cal = run(4) / 4
    n_iter = int(max(4, min(190, target_s / max(t_cal, 1e-6))))      # < 200: the early-stop gate stays shut
    dt = run(n_iter)
    evals = n_restarts * (lookahead + 1) * n_iter
    return dict(value=evals / dt, ms_per_step=dt / n_iter * 1e3, n_iter=n_iter, seconds=dt,
                sample=f"{n_restarts} chains x {lookahead + 1} levels (one Thompson path), {n_iter} Adam iterations "
                       f"({dt:.1f} s) of the same Ackley-2D HES-TS-AM-{lookahead} workload")


def cpu_baseline_config(cfg: int):
    """cpu_baseline leg of BASELINE configs[0], [2] and [3] (SURVEY 8d cfg 1, 3, 4): the oracle port timed on a bounded
    sample of the workload on all host cores; MC trees are extrapolated per node (the dense fantasy models of the
    reference need ~4.5 GB of host memory per restart at n = 500)."""
    from nonmyopia_b200.synthetic import make_problem
    from oracle import gp_math, hes_oracle as ho
    F64 = torch.float64
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if cfg == 1:
        fx = torch.load(os.path.join(ROOT, "tests", "golden", "query_ts_am_ackley_L1_full.pt"), weights_only=False)
        g = fx["gp"]
        d = fx["d"]
        ls = torch.full((1, d), g["lengthscale"], dtype=F64)
        gp = gp_math.DenseGP.fit_caches(g["train_X"], g["train_Y"], kind=g["kernel"], lengthscale=ls,
                                        outputscale=g["outputscale"], noise=g["noise"], mean_const=g["mean_const"])
        pth = gp_math.MatheronPath(fx["path"]["W"], fx["path"]["w"], fx["path"]["v"], g["train_X"], kind=g["kernel"],
                                   lengthscale=ls, outputscale=g["outputscale"], mean_const=g["mean_const"])
        c = fx["cost"]
        spec = ho.HESSpec(fx["L"], [1] * fx["L"], 1, ho.CostSpec(radius=c["radius"], k=c["k"], max_noise=c["max_noise"],
                                                                 p_norm=c["p_norm"]), enable_ts=True)
        torch.set_rng_state(fx["rng_state_in"])
        from nonmyopia_b200.amortized_network import AmortizedNetwork
        for _ in range(2):          # the two throw-away networks of actor.py:524-555 consume the generator
            AmortizedNetwork(d + 1, d, fx["H"], 1, torch.tensor([[0.0, 1.0]] * d, dtype=F64))
        P = {k: v.clone() for k, v in fx["policy"].items()}
        t0 = time.perf_counter()
        o = ho.query_loop(spec, fx["prev_X"], fx["prev_y"], fx["prev_h"], P, budget=fx["budget"], lr=fx["lr"],
                          n_iter=fx["max_iter"], past_initial=False, path=pth)
        dt = time.perf_counter() - t0
        nodes = fx["n_r"] * (fx["L"] + 1) * o["n_iter"]
        ok = int(o["idx"]) == fx["out"]["chosen_idx"]
        return dict(config="configs[0]: Ackley 2D HES-TS-AM-1, seed-pinned, 64 restarts, H=64, early stop",
                    cpu_baseline=dict(value=nodes / dt, unit=UNIT, cores=cores, kind="port",
                                      sample=f"the whole Actor.query loop of the BO step: {o['n_iter']} Adam iterations x 64 "
                                             f"restarts x 2 levels in {dt:.1f} s; chosen restart {int(o['idx'])} "
                                             f"({'=' if ok else '!='} reference {fx['out']['chosen_idx']})"),
                    seconds=dt, n_iter=o["n_iter"], reference_loop_seconds_in_build_container=fx.get("ref_cpu_seconds"))
    if cfg == 3:
        prob = make_problem("Hartmann", seed=2, lengthscale=0.5)
        d, n = prob.d, prob.train_X.shape[0]
        ls = torch.full((1, d), prob.lengthscale, dtype=F64)
        gp = gp_math.DenseGP.fit_caches(prob.train_X, prob.train_Y, kind=prob.kernel, lengthscale=ls,
                                        outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
        nf, L, n_r = [64, 4, 2], 3, 1
        spec = ho.HESSpec(L, nf, 1, ho.CostSpec(radius=prob.radius, k=1, max_noise=1e-5, p_norm=2.0), enable_ts=False)
        gen = torch.Generator().manual_seed(1)
        cnt = [n_r * math.prod(nf[:s]) for s in range(L + 1)]
        maps = [torch.rand(c, d, dtype=F64, generator=gen).requires_grad_(True) for c in cnt]
        base = [gp_math.sobol_normal_base_samples(f, 1, 0).reshape(-1) for f in nf]
        abase = gp_math.sobol_normal_base_samples(64, 1, 1).reshape(-1)
        prev_X, prev_y = prob.train_X[-1:].expand(n_r, -1), prob.train_Y[-1:].expand(n_r, -1)
        times = []
        for _ in range(3):
            t0 = time.perf_counter()
            o = ho.hes_forward(spec, prev_X, prev_y, None, [torch.sigmoid(m) for m in maps], gp=gp, design_base=base,
                               action_base=abase, enable_noise=False)
            tot = (o["acqf_loss"] + (2 * math.sqrt(d) - o["acqf_cost"]).abs()).mean()
            torch.autograd.grad(tot, maps)
            times.append(time.perf_counter() - t0)
        dt = min(times[1:])
        return dict(config="configs[2]: Hartmann 6D HES-MC tree nf=[64,4,2], n=500",
                    cpu_baseline=dict(value=sum(cnt) / dt, unit=UNIT, cores=cores, kind="port",
                                      sample=f"1 restart = {sum(cnt)} tree nodes, forward + autograd backward, best of 2 "
                                             f"after a warm-up ({dt:.1f} s each); extrapolated per node to 256 restarts "
                                             "(the dense fantasy models need ~4.5 GB of host memory per restart)"),
                    seconds=dt)
    if cfg == 4:
        prob = make_problem("Cosine8", seed=2)
        d, n, Hh, L, N = prob.d, prob.train_X.shape[0], 64, 3, 64
        ls = torch.full((1, d), prob.lengthscale, dtype=F64)
        gp = gp_math.DenseGP.fit_caches(prob.train_X, prob.train_Y, kind=prob.kernel, lengthscale=ls,
                                        outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
        torch.manual_seed(2)
        pth = gp_math.draw_matheron_path(gp, prob.train_Y, 1)
        from nonmyopia_b200.amortized_network import AmortizedNetwork
        net = AmortizedNetwork(d + 1, d, Hh, 1, torch.tensor([[0.0, 1.0]] * d, dtype=F64)).double()
        spec = ho.HESSpec(L, [1] * L, 1, ho.CostSpec(radius=prob.radius, k=1, max_noise=1e-5, p_norm=2.0), enable_ts=True)
        prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
        prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
        prev_h = torch.randn(N, Hh, dtype=F64)

        def run(n_iter):
            P = {k: v.detach().clone() for k, v in net.state_dict().items()}
            t0 = time.perf_counter()
            ho.query_loop(spec, prev_x, prev_y, prev_h, P, budget=2 * math.sqrt(d), lr=1e-3, n_iter=n_iter,
                          past_initial=False, path=pth)
            return time.perf_counter() - t0

        run(2)
        n_iter = int(max(4, min(150, 10.0 / max(run(4) / 4, 1e-6))))
        dt = run(n_iter)
        return dict(config="configs[3]: Cosine8 8D HES-TS-AM-3, n=200, 64 restarts, FP64",
                    cpu_baseline=dict(value=N * (L + 1) * n_iter / dt, unit=UNIT, cores=cores, kind="port",
                                      sample=f"64 restarts x 4 levels, {n_iter} Adam iterations ({dt:.1f} s)"),
                    seconds=dt)
    raise ValueError(cfg)


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path for this metric.  BoTorch/GPyTorch are not installable
    here (no network, absent from /opt/wheelhouse) and /root/reference does not exist on the GPU box, so the
    arm times the oracle port - the reference's algorithm restated in plain PyTorch and pinned against the
    reference's own files in the build container (oracle/make_golden.py) - on all host cores."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step_s = 3.0
    rates, ms = [], []
    chains = args.restarts * args.paths
    t_all = time.perf_counter()
    for _ in range(args.steps):
        r = cpu_oracle_rate(chains, args.lookahead, per_step_s, cores)
        rates.append(r["value"])
        ms.append(r["ms_per_step"])
        if time.perf_counter() - t_all > 240:          # keep the whole arm within a few minutes
            break
    rates.sort()
    value = rates[len(rates) // 2]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(rates), "warmup": args.warmup, "ms_per_step": sorted(ms)[len(ms) // 2],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"Ackley-2D HES-TS-AM-{args.lookahead}, n=50 GP points, H=64, FP64; each step = a "
                               f"bounded sample ({chains} chains, ~{per_step_s:.0f} s of Adam iterations) of the "
                               f"{args.restarts}x{args.paths}-chain workload, CPU oracle port on all host cores"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": r["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- B200 arm
def algorithmic_flops_per_node(d, hidden, n_train, n_freq):
    """FP64 flops per node eval, FMA = 2, transcendental = 1: Thompson path fwd+bwd
    (value and input gradient share one pass) + policy step fwd, bwd (dgrad) and weight gradient."""
    M = 2 * n_freq
    ts = 2 * M * d + 6 * M + 2 * n_train * (3 * d + 12)
    net = 2 * (10 * hidden * hidden + (2 * d + 1) * hidden)
    return ts, net


def run_b200(args, rank, world, local_rank):
    import ctypes as C
    import torch.distributed as dist
    from nonmyopia_b200 import _lib, ts_engine as te
    from nonmyopia_b200.gp import SurrogateGP, draw_thompson_path
    from nonmyopia_b200.ts_loop import TSAMLoop

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback exists)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    n_r, S, Lk = args.restarts, args.paths, args.lookahead
    N = n_r * S
    selfcheck = None
    if world > 1 and not args.no_selfcheck:
        selfcheck = parity_selfcheck(group, rank, world, args.exchange)
        if not selfcheck["ok"]:
            if rank == 0:
                print(json.dumps({"parity_selfcheck": selfcheck}), file=sys.stderr, flush=True)
            raise SystemExit("multi-GPU parity self-check failed: nothing is timed on a wrong result")
    prob, prev_x, prev_y, prev_h = host_workload(n_r, S, rank)
    gp = SurrogateGP(prob.train_X, prob.train_Y, kernel=prob.kernel, lengthscale=torch.full((D,), prob.lengthscale),
                     outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const, device=dev)
    torch.manual_seed(2)
    path = draw_thompson_path(gp, S)
    net = init_policy(D, H)
    theta0 = net.flat_parameters().detach()
    budget = 2 * math.sqrt(D)
    eng = te.TSRollout(path, D, H, Lk, N, cost_hypers(), device=dev, chains_per_path=n_r, seed=12345 + rank)
    loop = TSAMLoop(eng, theta0, prev_x, prev_y, prev_h, budget=budget, lr=1e-3, past_initial=False,
                    n_global=N * world, max_iter=10**9, patience=-1, group=group, exchange=args.exchange)
    stream = torch.cuda.Stream(device=dev)
    nodes_per_step = N * (Lk + 1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident loop, inputs already in HBM, one CUDA graph replay per step
    with torch.cuda.stream(stream):
        try:
            loop.capture()
            replay = loop.graph.replay
            mode = "cuda-graph"
        except Exception as e:                                  # NCCL capture unsupported: eager launches
            if world == 1:
                raise
            loop.graph = None
            replay = loop._iteration
            mode = f"eager ({type(e).__name__})"
        for _ in range(max(args.warmup, 3)):
            replay()
        barrier()
        clocks = ClockSampler(local_rank)
        if rank == 0:
            clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            replay()
        e1.record(stream)
        barrier()
        ms_total = e0.elapsed_time(e1)
        clk = clocks.stop() if rank == 0 else None
        barrier()
        # ---- sustained: the same replay back to back for >= args.sustained_s seconds (a real Actor.query is 500
        # iterations; the 100-step number above never leaves the burst clock state)
        sustained = None
        if args.sustained_s > 0:
            n_sus_t = torch.tensor([max(args.steps, int(args.sustained_s * 1e3 / max(ms_total / args.steps, 1e-3)))],
                                   device=dev, dtype=torch.int64)
            if world > 1:               # every rank must replay the SAME number of iterations (they exchange)
                dist.broadcast(n_sus_t, src=0)
            n_sus = int(n_sus_t.item())
            sclk = ClockSampler(local_rank)
            if rank == 0:
                sclk.start()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            s0.record(stream)
            for _ in range(n_sus):
                replay()
            s1.record(stream)
            barrier()
            sustained = {"steps": n_sus, "ms_total": s0.elapsed_time(s1), "clocks": sclk.stop() if rank == 0 else None}
            barrier()       # rank 0 may spend a moment reaping nvidia-smi: nobody enters the next exchange without it
        # per-kernel timing of the same steps (eager launches, events on the launching stream)
        # (every rank runs them when sharded: the iteration holds the two all-reduces)
        kern = time_kernels(loop, stream, min(args.steps, 20)) if (rank == 0 or world > 1) else {}
        # ---- e2e: the public class API with HOST buffers, copies inside the timed region
        e2e = run_e2e_loop(args, loop, prev_x, prev_y, prev_h, stream, barrier)
        e2e_api = run_e2e(args, gp, path, net, prev_x, prev_y, prev_h, budget, dev, stream, barrier)
        aq = run_actor_query(args, gp, path, net, prob, n_r, S, Lk, dev, group) if world == 1 else None
    t = torch.tensor([ms_total, e2e["ms_total"], e2e_api["ms_total"], sustained["ms_total"] if sustained else 0.0],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, api_ms, sus_ms = float(t[0]), float(t[1]), float(t[2]), float(t[3])
    st = loop.read_state()
    assert st.step >= args.steps and not st.done, "timed steps did not run"
    if rank != 0:
        _finish(world)
        return
    value = nodes_per_step * world * args.steps / (ms_total * 1e-3)
    f_ts, f_net = algorithmic_flops_per_node(D, H, gp.n, path.W.shape[0])
    # dominant kernel = largest share of the step
    dom = max(kern, key=lambda k: kern[k]["ms"]) if kern else None
    peaks = json.load(open(FP64_PEAK_FILE)) if os.path.isfile(FP64_PEAK_FILE) else {}
    fp64_peak = peaks.get("f64_fma_flops", 36.4e12) / 1e12
    roof = None
    if dom:
        flops = kern[dom]["flops"]
        ach = flops / (kern[dom]["ms"] * 1e-3) / 1e12
        roof = {"bound": "fp64", "kernel": dom, "achieved": ach, "peak": fp64_peak, "unit": "TFLOP/s",
                "frac": ach / fp64_peak, "traffic": ncu_dram_traffic(dom),
                "traffic_note": "DRAM read + write bytes per launch from profiles/ncu_full_r02c_rollout_wgrad.csv (the backward "
                                "workspace, not algorithmic bytes: the kernel is FP64-pipe bound)",
                "peak_source": "own-measured FP64 DFMA peak on this pool's B200 (profiles/pipe_peaks_r01.json); "
                               "MEASURED_PEAKS.json has no FP64 entry",
                "per_kernel_ms": {k: round(v["ms"], 4) for k, v in kern.items()}}
    cpu = None
    if not args.no_cpu_baseline and world == 1:            # reported on rank 0 at N = 1 only
        cores = os.cpu_count() or 1
        r = cpu_oracle_rate(N, Lk, 12.0, cores)
        cpu = {"value": r["value"], "unit": UNIT, "cores": cores, "kind": "port", "sample": r["sample"]}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"Ackley-2D HES-TS-AM-{Lk}: {n_r} restarts x {S} Thompson paths = {N} chains/GPU x "
                               f"{Lk + 1} levels, n=50 GP points, M=1024 features, H=64, FP64 (BASELINE configs[1])",
                   "step": "one AdamW iteration of Actor.query: 2 rollouts fwd, objective, snapshot, rollout bwd, "
                           "weight gradients, AdamW", "launch_mode": mode,
                   "exchange": {"none": "single GPU", "peer": "gradient SUM + decision MIN by own kernels over NVLink peer "
                                "memory, fused with AdamW", "nccl": "two NCCL all-reduces per iteration"}[loop.exchange],
                   "l2": "no flush: the per-step working set (backward workspace "
                         f"{loop.main.save.numel() * 8 / 1e6:.0f} MB + partials) exceeds the 126 MB L2",
                   "nodes_per_step": nodes_per_step * world, "flops_per_node": {"thompson_path": f_ts, "policy_fwd": f_net,
                                                                               "policy_fwd_bwd_wgrad": 3 * f_net}},
        "roofline": roof, "cpu_baseline": cpu,
        "e2e": {"value": nodes_per_step * world * e2e["steps"] / (e2e_ms * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"], "ms_per_step": e2e_ms / e2e["steps"],
                "api": "TSAMLoop.step_host (the following step, whose inputs the caller has already handed over, is launched behind this step's result copies) - the per-iteration engine of Actor.query (HES-TS-AM): per-step H2D of "
                       "prev_X/prev_y/prev_hid_state from pinned memory (each step's copy is issued on a side stream "
                       "during the previous step's kernels, two input-buffer sets), one iteration (same launches as "
                       "`value`), D2H of the per-restart loss/cost, stream synchronise",
                "actor_query": None if aq is None else {
                    "value": nodes_per_step * aq["iterations"] / aq["seconds"], "unit": UNIT,
                    "seconds": aq["seconds"], "iterations": aq["iterations"], "ms_per_iteration": 1e3 * aq["seconds"] / aq["iterations"],
                    "h2d_bytes_per_step": aq["h2d"],
                    "api": "Actor.query (reference actor.py:439-734) in the default seed-stream mode on the same workload: "
                           "wall time of the whole call (graph capture, host-RNG ring uploads from pinned memory every 16 "
                           "iterations, state polls, final selection) / iterations"},
                "class_api_eager": {
                    "value": nodes_per_step * world * e2e_api["steps"] / (api_ms * 1e-3), "unit": UNIT,
                    "ms_per_step": api_ms / e2e_api["steps"],
                    "api": "acqfs.qMultiStepHEntropySearch.forward x2 + torch.autograd.grad + torch AdamW, eager, "
                           "same per-step H2D / D2H"}},
        "gpu_launches": loop.n_launches_per_iter * args.steps, "clocks": clk,
    }
    if sustained:
        line["sustained"] = {"value": nodes_per_step * world * sustained["steps"] / (sus_ms * 1e-3), "unit": UNIT,
                             "steps": sustained["steps"], "ms_per_step": sus_ms / sustained["steps"],
                             "seconds": sus_ms * 1e-3, "clocks": sustained["clocks"]}
    if selfcheck is not None:
        line["parity_selfcheck"] = selfcheck
    print(json.dumps(line), flush=True)
    _finish(world)


def ncu_dram_traffic(kernel_substr):
    """DRAM bytes per launch (read + write) of the named kernel from the committed `ncu --set full` capture
    (profiles/ncu_full_r02c_rollout_wgrad.csv, raw page); None when the file or the kernel is missing."""
    import csv
    path = os.path.join(ROOT, "profiles", "ncu_full_r02c_rollout_wgrad.csv")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        kn, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        for r in rows[2:]:
            if kernel_substr in r[kn]:
                return float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
    except Exception:
        pass
    return None


def parity_selfcheck(group, rank, world, exchange):
    """N > 1: before anything is timed, run the restart-sharded PRODUCT path (``Actor.query`` with
    ``parms.process_group``: CUDA-graph loop, host-stream ring, the active exchange mode, final cross-rank selection)
    on the committed golden of the reference's own ``Actor.query`` at BASELINE configs[0] full size (64 restarts,
    H = 64, <= 500 iterations, pinned by seed only; tests/golden/query_ts_am_ackley_L1_full.pt, generated by
    oracle/make_golden.py from /root/reference) and compare chosen_idx / next_X / per-restart losses / final
    weights.  The fixture is data; nothing under oracle/ is imported."""
    import tempfile
    import types
    from nonmyopia_b200 import acqfs, actor as A
    from nonmyopia_b200.gp import SurrogateGP
    from nonmyopia_b200.path import ThompsonPath
    F64 = torch.float64
    fx = torch.load(os.path.join(ROOT, "tests", "golden", "query_ts_am_ackley_L1_full.pt"), weights_only=False)
    g, pth, d = fx["gp"], fx["path"], fx["d"]
    n = g["train_X"].shape[0]
    ls = torch.full((d,), g["lengthscale"], dtype=F64)
    parms = types.SimpleNamespace(
        algo="HES", algo_ts=True, amortized=True, algo_lookahead_steps=fx["L"], x_dim=d, y_dim=1, hidden_dim=fx["H"],
        n_actions=1, n_restarts=fx["n_r"], n_samples=64, env_discretized=False, num_categories=None,
        bounds=torch.tensor([[0.0, 1.0]] * d, dtype=F64), torch_dtype=F64, device="cuda", acq_opt_lr=fx["lr"],
        acq_opt_iter=fx["max_iter"], n_initial_points=n, save_dir=tempfile.mkdtemp(), budget=fx["budget"], plot=False,
        loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
        cost_function_class=acqfs.qCostFunction, cost_func_hypers=fx["cost"], rng_mode="host", process_group=group,
        exchange=exchange)
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        act = A.Actor(parms)
    act.maps.load_state_dict({k: v.clone() for k, v in fx["policy"].items()})
    buffer = dict(x=g["train_X"].cuda(), y=g["train_Y"].cuda(), h=torch.zeros(n, fx["H"], dtype=F64, device="cuda"),
                  cost=torch.zeros(n, dtype=F64, device="cuda"))
    model = SurrogateGP(g["train_X"], g["train_Y"], kernel=g["kernel"], lengthscale=ls, outputscale=g["outputscale"],
                        noise=g["noise"], mean_const=g["mean_const"])
    act.construct_acqf(model, buffer)
    act.hidden_noise = fx["hidden_noise"].cuda()
    path = ThompsonPath(pth["W"].cuda(), pth["w"].cuda(), pth["v"].cuda(), g["train_X"].cuda(), ls.cuda(),
                        kernel=g["kernel"], outputscale=g["outputscale"], mean_const=g["mean_const"])
    torch.set_rng_state(fx["rng_state_in"])
    t0 = time.perf_counter()
    out = act.query(buffer, iteration=n, f_posterior=path, save_trajectory=False)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    ref = fx["out"]

    def rel(a, b, atol):
        a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
        return float(((a - b).abs() / b.abs().clamp_min(atol)).max())

    errs = {"next_X": rel(out["next_X"], ref["next_X"], 1e-3), "all_losses": rel(out["all_losses"], ref["all_losses"], 1e-2),
            "all_costs": rel(out["all_costs"], ref["all_costs"], 1e-3), "actions": rel(out["actions"], ref["actions"], 1e-3)}
    sd = act.maps.state_dict()
    errs["final_weights"] = max(float((sd[k].cpu() - v).abs().max() / max(1.0, float(v.abs().max())))
                                for k, v in ref["final_policy"].items())
    same_idx = int(out["chosen_idx"]) == int(ref["chosen_idx"])
    rng_ok = bool(torch.equal(torch.get_rng_state(), fx["rng_state_out"]))
    ok = same_idx and rng_ok and max(errs.values()) < 1e-7
    flag = torch.tensor([1.0 if ok else 0.0], device="cuda", dtype=F64)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return {"ok": bool(flag.item() == 1.0), "max_rel": max(errs.values()), "errors": errs, "chosen_idx": int(out["chosen_idx"]),
            "chosen_idx_reference": int(ref["chosen_idx"]), "generator_state_matches_reference": rng_ok,
            "golden": "query_ts_am_ackley_L1_full (reference Actor.query, 64 restarts, H=64, seed-pinned)",
            "n_ranks": world, "query_seconds": dt, "reference_cpu_seconds": fx.get("ref_cpu_seconds")}


def _finish(world):
    """End of a rank.  Sharded runs leave without tearing NCCL down: the communicator is referenced by the
    captured per-iteration graph, and its teardown at interpreter exit was seen to hang after the result line
    had been printed; the work is done and synchronised at this point."""
    if world > 1:
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def time_kernels(loop, stream, n_steps):
    """Average device time of every kernel of the step, CUDA events on the launching stream (eager
    launches of the same sequence the graph replays) + the algorithmic FP64 flops of each."""
    eng = loop.eng
    names, evs = [], []

    def mark(name):
        e = torch.cuda.Event(enable_timing=True)
        e.record(stream)
        names.append(name)
        evs.append(e)

    for _ in range(3):
        loop._iteration()
    torch.cuda.synchronize()
    for _ in range(n_steps):
        mark("start")
        loop._iteration(mark=mark)
    torch.cuda.synchronize()
    acc = {}
    for i in range(1, len(evs)):
        if names[i] == "start":
            continue
        acc.setdefault(names[i], []).append(evs[i - 1].elapsed_time(evs[i]))
    nodes = eng.N * (eng.L + 1)
    f_ts, f_net = algorithmic_flops_per_node(eng.d, eng.H, eng.path.train_X.shape[0], eng.path.W.shape[0])
    flops = {"ts_rollout_fwd": nodes * (2 * f_net + 1.5 * f_ts), "ts_rollout_bwd": nodes * f_net,
             "policy_wgrad": nodes * f_net, "objective+snapshot": 0.0, "adamw": 0.0}
    return {k: {"ms": sum(v) / len(v), "flops": flops.get(k, 0.0)} for k, v in acc.items()}


def run_actor_query(args, gp, path, net, prob, n_r, S, Lk, dev, group):
    """The call a user of the reference makes - ``Actor.query`` (actor.py:439-734) - on the bench workload, in the
    default seed-stream mode: the CUDA-graph loop fed by the host-RNG ring (every chunk of iterations: H2D of the
    reference-order draws from pinned memory, D2H poll of the loop state), final selection.  Wall time around the call
    (second call; 120 iterations, early stop closed by ``past_initial = False`` and < 200 iterations)."""
    import contextlib
    import io
    import tempfile
    import types
    from nonmyopia_b200 import acqfs, actor as A
    F64 = torch.float64
    N, n_it = n_r * S, 120
    world = 1 if group is None else torch.distributed.get_world_size(group)
    n = prob.train_X.shape[0]
    parms = types.SimpleNamespace(
        algo="HES", algo_ts=True, amortized=True, algo_lookahead_steps=Lk, x_dim=D, y_dim=1, hidden_dim=H, n_actions=1,
        n_restarts=N * world, n_samples=64, env_discretized=False, num_categories=None,
        bounds=torch.tensor([[0.0, 1.0]] * D, dtype=F64), torch_dtype=F64, device=dev, acq_opt_lr=1e-3, acq_opt_iter=n_it,
        n_initial_points=n, save_dir=tempfile.mkdtemp(), budget=2 * math.sqrt(D), plot=False,
        loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
        cost_function_class=acqfs.qCostFunction, cost_func_hypers=cost_hypers(), rng_mode="host", process_group=group,
        exchange=args.exchange)
    with contextlib.redirect_stdout(io.StringIO()):
        act = A.Actor(parms)
        act.maps.load_state_dict(net.state_dict())
        buffer = dict(x=prob.train_X.to(dev), y=prob.train_Y.to(dev), h=torch.zeros(n, H, dtype=F64, device=dev),
                      cost=torch.zeros(n, dtype=F64, device=dev))
        act.construct_acqf(gp, buffer)
        g = torch.Generator().manual_seed(1000)
        act.hidden_noise = torch.randn(N * world, H, generator=g).to(F64).to(dev)
        secs = []
        for _ in range(2):
            torch.manual_seed(5)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = act.query(buffer, iteration=n, f_posterior=path, save_trajectory=False)
            torch.cuda.synchronize()
            secs.append(time.perf_counter() - t0)
    assert bool(torch.isfinite(out["next_X"]).all())
    ring = N * (Lk + 1) * (2 * 16 + 2 * 8)          # bytes of draws per iteration and rank (cs, dims, noise x 2 rollouts)
    return dict(seconds=secs[-1], iterations=n_it, h2d=ring, d2h=72 // 16)


def run_e2e_loop(args, loop, prev_x, prev_y, prev_h, stream, barrier):
    """The step of `value` driven from HOST buffers: pinned inputs -> HBM, one iteration, loss/cost -> host."""
    hx, hy, hh = prev_x.pin_memory(), prev_y.pin_memory(), prev_h.pin_memory()
    N = hx.shape[0]
    out_host = torch.empty(2, N, dtype=torch.float64).pin_memory()
    nxt = (hx, hy, hh)                      # inputs of the following step: copied under this step's kernels
    for k in range(3):
        loop.step_host(hx, hy, hh, out_host, next_inputs=nxt if k < 2 else None)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(args.steps):
        # (the last call promises no following step: step_host launches a promised step ahead)
        loop.step_host(hx, hy, hh, out_host, next_inputs=nxt if k + 1 < args.steps else None)
    e1.record(stream)
    barrier()
    assert torch.isfinite(out_host).all()
    h2d = (hx.numel() + hy.numel() + hh.numel()) * 8
    return dict(ms_total=e0.elapsed_time(e1), steps=args.steps, h2d=h2d, d2h=out_host.numel() * 8)


def run_e2e(args, gp, path, net, prev_x, prev_y, prev_h, budget, dev, stream, barrier):
    """Same step through the public class API, host buffers in, host results out."""
    import copy
    from nonmyopia_b200 import acqfs
    Lk = args.lookahead
    acqf = acqfs.qMultiStepHEntropySearch(
        model=gp, loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
        cost_function_class=acqfs.qCostFunction, cost_func_hypers=cost_hypers(), lookahead_steps=Lk, n_actions=1,
        n_fantasy_at_design_pts=[1] * Lk, n_fantasy_at_action_pts=64, enable_ts=True, rng_mode="device")
    acqf.dump_model(f=path)
    maps = copy.deepcopy(net).to(dev)
    frozen = copy.deepcopy(net).to(dev)
    params = list(maps.parameters())
    opt = torch.optim.AdamW(params, lr=1e-3)
    hx, hy, hh = prev_x.pin_memory(), prev_y.pin_memory(), prev_h.pin_memory()
    N = hx.shape[0]
    out_host = torch.empty(2, N, dtype=torch.float64).pin_memory()
    cos = torch.nn.CosineSimilarity(dim=-1, eps=1e-6)

    def step():
        x = hx.to(dev, non_blocking=True)
        y = hy.to(dev, non_blocking=True)
        h = hh.to(dev, non_blocking=True)
        rd = acqf.forward(prev_X=x, prev_y=y, prev_hid_state=h, maps=maps, prev_cost=0.0, enable_noise=False)
        with torch.no_grad():
            od = acqf.forward(prev_X=x, prev_y=y, prev_hid_state=h, maps=frozen, prev_cost=0.0)
        KL = 0
        for s in range(Lk):
            KL = KL + 1 - cos(od["X"][s], rd["X"][s]).flatten()
        KL = KL + 1 - cos(od["actions"], rd["actions"]).flatten()
        loss = rd["acqf_loss"] + abs(budget - rd["acqf_cost"]) + KL
        grads = torch.autograd.grad(loss.mean(), params, allow_unused=True)
        for p_, g_ in zip(params, grads):
            p_.grad = g_
        opt.step()
        opt.zero_grad()
        out_host[0].copy_(rd["acqf_loss"].detach(), non_blocking=True)
        out_host[1].copy_(rd["acqf_cost"].detach(), non_blocking=True)
        stream.synchronize()                    # the reference reads loss / cost on the host every iteration

    steps = max(5, min(args.steps, 20))
    for _ in range(3):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    prof = None
    if os.environ.get("NMB_PROFILE_CLASS_API"):          # host profile of the eager class-API step (stderr)
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    for _ in range(steps):
        step()
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(45)
    e1.record(stream)
    barrier()
    h2d = (hx.numel() + hy.numel() + hh.numel()) * 8
    return dict(ms_total=e0.elapsed_time(e1), steps=steps, h2d=h2d, d2h=out_host.numel() * 8)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.cpu_config is not None:
        if rank == 0:
            print(json.dumps(cpu_baseline_config(args.cpu_config)), flush=True)
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
