#!/usr/bin/env python
"""Secondary measurements for BASELINE.json configs[2], [3] and [4] (SURVEY 8d cfg 3-5); bench.py stays the
headline (configs[1]).  One JSON line per case on stdout.

    python tools/bench_configs.py mc     [--restarts 256] [--steps 10]   # cfg 3: Hartmann-6D HES-MC tree
    python tools/bench_configs.py tf32                                     # cfg 4: Cosine8 FP64 vs FP32 variant
    python tools/bench_configs.py sweep  [--max-restarts 8192]            # cfg 5: restarts x lookahead sweep

Every case runs this repo's CUDA path only (no oracle, no CPU fallback); timing = CUDA events on the launching
stream after warm-up.  Multi-GPU: launch under torchrun, every rank runs ``restarts / world`` restarts.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch

F64 = torch.float64
PEAKS = os.path.join(ROOT, "profiles", "pipe_peaks_r01.json")


def fp64_peak():
    try:
        return float(json.load(open(PEAKS))["f64_fma_flops"])
    except Exception:
        return None


def timed(fn, steps, warmup=3):
    stream = torch.cuda.current_stream()
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def make_gp(env, dev, **hyp):
    from nonmyopia_b200.gp import SurrogateGP
    from nonmyopia_b200.synthetic import make_problem
    prob = make_problem(env, seed=2, **hyp)
    ls = torch.full((prob.d,), prob.lengthscale, dtype=F64)
    gp = SurrogateGP(prob.train_X.to(dev), prob.train_Y.to(dev), kernel=prob.kernel, lengthscale=ls.to(dev),
                     outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
    return prob, gp


def cost_hypers(radius):
    return dict(radius=radius, k=1.0, max_noise=1e-5, p_norm=2.0, discount=0.0, discount_threshold=-1.0)


# ----------------------------------------------------------------------------- cfg 3: HES-MC tree
def run_mc(args, rank, world):
    """Hartmann-6D, n = 500, nf = [64, 4, 2] (reference actor.py:98-99), S_a = 64 leaf samples: one step = the
    free-parameter Adam iteration of Actor.query (acqf.forward over the fantasy tree, autograd backward to every
    design, fused AdamW)."""
    from nonmyopia_b200 import acqfs, sharding as sh
    from nonmyopia_b200.actor import _FlatAdamW, _logit, generate_random_points_batch
    dev = torch.device("cuda", torch.cuda.current_device())
    prob, gp = make_gp("Hartmann", dev, lengthscale=0.5)
    d, n = prob.d, prob.train_X.shape[0]
    nf = [int(v) for v in args.widths.split(",")]
    L = len(nf)
    lo, hi = sh.restart_range(args.restarts, rank, world)
    n_r = hi - lo
    torch.manual_seed(2)
    acqf = acqfs.qMultiStepHEntropySearch(
        model=gp, loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
        cost_function_class=acqfs.qCostFunction, cost_func_hypers=cost_hypers(prob.radius), lookahead_steps=L,
        n_actions=1, n_fantasy_at_design_pts=nf, n_fantasy_at_action_pts=64, enable_ts=False, rng_mode="device")
    acqf.dump_model()
    prev_X = prob.train_X[-1:].expand(n_r, -1).contiguous().to(dev)
    prev_y = prob.train_Y[-1:].expand(n_r, -1).contiguous().to(dev)
    prev_h = torch.zeros(n_r, 4, dtype=F64, device=dev)
    levels, parent = [], prob.train_X[-1:].expand(n_r, -1)
    for s in range(L + 1):
        parent = generate_random_points_batch(parent, prob.radius, 1 if s == 0 else nf[s - 1])
        parent = parent.reshape(-1, d) if s == 0 else parent
        levels.append(parent.reshape(-1, d))
    params = [_logit(torch.clamp(x.to(dev), 0.01, 0.99)).requires_grad_(True) for x in levels]
    opt = _FlatAdamW(params, lr=1e-2)
    budget = 2 * math.sqrt(d)

    def objective(maps):
        return acqf.forward(prev_X=prev_X, prev_y=prev_y, prev_hid_state=prev_h, maps=maps, prev_cost=0.0, enable_noise=False)

    group = None
    if world > 1:
        import torch.distributed as dist
        group = dist.group.WORLD

    logged = []

    def run(n_iter):
        """The product loop (sharding.free_restart_loop = Actor._query_free): graph-replayed iteration, two decision
        minima read on the host every iteration (MIN-reduced over the ranks when sharded); CUDA-event time of the
        replayed iterations."""
        t = {}
        sh.free_restart_loop(objective, params, opt.step, budget=budget, n_iter=n_iter, past_initial=False,
                             n_global=args.restarts, group=group, log=logged.append, use_graph=not args.eager, timing=t)
        return t.get("ms_per_iter")

    run(6)                                    # allocator growth, module loading
    vals = [run(4 + args.steps) for _ in range(3)]
    fell_back = [m for m in logged if "eagerly" in m]
    mode = "eager" if (args.eager or fell_back) else "cuda-graph"
    ms = min(v for v in vals if v is not None) if any(v is not None for v in vals) else float("nan")
    print(f"rank {rank}: {[v and round(v, 3) for v in vals]} ms/step ({mode}) {fell_back[:1]}", file=sys.stderr, flush=True)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=dev, dtype=F64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    per_restart = sum(math.prod(nf[:s]) for s in range(L + 1))
    nodes = args.restarts * per_restart
    # F_MC(d, n, s) ~ 2 n^2 + 2 n_s (3d + 12) + 8 n s per node at depth s
    flops = args.restarts * sum(math.prod(nf[:s]) * (2 * n * n + 2 * (n + s) * (3 * d + 12) + 8 * n * s)
                                for s in range(L + 1))
    peak = fp64_peak()
    out = dict(case="cfg3_hartmann_hes_mc", metric="hes_lookahead_node_posterior_evals_per_s_fwd_bwd",
               value=nodes / (ms * 1e-3), unit="node evals/s", n_gpus=world, ms_per_step=ms, steps=args.steps, launch_mode=mode,
               dtype="f64", config=dict(workload=f"Hartmann-6D HES-MC tree nf={nf}, n={n} GP points, {args.restarts} restarts "
                                        f"({n_r} per rank), S_a=64, FP64 (BASELINE configs[2])",
                                        nodes_per_step=nodes, flops_per_step=flops),
               roofline=dict(bound="fp64", achieved=flops / (ms * 1e-3) / 1e12, peak=peak and world * peak / 1e12,
                             unit="TFLOP/s", peak_note=f"{world} x the measured single-GPU FP64 peak",
                             frac=peak and flops / (ms * 1e-3) / (world * peak)))
    if rank == 0:
        print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------- cfg 4: FP64 vs FP32 variant
def run_tf32(args, rank, world):
    """Cosine8 8-D, n = 200, HES-TS-AM: identical inputs and in-kernel draws through the FP64 and the FP32 kernels;
    reports the relative deviations (north_star: 1e-3) and whether the arg-min restart agrees, plus both rates."""
    from nonmyopia_b200 import ts_engine as te
    from nonmyopia_b200.amortized_network import AmortizedNetwork
    from nonmyopia_b200.gp import draw_thompson_path
    from nonmyopia_b200.path import ThompsonPath
    from nonmyopia_b200.ts_loop import TSAMLoop
    dev = torch.device("cuda", torch.cuda.current_device())
    prob, gp = make_gp("Cosine8", dev)
    d, H = prob.d, 64
    for L in (3, 20):
        N = args.restarts
        torch.manual_seed(2)
        path64 = draw_thompson_path(gp, 1)
        net = AmortizedNetwork(d + 1, d, H, 1, torch.tensor([[0.0, 1.0]] * d, dtype=F64)).double()
        theta = net.flat_parameters().detach()
        g = torch.Generator().manual_seed(7)
        prev_h = torch.randn(N, H, generator=g).to(F64)
        prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
        prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
        res = {}
        for dt in (torch.float64, torch.float32):
            path = path64 if dt == F64 else ThompsonPath(
                path64.W.to(dt), path64.w.to(dt), path64.v.to(dt), path64.train_X.to(dt), path64.lengthscale.to(dt),
                kernel=path64.kernel, outputscale=path64.outputscale, mean_const=path64.mean_const)
            eng = te.TSRollout(path, d, H, L, N, cost_hypers(prob.radius), dtype=dt, device=dev, seed=99)
            loop = TSAMLoop(eng, theta, prev_x, prev_y, prev_h, budget=2 * math.sqrt(d), lr=1e-3, past_initial=False,
                            max_iter=10**9, patience=-1)
            loop.run(1, use_graph=False)
            first = dict(loss=loop.main.loss.double().clone(), cost=loop.main.cost.double().clone(),
                         A=loop.main.A.double().clone(), Y=loop.main.Y.double().clone())
            loop.capture()
            ms = timed(loop.graph.replay, args.steps * 5)
            res[dt] = dict(first=first, ms=ms)
        a, b = res[torch.float64]["first"], res[torch.float32]["first"]
        rel = lambda k: float(((a[k] - b[k]).abs() / a[k].abs().clamp_min(1.0)).max())
        tot64, tot32 = a["loss"] + (2 * math.sqrt(d) - a["cost"]).abs(), b["loss"] + (2 * math.sqrt(d) - b["cost"]).abs()
        srt = torch.sort(tot64).values
        out = dict(case=f"cfg4_cosine8_fp64_vs_fp32_L{L}", restarts=N, lookahead=L, n_train=prob.train_X.shape[0],
                   rel_dev=dict(acqf_loss=rel("loss"), acqf_cost=rel("cost"), actions=rel("A"), path_values=rel("Y")),
                   tolerance=1e-3, argmin_fp64=int(tot64.argmin()), argmin_fp32=int(tot32.argmin()),
                   argmin_margin_fp64=float(srt[1] - srt[0]),
                   ms_per_step=dict(f64=res[torch.float64]["ms"], f32=res[torch.float32]["ms"]),
                   node_evals_per_s=dict(f64=N * (L + 1) / res[torch.float64]["ms"] * 1e3,
                                         f32=N * (L + 1) / res[torch.float32]["ms"] * 1e3))
        if rank == 0:
            print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------- cfg 5: sweep
def run_sweep(args, rank, world):
    """Alpine (n = 100) and SynGP (n = 25), HES-TS-AM, restarts 64 ... max x lookahead 1-4: graph-replayed
    Adam iterations per case."""
    from nonmyopia_b200 import ts_engine as te
    from nonmyopia_b200.amortized_network import AmortizedNetwork
    from nonmyopia_b200.gp import draw_thompson_path
    from nonmyopia_b200.ts_loop import TSAMLoop
    dev = torch.device("cuda", torch.cuda.current_device())
    group = None
    if world > 1:
        import torch.distributed as dist
        group = dist.group.WORLD
    peak = fp64_peak()
    for env in ("Alpine", "SynGP"):
        prob, gp = make_gp(env, dev)
        d, H, n = prob.d, 64, prob.train_X.shape[0]
        torch.manual_seed(2)
        path = draw_thompson_path(gp, 1)
        net = AmortizedNetwork(d + 1, d, H, 1, torch.tensor([[0.0, 1.0]] * d, dtype=F64)).double()
        theta = net.flat_parameters().detach()
        n_r = 64
        while n_r <= args.max_restarts:
            for L in (1, 2, 3, 4):
                N = n_r                                            # per rank (weak scaling)
                g = torch.Generator().manual_seed(7 + rank)
                prev_h = torch.randn(N, H, generator=g).to(F64)
                prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
                prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
                eng = te.TSRollout(path, d, H, L, N, cost_hypers(prob.radius), device=dev, seed=99 + rank)
                loop = TSAMLoop(eng, theta, prev_x, prev_y, prev_h, budget=2 * math.sqrt(d), lr=1e-3,
                                past_initial=False, n_global=N * world, max_iter=10**9, patience=-1, group=group)
                loop.capture()
                ms = timed(loop.graph.replay, args.steps * 5, warmup=5)
                if world > 1:
                    t = torch.tensor([ms], device=dev, dtype=F64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    ms = float(t)
                nodes = N * world * (L + 1)
                f_node = 2 * 1024 * d + 6 * 1024 + 2 * n * (3 * d + 12) + 3 * 2 * (10 * H * H + (2 * d + 1) * H)
                out = dict(case="cfg5_sweep", env=env, n_train=n, restarts_per_gpu=N, n_gpus=world, lookahead=L,
                           ms_per_step=ms, node_evals_per_s=nodes / ms * 1e3,
                           fp64_frac=peak and nodes * f_node / (ms * 1e-3) / (world * peak))
                if rank == 0:
                    print(json.dumps(out), flush=True)
                del loop, eng
            n_r *= 4 if n_r < 4096 else 2


# ----------------------------------------------------------------------------- cfg 5, MC half
def run_sweep_mc(args, rank, world):
    """Alpine (n = 100) and SynGP (n = 25), HES-MC with the reference's widths (actor.py:92-102), lookahead 1-4,
    16 (the reference's own setting, main.py:71-72) ... max restarts SHARDED over the ranks (strong scaling): the
    product free-parameter loop (graph replay, two decision minima per iteration)."""
    from nonmyopia_b200 import acqfs, sharding as sh
    from nonmyopia_b200.actor import _FlatAdamW, _logit, generate_random_points_batch
    dev = torch.device("cuda", torch.cuda.current_device())
    group = None
    if world > 1:
        import torch.distributed as dist
        group = dist.group.WORLD
    peak = fp64_peak()
    widths = {1: [64], 2: [64, 8], 3: [64, 4, 2], 4: [64, 4, 2, 1]}
    for env in ("Alpine", "SynGP"):
        prob, gp = make_gp(env, dev)
        d, n = prob.d, prob.train_X.shape[0]
        n_all = 16
        while n_all <= args.max_restarts:
            for L in (1, 2, 3, 4):
                nf = widths[L]
                lo, hi = sh.restart_range(n_all, rank, world)
                n_r = hi - lo
                torch.manual_seed(2)
                acqf = acqfs.qMultiStepHEntropySearch(
                    model=gp, loss_function_class=acqfs.qLossFunctionTopK,
                    loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5), cost_function_class=acqfs.qCostFunction,
                    cost_func_hypers=cost_hypers(prob.radius), lookahead_steps=L, n_actions=1,
                    n_fantasy_at_design_pts=nf, n_fantasy_at_action_pts=64, enable_ts=False, rng_mode="device")
                acqf.dump_model()
                prev_X = prob.train_X[-1:].expand(n_r, -1).contiguous().to(dev)
                prev_y = prob.train_Y[-1:].expand(n_r, -1).contiguous().to(dev)
                prev_h = torch.zeros(n_r, 4, dtype=F64, device=dev)
                levels, parent = [], prob.train_X[-1:].expand(n_r, -1)
                for s_ in range(L + 1):
                    parent = generate_random_points_batch(parent, prob.radius, 1 if s_ == 0 else nf[s_ - 1])
                    levels.append(parent.reshape(-1, d))
                params = [_logit(torch.clamp(x.to(dev), 0.01, 0.99)).requires_grad_(True) for x in levels]
                opt = _FlatAdamW(params, lr=1e-2)
                objective = lambda maps: acqf.forward(prev_X=prev_X, prev_y=prev_y, prev_hid_state=prev_h, maps=maps,
                                                      prev_cost=0.0, enable_noise=False)

                def run(n_iter):
                    t = {}
                    sh.free_restart_loop(objective, params, opt.step, budget=2 * math.sqrt(d), n_iter=n_iter,
                                         past_initial=False, n_global=n_all, group=group, log=None, use_graph=True,
                                         timing=t)
                    return t["ms_per_iter"]

                run(6)
                ms = min(run(4 + args.steps) for _ in range(2))
                if world > 1:
                    t = torch.tensor([ms], device=dev, dtype=F64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    ms = float(t)
                per_restart = sum(math.prod(nf[:s_]) for s_ in range(L + 1))
                nodes = n_all * per_restart
                flops = n_all * sum(math.prod(nf[:s_]) * (2 * n * n + 2 * (n + s_) * (3 * d + 12) + 8 * n * s_)
                                    for s_ in range(L + 1))
                out = dict(case="cfg5_sweep_mc", env=env, n_train=n, restarts=n_all, restarts_per_gpu=n_r, n_gpus=world,
                           lookahead=L, widths=nf, ms_per_step=ms, node_evals_per_s=nodes / ms * 1e3,
                           fp64_frac=peak and flops / (ms * 1e-3) / (world * peak))
                if rank == 0:
                    print(json.dumps(out), flush=True)
                del acqf, params, opt
            n_all *= 4


# ----------------------------------------------------------------------------- cfg 1: one BO step
def run_cfg1(args, rank, world):
    """BASELINE configs[0]: one BO step of Ackley 2D HES-TS-AM-1 at the reference's operating point (64 restarts,
    H = 64, <= 300 pre-training epochs, <= 500 Adam iterations with early stop) through the product ``Actor``:
    ``reset_parameters`` + ``query`` in the seed-stream mode, from the committed goldens of the reference's own run
    (wall time on the host around each call; chosen restart compared with the reference's)."""
    import tempfile
    import types
    from nonmyopia_b200 import acqfs, actor as A
    from nonmyopia_b200.gp import SurrogateGP
    from nonmyopia_b200.path import ThompsonPath
    fq = torch.load(os.path.join(ROOT, "tests", "golden", "query_ts_am_ackley_L1_full.pt"), weights_only=False)
    fp = torch.load(os.path.join(ROOT, "tests", "golden", "pretrain_ackley_L1.pt"), weights_only=False)
    import contextlib
    import io
    res = {}
    for rep in range(3):
        for tag, fx in (("reset_parameters", fp), ("query", fq)):
            g, d = fx["gp"], fx["d"]
            n = g["train_X"].shape[0]
            ls = torch.full((d,), g["lengthscale"], dtype=F64)
            parms = types.SimpleNamespace(
                algo="HES", algo_ts=True, amortized=True, algo_lookahead_steps=fx["L"], x_dim=d, y_dim=1,
                hidden_dim=fx["H"], n_actions=1, n_restarts=fx["n_r"], n_samples=64, env_discretized=False,
                num_categories=None, bounds=torch.tensor([[0.0, 1.0]] * d, dtype=F64), torch_dtype=F64, device="cuda",
                acq_opt_lr=fx["lr"], acq_opt_iter=fx.get("max_iter", 500), n_initial_points=n,
                save_dir=tempfile.mkdtemp(), budget=fx["budget"], plot=False,
                loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
                cost_function_class=acqfs.qCostFunction, cost_func_hypers=fx["cost"], rng_mode="host")
            with contextlib.redirect_stdout(io.StringIO()):
                act = A.Actor(parms)
                act.maps.load_state_dict({k: v.clone() for k, v in fx["policy"].items()})
                buffer = dict(x=g["train_X"].cuda(), y=g["train_Y"].cuda(),
                              h=torch.zeros(n, fx["H"], dtype=F64, device="cuda"), cost=torch.zeros(n, dtype=F64, device="cuda"))
                model = SurrogateGP(g["train_X"], g["train_Y"], kernel=g["kernel"], lengthscale=ls,
                                    outputscale=g["outputscale"], noise=g["noise"], mean_const=g["mean_const"])
                act.construct_acqf(model, buffer)
                torch.set_rng_state(fx["rng_state_in"])
                torch.cuda.synchronize()
                prof = None
                if rep == 2 and os.environ.get("NMB_PROFILE_CFG1"):       # host profile of the two calls (stderr)
                    import cProfile
                    prof = cProfile.Profile()
                    prof.enable()
                t0 = time.perf_counter()
                if tag == "reset_parameters":
                    act.reset_parameters(buffer)
                    torch.cuda.synchronize()
                    dt = time.perf_counter() - t0
                    ok = bool(act.pretrain_losses.numel() == fx["out"]["losses"].numel())
                    info = dict(epochs=int(act.pretrain_losses.numel()), same_early_stop_epoch_as_reference=ok)
                else:
                    act.hidden_noise = fx["hidden_noise"].cuda()
                    pth = fx["path"]
                    path = ThompsonPath(pth["W"].cuda(), pth["w"].cuda(), pth["v"].cuda(), g["train_X"].cuda(), ls.cuda(),
                                        kernel=g["kernel"], outputscale=g["outputscale"], mean_const=g["mean_const"])
                    t0 = time.perf_counter()
                    out = act.query(buffer, iteration=n, f_posterior=path, save_trajectory=False)
                    torch.cuda.synchronize()
                    dt = time.perf_counter() - t0
                    info = dict(chosen_idx=int(out["chosen_idx"]), chosen_idx_reference=fx["out"]["chosen_idx"],
                                next_X=out["next_X"].tolist(), next_X_reference=fx["out"]["next_X"].tolist(),
                                iterations=fx["n_iter"])
            if prof is not None:
                import pstats
                prof.disable()
                sys.stderr.write(f"==== host profile of {tag}\n")
                pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(28)
            res[tag] = dict(seconds=dt, reference_cpu_seconds_build_container=fx.get("ref_cpu_seconds"), **info)
    out = dict(case="cfg1_bo_step_ackley_hes_ts_am_1", note="third repetition (the first two pay library load, CUDA "
               "context and pinned-memory set-up)", **res)
    if rank == 0:
        print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("case", choices=["mc", "tf32", "sweep", "sweep_mc", "cfg1"])
    ap.add_argument("--restarts", type=int, default=None)
    ap.add_argument("--widths", default="64,4,2")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--max-restarts", type=int, default=8192)
    ap.add_argument("--eager", action="store_true", help="mc: eager launch sequence instead of the graph replay")
    args = ap.parse_args()
    if args.restarts is None:
        args.restarts = 256 if args.case == "mc" else 64
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    {"mc": run_mc, "tf32": run_tf32, "sweep": run_sweep, "sweep_mc": run_sweep_mc, "cfg1": run_cfg1}[args.case](args, rank, world)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
