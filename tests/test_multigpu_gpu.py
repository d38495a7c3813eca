"""Two ranks over NCCL (one process per GPU): the restart-sharded loops must reproduce the single-GPU results -
HES-TS-AM against the reference golden of Actor.query (same recorded draws, sliced per rank), the free-parameter
loop against the same Actor run unsharded.  Skipped on a one-GPU box (run with ``gpurun --gpus 2``)."""
import os
import tempfile
import types

import pytest
import torch
import torch.multiprocessing as mp

from tests import helpers as H

pytestmark = pytest.mark.gpu
F64 = torch.float64


def _need_two():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")


def _spawn(fn, *args):
    tmp = tempfile.mkdtemp()
    mp.spawn(_entry, args=(fn, os.path.join(tmp, "store"), tmp, *args), nprocs=2, join=True)


def _entry(rank, fn, store, tmp, *args):
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"file://{store}", rank=rank, world_size=2,
                            device_id=torch.device("cuda", rank))
    try:
        fn(rank, dist.group.WORLD, tmp, *args)
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- HES-TS-AM (device loop)
def _ts_am_worker(rank, group, tmp, exchange):
    from nonmyopia_b200 import sharding as sh, ts_engine as te
    from nonmyopia_b200.ts_loop import TSAMLoop
    from tests.test_path_gpu import _product_path
    from tests.test_ts_rollout_gpu import pack_draws
    fx = H.load("query_ts_am_ackley_L1")
    d, Hh, L, N = fx["d"], fx["H"], fx["L"], fx["n_r"]
    lo, hi = sh.restart_range(N, rank, 2)
    eng = te.TSRollout(_product_path(fx), d, Hh, L, hi - lo, fx["cost"])
    loop = TSAMLoop(eng, te.flatten_params(fx["policy"]), fx["prev_X"][lo:hi], fx["prev_y"][lo:hi], fx["prev_h"][lo:hi],
                    budget=fx["budget"], lr=fx["lr"], past_initial=False, max_iter=fx["n_iter"], n_global=N, group=group,
                    exchange=exchange)
    assert loop.exchange == exchange
    draws = []
    for dr in fx["draws"]:
        full = (*pack_draws(dr["main_rot"], dr["main_noise"], L, N), *pack_draws(dr["orig_rot"], dr["orig_noise"], L, N))
        draws.append(tuple(t[:, lo:hi].contiguous() for t in full))
    st = loop.run(fx["n_iter"], draws_per_iter=draws)
    assert st.step == fx["n_iter"]
    out = fx["out"]
    total = loop.best_loss + loop.best_cost + loop.best_KL
    _, idx, owner, li = sh.select_best(total, N, group)
    assert idx == out["chosen_idx"]
    assert H.rel_err(sh.gather_restarts(loop.best_loss, N, group), out["all_losses"], atol=1e-2) < 1e-8
    assert H.rel_err(sh.gather_restarts(loop.best_cost, N, group), out["all_costs"], atol=1e-3) < 1e-8
    nx = sh.broadcast_from_owner(loop.best_X[0][li].reshape(1, -1) if rank == owner else None, (1, d), F64, "cuda",
                                 owner, group)
    assert H.rel_err(nx, out["next_X"], atol=1e-3) < 1e-8
    final = te.unflatten_params(loop.theta.cpu(), d, Hh)
    for k, ref in out["final_policy"].items():
        assert H.rel_err(final[k], ref, atol=1e-3) < 1e-8, k


@pytest.mark.parametrize("exchange", ["nccl", "peer"])
def test_ts_am_loop_sharded_matches_reference_golden(exchange):
    """exchange = "peer": gradient SUM / decision MIN through the nmb_peer_* kernels over NVLink peer memory."""
    _need_two()
    _spawn(_ts_am_worker, exchange)


def _graph_worker(rank, group, tmp, exchange):
    """Default mode: collectives captured in the per-iteration CUDA graph; the replicated policy must stay
    bit-identical on both ranks and the loop state must advance in lock step."""
    import torch.distributed as dist
    from nonmyopia_b200 import sharding as sh, ts_engine as te
    from nonmyopia_b200.ts_loop import TSAMLoop
    from tests.test_path_gpu import _product_path
    fx = H.load("ts_am_ackley_L3")
    d, Hh, L, N = fx["d"], fx["H"], fx["L"], fx["n_r"]
    lo, hi = sh.restart_range(N, rank, 2)
    eng = te.TSRollout(_product_path(fx), d, Hh, L, hi - lo, fx["cost"], seed=77 + rank)
    loop = TSAMLoop(eng, te.flatten_params(fx["policy"]), fx["prev_X"][lo:hi], fx["prev_y"][lo:hi], fx["prev_h"][lo:hi],
                    budget=fx["budget"], lr=1e-3, past_initial=True, max_iter=64, n_global=N, group=group,
                    exchange=exchange)
    st = loop.run(32, poll_every=8)
    assert st.step == 32
    both = [torch.empty_like(loop.theta) for _ in range(2)]
    dist.all_gather(both, loop.theta, group=group)
    assert torch.equal(both[0], both[1])
    assert not torch.equal(loop.theta, loop.theta0) and bool(torch.isfinite(loop.theta).all())
    flags = torch.tensor([float(st.step), float(st.done)], device="cuda", dtype=F64)
    other = [torch.empty_like(flags) for _ in range(2)]
    dist.all_gather(other, flags, group=group)
    assert torch.equal(other[0], other[1])


@pytest.mark.parametrize("exchange", ["nccl", "peer"])
def test_ts_am_graph_loop_two_ranks_lockstep(exchange):
    _need_two()
    _spawn(_graph_worker, exchange)


# ----------------------------------------------------------------------------- free parameters through Actor
def _parms(fx, tmp, algo_ts, group, n_iter):
    from nonmyopia_b200 import acqfs
    d = fx["d"]
    cost = dict(fx["cost"], max_noise=0.0)       # per-call device randn is not a shardable stream
    return types.SimpleNamespace(
        algo="HES", algo_ts=algo_ts, amortized=False, algo_lookahead_steps=fx["L"], x_dim=d, y_dim=1, hidden_dim=4,
        n_actions=1, n_restarts=fx["n_r"], n_samples=fx.get("S_a", 64), env_discretized=False, num_categories=None,
        bounds=torch.tensor([[0.0, 1.0]] * d, dtype=F64), torch_dtype=F64, device="cuda", acq_opt_lr=1e-2,
        acq_opt_iter=n_iter, n_initial_points=fx["gp"]["train_X"].shape[0], save_dir=tmp, budget=fx["budget"], plot=False,
        loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
        cost_function_class=acqfs.qCostFunction, cost_func_hypers=cost, rng_mode="device", process_group=group)


def _free_worker(rank, group, tmp, name):
    from nonmyopia_b200 import actor as A, sharding as sh
    from tests.test_dropin_gpu import product_gp
    from tests.test_path_gpu import _product_path
    fx = H.load(name)
    ts = fx["kind"] == "ts_free"
    g = fx["gp"]
    n = g["train_X"].shape[0]
    buffer = dict(x=g["train_X"].cuda(), y=g["train_Y"].cuda(), h=torch.zeros(n, 4, dtype=F64, device="cuda"),
                  cost=torch.zeros(n, dtype=F64, device="cuda"))
    outs = []
    for grp in (None, group):
        parms = _parms(fx, tmp, ts, grp, 12)
        act = A.Actor(parms)
        torch.manual_seed(4)                     # Sobol sampler seeds come from the host generator
        act.construct_acqf(product_gp(g), buffer)
        torch.manual_seed(5)                     # same host draws on every rank, then sliced
        act.reset_parameters(buffer)
        if grp is not None:
            lo, hi = sh.restart_range(fx["n_r"], rank, 2)
            assert act.maps[0].shape[0] == hi - lo
        outs.append(act.query(buffer, iteration=n, f_posterior=_product_path(fx) if ts else None,
                              save_trajectory=False))
    one, two = outs
    assert int(one["chosen_idx"]) == int(two["chosen_idx"])
    assert two["all_losses"].shape == one["all_losses"].shape
    assert H.rel_err(two["all_losses"], one["all_losses"], atol=1e-2) < 1e-8
    assert H.rel_err(two["all_costs"], one["all_costs"], atol=1e-3) < 1e-8
    assert H.rel_err(two["next_X"], one["next_X"], atol=1e-3) < 1e-8
    assert two["actions"].shape == one["actions"].shape
    assert H.rel_err(two["actions"], one["actions"], atol=1e-3) < 1e-8


@pytest.mark.parametrize("name", ["ts_free_ackley_L2", "mc_ackley_L1"])
def test_actor_query_free_sharded_equals_unsharded(name):
    _need_two()
    _spawn(_free_worker, name)


# ----------------------------------------------------------------------------- HES-TS-AM through Actor, both rng modes
def _actor_am_worker(rank, group, tmp, rng_mode):
    """Actor.query (HES-TS-AM) sharded over two ranks == the same query unsharded, in the seed-stream mode ("host":
    every rank walks the full generator stream and keeps its restarts) and in the in-kernel mode ("device": the Philox
    counter is keyed by the GLOBAL chain index)."""
    from nonmyopia_b200 import actor as A
    from tests.test_dropin_gpu import _parms, product_gp
    from tests.test_path_gpu import _product_path
    fx = H.load("query_ts_am_ackley_L1")
    n = fx["gp"]["train_X"].shape[0]
    g = fx["gp"]
    buffer = dict(x=g["train_X"].cuda(), y=g["train_Y"].cuda(), h=torch.zeros(n, fx["H"], dtype=F64, device="cuda"),
                  cost=torch.zeros(n, dtype=F64, device="cuda"))
    outs = []
    for grp in (None, group):
        parms = _parms(fx, tmp, 40, rng_mode)
        parms.process_group = grp
        act = A.Actor(parms)
        act.maps.load_state_dict({k: v.clone() for k, v in fx["policy"].items()})
        act.construct_acqf(product_gp(g), buffer)
        act.hidden_noise = fx["prev_h"].cuda()
        torch.manual_seed(11)
        outs.append(act.query(buffer, iteration=n, f_posterior=_product_path(fx), save_trajectory=False))
        state = torch.get_rng_state()
        if grp is None:
            state_one = state
    one, two = outs
    assert int(one["chosen_idx"]) == int(two["chosen_idx"])
    assert H.rel_err(two["all_losses"], one["all_losses"], atol=1e-2) < 1e-8
    assert H.rel_err(two["all_costs"], one["all_costs"], atol=1e-3) < 1e-8
    assert H.rel_err(two["next_X"], one["next_X"], atol=1e-3) < 1e-8
    assert H.rel_err(two["hidden_state"], one["hidden_state"], atol=1e-3) < 1e-8
    if rng_mode == "host":
        assert torch.equal(state, state_one)        # both leave the generator where the reference would


@pytest.mark.parametrize("rng_mode", ["host", "device"])
def test_actor_query_amortized_sharded_equals_unsharded(rng_mode):
    _need_two()
    _spawn(_actor_am_worker, rng_mode)
