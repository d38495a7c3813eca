"""Host-side restart sharding (nonmyopia_b200/sharding.py) on CPU: world_size-2 ``gloo`` process
groups drive the sharded AdamW loop with the CPU oracle as the objective and must reproduce the unsharded loop
(same snapshot decisions, same arg-min restart, parameters / best values to 1e-12)."""
import os
import tempfile

import pytest
import torch
import torch.multiprocessing as mp

from nonmyopia_b200 import sharding as sh
from oracle import hes_oracle as ho
from tests import helpers as H

F64 = torch.float64
N_ITER = 8


def test_restart_range_covers_ragged():
    for n, w in [(8, 2), (7, 2), (64, 8), (13, 4), (3, 3)]:
        spans = [sh.restart_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.restart_range(1, 0, 2)
    with pytest.raises(ValueError):
        sh.restart_range(8, 2, 2)


def test_shard_level_is_strided_and_invertible():
    n_r, d, nf = 5, 3, [4, 2]
    t = torch.arange(nf[1] * nf[0] * n_r * d, dtype=F64).reshape(-1, d)          # [*nf_rev, n_r] flattened
    spans = [sh.restart_range(n_r, r, 2) for r in range(2)]
    parts = [sh.shard_level(t, n_r, lo, hi) for lo, hi in spans]
    tree = t.reshape(nf[1], nf[0], n_r, d)
    for (lo, hi), p in zip(spans, parts):
        assert torch.equal(p, tree[:, :, lo:hi].reshape(-1, d))
    assert torch.equal(sh.unshard_level(parts, [hi - lo for lo, hi in spans]), t)
    with pytest.raises(ValueError):
        sh.shard_level(t[:-1], n_r, 0, 2)


# ----------------------------------------------------------------------------- the loop, sharded vs not
def _problem(name, n_keep):
    """Oracle objective + initial logits of a golden fixture, restricted to its first ``n_keep`` restarts."""
    fx = H.load(name)
    n_r, L = fx["n_r"], fx["L"]
    keep = slice(0, n_keep)
    logits = [sh.shard_level(l, n_r, 0, n_keep) for l in fx["logits"]]
    prev_X, prev_y = fx["prev_X"][keep], fx["prev_y"][keep]
    ts = fx["kind"] == "ts_free"
    nf = [1] * L if ts else list(fx["nf"])
    spec = ho.HESSpec(lookahead=L, nf=nf, cost=H.cost_spec(fx["cost"]), enable_ts=ts)
    path = H.oracle_path(fx) if ts else None
    gp = None if ts else H.dense_gp(fx["gp"])

    def objective_for(lo, hi):
        px, py = prev_X[lo:hi], prev_y[lo:hi]

        def objective(maps):
            # zero cost noise: the reference draws it per call from the device generator (not shardable state)
            shapes = [(*nf[:s][::-1], hi - lo, 1, 1) for s in range(L + 1)]
            draws = ho.HESDraws(cost_noise=[torch.zeros(*s, dtype=F64) for s in shapes])
            return ho.hes_forward(spec, px, py, None, maps, path=path, gp=gp,
                                  design_base=fx.get("design_base"), action_base=fx.get("action_base"),
                                  enable_noise=False, draws=draws)
        return objective
    return fx, logits, objective_for


def _run_loop(logits, objective, n_global, group, budget):
    params = [l.clone().requires_grad_(True) for l in logits]
    opt = torch.optim.AdamW(params, lr=1e-2)

    def opt_step(grads):
        for p, g in zip(params, grads):
            p.grad = g
        opt.step()
        opt.zero_grad()
    res = sh.free_restart_loop(objective, params, opt_step, budget=budget, n_iter=N_ITER, past_initial=True,
                               n_global=n_global, group=group, log=None)
    return params, res


def _worker(rank, world, store, name, n_keep, ref_file):
    import torch.distributed as dist
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{store}", rank=rank, world_size=world)
    try:
        group = dist.group.WORLD
        ref = torch.load(ref_file, weights_only=False)
        fx, logits, objective_for = _problem(name, n_keep)
        lo, hi = sh.restart_range(n_keep, rank, world)
        mine = [sh.shard_level(l, n_keep, lo, hi) for l in logits]
        params, res = _run_loop(mine, objective_for(lo, hi), n_keep, group, fx["budget"])
        # parameters after N_ITER sharded AdamW steps == this rank's shard of the unsharded run
        for p, full in zip(params, ref["params"]):
            want = sh.shard_level(full, n_keep, lo, hi)
            assert float((p.detach() - want).abs().max()) < 1e-12
            assert float((sh.gather_level(p.detach(), n_keep, group) - full).abs().max()) < 1e-12
        # same snapshot (taken at the same iteration on every rank) and the same winner
        total = res["best_loss"] + res["best_cost"]
        totals, idx, owner, li = sh.select_best(total, n_keep, group)
        assert idx == ref["idx"], (idx, ref["idx"])
        assert float((totals - ref["totals"]).abs().max()) < 1e-12
        d = fx["d"]
        nx = sh.broadcast_from_owner(res["best_next_X"][0].reshape(-1, d)[li].reshape(1, d) if rank == owner else None,
                                     (1, d), F64, "cpu", owner, group)
        assert float((nx - ref["next_X"]).abs().max()) < 1e-12
        assert res["n_iter"] == ref["n_iter"]
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,n_keep", [("ts_free_ackley_L2", 8), ("ts_free_ackley_L2", 7), ("mc_ackley_L1", 4),
                                         ("mc_hartmann_L3", 3)])
def test_sharded_loop_matches_unsharded_gloo_world2(name, n_keep):
    fx, logits, objective_for = _problem(name, n_keep)
    params, res = _run_loop(logits, objective_for(0, n_keep), n_keep, None, fx["budget"])
    total = res["best_loss"] + res["best_cost"]
    idx = int(torch.argmin(total))
    d = fx["d"]
    tmp = tempfile.mkdtemp()
    ref_file = os.path.join(tmp, "ref.pt")
    torch.save(dict(params=[p.detach() for p in params], totals=total, idx=idx, n_iter=res["n_iter"],
                    next_X=res["best_next_X"][0].reshape(-1, d)[idx].reshape(1, d)), ref_file)
    mp.spawn(_worker, args=(2, os.path.join(tmp, "store"), name, n_keep, ref_file), nprocs=2, join=True)
