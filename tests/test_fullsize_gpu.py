"""BASELINE.json full sizes, through size-independent properties (the oracle cannot run them whole in seconds):

* configs[1] - Ackley 2D HES-TS-AM, L = 3, 64 restarts x 64 Thompson paths = 4096 chains, H = 64, FP64:
  chains are independent given their draws, so (1) a sample of chains taken from the full-size launch equals the
  oracle run on just those chains, (2) the two halves launched on their own reproduce the full launch's per-chain
  outputs, (3) the weight gradient of the full launch is the sum of the halves' gradients (linearity), and (4) the
  gradient of the sampled chains alone equals the oracle's autograd gradient.
* configs[2] - Hartmann 6D HES-MC, n = 500, widths [64, 4, 2], 256 restarts: restarts are independent, so a sample of
  restarts from the full-size tree equals the same restarts launched alone (values and design gradients), and the
  mean of the per-restart values is permutation invariant.
"""
import math

import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
F64 = torch.float64


# ------------------------------------------------------------------------------------------------ configs[1]
def _ts_setup(n_r=64, S=64, L=3, Hd=64):
    from nonmyopia_b200.gp import SurrogateGP, draw_thompson_path
    from nonmyopia_b200.synthetic import make_problem
    from oracle import hes_oracle as ho
    prob = make_problem("Ackley", seed=2)
    d = prob.d
    gp = SurrogateGP(prob.train_X, prob.train_Y, kernel=prob.kernel, lengthscale=torch.full((d,), prob.lengthscale),
                     outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
    torch.manual_seed(2)
    path = draw_thompson_path(gp, S)
    gen = torch.Generator().manual_seed(7)
    P = ho.init_policy_params(d, Hd, generator=gen)
    N = n_r * S
    hidden_noise = torch.randn(n_r, Hd, dtype=F64, generator=gen)
    prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
    prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
    prev_h = hidden_noise.repeat(S, 1).contiguous()                 # chain = path * n_r + restart (bench.py layout)
    # draws of the "orig" rollout (every call rotated): float32 angles like amortized_network.py:136,144-153
    rot, noise = [], []
    for _ in range(L + 1):
        th = torch.rand(N, generator=gen) * 2 * math.pi - math.pi
        c, s = torch.cos(th), torch.sin(th)
        Q = torch.stack([torch.stack([c, -s], -1), torch.stack([s, c], -1)], -2).to(F64)
        dims = torch.stack([torch.randperm(d, generator=gen)[:2] for _ in range(N)])
        rot.append((Q, dims))
        noise.append(torch.randn(N, 1, 1, dtype=F64, generator=gen))
    cost = dict(radius=0.075, k=1.0, max_noise=1e-5, p_norm=2.0, discount=0.0, discount_threshold=-1.0)
    return dict(prob=prob, gp=gp, path=path, P=P, d=d, H=Hd, L=L, n_r=n_r, S=S, N=N, prev_x=prev_x, prev_y=prev_y,
                prev_h=prev_h, rot=rot, noise=noise, cost=cost, budget=2 * math.sqrt(d))


def _ts_run(w, chains: torch.Tensor, paths, cpp: int, n_total: int):
    """Launch forward + backward + weight gradient on ``chains`` (indices into the full problem), whose Thompson
    paths are ``paths`` (``cpp`` consecutive chains per path); seeds are those of the full-size mean."""
    from nonmyopia_b200 import ts_engine as te
    from nonmyopia_b200.path import ThompsonPath
    from tests.test_ts_rollout_gpu import pack_draws
    d, Hd, L = w["d"], w["H"], w["L"]
    n = chains.numel()
    full = w["path"]
    sub = ThompsonPath(full.W, full.w[paths].contiguous(), full.v[paths].contiguous(), full.train_X, full.lengthscale,
                       kernel=full.kernel, outputscale=full.outputscale, mean_const=full.mean_const)
    eng = te.TSRollout(sub, d, Hd, L, n, w["cost"], dtype=F64, chains_per_path=cpp)
    theta = te.flatten_params(w["P"]).to("cuda", F64)
    theta_t = eng.transpose(theta)
    rq, rd, cn = pack_draws([(q[chains], dm[chains]) for q, dm in w["rot"]], [z[chains] for z in w["noise"]], L, n)
    buf = te.RolloutBuffers(d, Hd, L, n, F64, "cuda", with_save=True)
    prob = eng.problem(w["prev_x"][chains].cuda().contiguous(), w["prev_y"][chains].reshape(-1).cuda().contiguous(),
                       w["prev_h"][chains].cuda().contiguous())
    ro = eng.rollout_struct(buf, theta, theta_t, noisy_design=True, rot_q=rq, rot_dims=rd, cost_noise=cn)
    eng.forward(prob, [ro])
    g_loss = torch.full((n,), 1.0 / n_total, dtype=F64, device="cuda")
    g_cost = -torch.sign(w["budget"] - buf.cost) / n_total
    eng.backward(prob, ro, g_loss, g_cost)
    grad = eng.weight_grad(prob, buf)
    torch.cuda.synchronize()
    return buf, grad.clone()


def test_cfg2b_full_size_chain_independence_linearity_and_oracle_sample():
    from nonmyopia_b200 import ts_engine as te
    from oracle import gp_math, hes_oracle as ho
    w = _ts_setup()
    N, n_r, S, d, Hd, L = w["N"], w["n_r"], w["S"], w["d"], w["H"], w["L"]
    assert N == 4096
    allc = torch.arange(N)
    full, g_full = _ts_run(w, allc, torch.arange(S), n_r, N)
    assert torch.isfinite(full.loss).all() and torch.isfinite(g_full).all()

    # (2) + (3): halves reproduce the per-chain outputs, their gradients add up to the full gradient
    g_sum = torch.zeros_like(g_full)
    for h in range(2):
        ch = allc[h * N // 2:(h + 1) * N // 2]
        part, g = _ts_run(w, ch, torch.arange(h * S // 2, (h + 1) * S // 2), n_r, N)
        assert float((part.loss - full.loss[ch.cuda()]).abs().max()) < 1e-13
        assert float((part.cost - full.cost[ch.cuda()]).abs().max()) < 1e-13
        assert float((part.A - full.A[ch.cuda()]).abs().max()) < 1e-13
        g_sum += g
    assert float((g_sum - g_full).abs().max()) / float(g_full.abs().max()) < 1e-11

    # (1) + (4): a sample of chains (first / middle / last path, first / last tile rows) against the oracle
    paths, restarts = [0, 17, 63], [0, 5, 31, 63]
    chains = torch.tensor([p * n_r + r for p in paths for r in restarts])
    sub, g_sub = _ts_run(w, chains, torch.tensor(paths), len(restarts), N)
    spec = ho.HESSpec(L, [1] * L, 1, H.cost_spec(w["cost"]), enable_ts=True)
    gpd = w["gp"]
    ls = gpd.lengthscale.reshape(1, -1).cpu().to(F64)
    Pq = {k: v.clone().requires_grad_(True) for k, v in w["P"].items()}
    tot, o_loss, o_cost, o_A = 0.0, [], [], []
    pth = w["path"]
    for p in paths:
        ch = torch.tensor([p * n_r + r for r in restarts])
        op = gp_math.MatheronPath(pth.W.cpu(), pth.w[p:p + 1].cpu(), pth.v[p:p + 1].cpu(), gpd.train_X.cpu(),
                                  kind=gpd.kernel, lengthscale=ls, outputscale=gpd.outputscale,
                                  mean_const=gpd.mean_const)
        draws = ho.HESDraws(rot=[(q[ch], dm[ch]) for q, dm in w["rot"]], cost_noise=[z[ch] for z in w["noise"]])
        o = ho.hes_forward(spec, w["prev_x"][ch], w["prev_y"][ch], w["prev_h"][ch], Pq, path=op, enable_noise=True,
                           draws=draws)
        tot = tot + (o["acqf_loss"] + (w["budget"] - o["acqf_cost"]).abs()).sum() / N
        o_loss.append(o["acqf_loss"].detach()), o_cost.append(o["acqf_cost"].detach())
        o_A.append(o["actions"].detach().reshape(-1, d))
    o_loss, o_cost, o_A = torch.cat(o_loss), torch.cat(o_cost), torch.cat(o_A)
    for buf, idx in ((sub, torch.arange(chains.numel())), (full, chains)):      # the sample alone AND inside 4096
        idx = idx.cuda()
        assert H.rel_err(buf.loss[idx], o_loss, atol=1e-2) < 1e-8
        assert H.rel_err(buf.cost[idx], o_cost, atol=1e-3) < 1e-8
        assert H.rel_err(buf.A[idx], o_A, atol=1e-3) < 1e-8
    gref = dict(zip(Pq.keys(), torch.autograd.grad(tot, list(Pq.values()))))
    gd = te.unflatten_params(g_sub.cpu(), d, Hd)
    for k, ref in gref.items():
        scale = max(float(ref.abs().max()), 1e-9)
        assert float((gd[k] - ref).abs().max()) / scale < 1e-8, k


# ------------------------------------------------------------------------------------------------ configs[2]
def _mc_objective(gp, prob, nf, levels, budget, n_total):
    """``acqf.forward`` over the given restarts' trees + the seed of the full-size mean; returns values and the
    gradient on every level."""
    from nonmyopia_b200 import acqfs
    n_r = levels[0].shape[0]
    torch.manual_seed(2)
    acqf = acqfs.qMultiStepHEntropySearch(
        model=gp, loss_function_class=acqfs.qLossFunctionTopK, loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
        cost_function_class=acqfs.qCostFunction,
        cost_func_hypers=dict(radius=prob.radius, k=1.0, max_noise=0.0, p_norm=2.0, discount=0.0,
                              discount_threshold=-1.0),
        lookahead_steps=len(nf), n_actions=1, n_fantasy_at_design_pts=nf, n_fantasy_at_action_pts=64, enable_ts=False)
    acqf.dump_model()
    dev = levels[0].device
    prev_X = prob.train_X[-1:].expand(n_r, -1).contiguous().to(dev)
    prev_y = prob.train_Y[-1:].expand(n_r, -1).contiguous().to(dev)
    maps = [x.clone().requires_grad_(True) for x in levels]
    rd = acqf.forward(prev_X=prev_X, prev_y=prev_y, prev_hid_state=None, maps=maps, prev_cost=0.0, enable_noise=False)
    total = (rd["acqf_loss"] + (budget - rd["acqf_cost"]).abs()).sum() / n_total
    grads = torch.autograd.grad(total, maps)
    torch.cuda.synchronize()
    return rd["acqf_loss"].detach(), rd["acqf_cost"].detach(), [g.detach() for g in grads]


def test_cfg3_full_size_restart_independence():
    from nonmyopia_b200.actor import generate_random_points_batch
    from nonmyopia_b200.gp import SurrogateGP
    from nonmyopia_b200.synthetic import make_problem
    prob = make_problem("Hartmann", seed=2, n_points=500)
    d, n_r, nf = prob.d, 256, [64, 4, 2]
    assert prob.train_X.shape[0] == 500
    gp = SurrogateGP(prob.train_X, prob.train_Y, kernel=prob.kernel, lengthscale=torch.full((d,), 0.5),
                     outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
    torch.manual_seed(3)
    levels, parent = [], prob.train_X[-1:].expand(n_r, -1)
    for s in range(len(nf) + 1):
        parent = generate_random_points_batch(parent, prob.radius, 1 if s == 0 else nf[s - 1])
        parent = parent.reshape(-1, d) if s == 0 else parent
        levels.append(torch.clamp(parent.reshape(-1, d), 0.01, 0.99).cuda())
    budget = 2 * math.sqrt(d)
    loss, cost, grads = _mc_objective(gp, prob, nf, levels, budget, n_r)
    assert loss.shape == (n_r,) and torch.isfinite(loss).all() and all(torch.isfinite(g).all() for g in grads)

    # a sample of restarts launched alone: restart index is the fastest axis of every level
    idx = torch.tensor([0, 100, 255], device="cuda")
    sub_levels = [x.reshape(-1, n_r, d)[:, idx, :].reshape(-1, d).contiguous() for x in levels]
    s_loss, s_cost, s_grads = _mc_objective(gp, prob, nf, sub_levels, budget, n_r)
    assert float((s_loss - loss[idx]).abs().max()) < 1e-11
    assert float((s_cost - cost[idx]).abs().max()) < 1e-12
    for g_full, g_sub in zip(grads, s_grads):
        ref = g_full.reshape(-1, n_r, d)[:, idx, :].reshape(-1, d)
        assert float((g_sub - ref).abs().max()) / max(float(ref.abs().max()), 1e-12) < 1e-8

    # permutation of the restarts permutes values and gradients
    perm = torch.randperm(n_r, generator=torch.Generator().manual_seed(1)).cuda()
    p_levels = [x.reshape(-1, n_r, d)[:, perm, :].reshape(-1, d).contiguous() for x in levels]
    p_loss, p_cost, p_grads = _mc_objective(gp, prob, nf, p_levels, budget, n_r)
    assert float((p_loss - loss[perm]).abs().max()) < 1e-11
    assert float((p_cost - cost[perm]).abs().max()) < 1e-12
    assert abs(float(p_loss.mean() - loss.mean())) < 1e-12
    g0 = grads[0][perm]
    assert float((p_grads[0] - g0).abs().max()) / max(float(g0.abs().max()), 1e-12) < 1e-8
