"""GPU parity of the fantasy-tree posterior kernels (HES-MC) and the myopic / KG baselines against the
golden vectors recorded from the reference's acqfs.py (oracle/make_golden.py) and the CPU oracle."""
import pytest
import torch

from tests import helpers as H
from tests.test_dropin_gpu import make_acqf, product_gp

pytestmark = pytest.mark.gpu
F64 = torch.float64


def _samplers(fx):
    from nonmyopia_b200 import acqfs
    ds = []
    for s, b in enumerate(fx["design_base"]):
        smp = acqfs.PosteriorMeanSampler(fx["nf"][s]) if fx["nf"][s] == 1 else acqfs.SobolQMCNormalSampler(fx["nf"][s])
        smp.base_samples = (torch.zeros(fx["nf"][s], dtype=F64) if b is None else b.clone())
        ds.append(smp)
    a = acqfs.SobolQMCNormalSampler(fx["S_a"])
    a.base_samples = fx["action_base"].clone()
    return ds, a


def test_whiten_matches_oracle():
    """u = L^-1 k, m0 = k.alpha, |u|^2 and the backward of a random linear functional of them."""
    import ctypes as C
    from nonmyopia_b200 import _lib
    from nonmyopia_b200.mc_tree import gp_struct
    from oracle import gp_math
    fx = H.load("mc_hartmann_L3")
    ref = H.dense_gp(fx["gp"])
    gp = product_gp(fx["gp"])
    g, keep = gp_struct(gp, F64)
    gen = torch.Generator().manual_seed(0)
    P, d, n = 77, fx["d"], gp.n
    x = torch.rand(P, d, dtype=F64, generator=gen)
    xr = x.clone().requires_grad_(True)
    k = gp_math.kernel_matrix(xr, ref.train_X, ref.kind, ref.lengthscale, ref.outputscale)
    u_ref = k @ ref.R
    m0_ref = k @ ref.alpha
    gU = torch.randn(P, g.ldu, dtype=F64, generator=gen)
    gm = torch.randn(P, dtype=F64, generator=gen)
    (gx_ref,) = torch.autograd.grad((u_ref * gU[:, :n]).sum() + (m0_ref * gm).sum(), xr)
    xc = x.cuda()
    U = torch.zeros(P, g.ldu, dtype=F64, device="cuda")
    m0, q = torch.zeros(P, dtype=F64, device="cuda"), torch.zeros(P, dtype=F64, device="cuda")
    lib = _lib.lib()
    _lib.check(lib.nmb_gp_whiten_fwd(C.byref(g), 0, _lib.ptr(xc), C.c_int64(P), _lib.ptr(U), _lib.ptr(m0), _lib.ptr(q),
                                     _lib.stream_ptr()))
    assert H.rel_err(U[:, :n], u_ref, atol=1e-4) < 1e-8
    assert H.rel_err(m0, m0_ref, atol=1e-3) < 1e-8
    assert H.rel_err(q, (u_ref * u_ref).sum(-1), atol=1e-3) < 1e-8
    gx = torch.zeros(P, d, dtype=F64, device="cuda")
    gUc, gmc = gU.cuda(), gm.cuda()            # held: a temporary's block could be reused before the kernel runs
    _lib.check(lib.nmb_gp_whiten_bwd(C.byref(g), 0, _lib.ptr(xc), C.c_int64(P), _lib.ptr(gUc), _lib.ptr(gmc),
                                     _lib.ptr(gx), _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert H.rel_err(gx, gx_ref, atol=1e-2) < 1e-8


@pytest.mark.parametrize("name", ["mc_ackley_L1", "mc_hartmann_L3", "mc_ackley_L4_m12"])
def test_hes_mc_matches_reference_golden(name):
    fx = H.load(name)
    L, n_r, d = fx["L"], fx["n_r"], fx["d"]
    ds, a = _samplers(fx)
    acqf = make_acqf(fx, fx["nf"], False, n_samples=fx["S_a"], design_samplers=ds, action_sampler=a)
    acqf.dump_model()
    lg = [l.clone().cuda().requires_grad_(True) for l in fx["logits"]]
    acqf._given_draws = torch.cat([z.reshape(-1) for z in fx["out"]["cost_noise"]]).cuda()
    out = acqf.forward(fx["prev_X"].cuda(), fx["prev_y"].cuda(), torch.zeros(n_r, 4, dtype=F64, device="cuda"),
                       [torch.sigmoid(x) for x in lg], enable_noise=False)
    # north_star: acquisition values within 1e-8 relative in FP64
    assert H.rel_err(out["acqf_loss"], fx["out"]["acqf_loss"], atol=1e-2) < 1e-8
    assert H.rel_err(out["acqf_cost"], fx["out"]["acqf_cost"], atol=1e-3) < 1e-8
    assert out["actions"].shape == tuple(fx["nf"][::-1]) + (n_r, 1, d)
    grads = torch.autograd.grad((out["acqf_loss"] + (fx["budget"] - out["acqf_cost"]).abs()).mean(), lg)
    for s, (gd, gr) in enumerate(zip(grads, fx["out"]["grads"])):
        scale = max(float(gr.abs().max()), 1e-6)
        assert float((gd.cpu() - gr).abs().max()) / scale < 1e-7, s


def test_node_posteriors_vs_oracle():
    """Posterior mean / variance of every tree node vs the dense fantasy models of the oracle."""
    from nonmyopia_b200.mc_tree import MCTree
    from oracle import gp_math
    fx = H.load("mc_hartmann_L3")
    ref = H.dense_gp(fx["gp"])
    gp = product_gp(fx["gp"])
    nf, n_r, d = fx["nf"], fx["n_r"], fx["d"]
    levels = [torch.sigmoid(l) for l in fx["logits"]]
    tree = MCTree(gp, nf, n_r, [b for b in fx["design_base"]], fx["action_base"])
    ws = tree.forward(torch.cat(levels).cuda())
    model = ref
    for s in range(len(nf) + 1):
        X = levels[s].reshape(*nf[:s][::-1], n_r, 1, d)
        mean, cov = model.posterior(X)
        o0, o1 = tree.offsets[s], tree.offsets[s + 1]
        assert H.rel_err(ws["mean"][o0:o1], mean.reshape(-1), atol=1e-2) < 1e-8, s
        assert H.rel_err(ws["var"][o0:o1], cov.reshape(-1), atol=1e-6) < 1e-8, s
        if s < len(nf):
            base = fx["design_base"][s].reshape(nf[s], *([1] * (mean.dim() - 1)), 1)
            ys = gp_math.sample_from_posterior(mean, cov, base)
            model = model.condition_on_observations(X, ys)


def test_baselines_match_reference_golden():
    from nonmyopia_b200 import acqfs
    fx = H.load("baselines_ackley")
    gp = product_gp(fx["gp"])
    for algo, g in fx["algos"].items():
        n_r = g["n_r"]
        smp = acqfs.SobolQMCNormalSampler(64, seed=0)
        smp.base_samples = g["base"].clone()
        acqf = acqfs.qBOAcqf(name=algo, model=gp, lookahead_steps=1 if algo == "qKG" else 0, n_actions=1,
                             n_fantasy_at_design_pts=[64], loss_function_class=acqfs.qLossFunctionTopK,
                             loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5),
                             cost_function_class=acqfs.qCostFunction, cost_func_hypers=fx["cost"], sampler=smp,
                             best_f=fx["best_f"])
        maps = [m.clone().cuda().requires_grad_(True) for m in g["maps"]]
        nz = [z.cuda() for z in g["cost_noise"]]
        acqf._given_draws = nz if algo == "qKG" else nz[0].reshape(-1)
        out = acqf.forward(g["prev_X"].cuda(), None, maps)
        assert H.rel_err(out["acqf_loss"], g["acqf_loss"], atol=1e-3) < 1e-8, algo
        assert H.rel_err(out["acqf_cost"], g["acqf_cost"], atol=1e-3) < 1e-8, algo
        grads = torch.autograd.grad((out["acqf_loss"] + out["acqf_cost"]).mean(), maps)
        for gd, gr in zip(grads, g["grads"]):
            scale = max(float(gr.abs().max()), 1e-6)
            assert float((gd.cpu() - gr).abs().max()) / scale < 1e-7, algo


@pytest.mark.parametrize("algo", ["qEI", "qUCB", "qKG"])
def test_baseline_loop_graph_replay_equals_eager(algo):
    """The free-parameter loop over a BoTorch-style baseline, replayed as a CUDA graph, equals the eager loop
    (cost noise off; q = 1 trees have no atomics: bitwise; qKG's one-level tree: 1e-9)."""
    from nonmyopia_b200 import acqfs, sharding as sh
    from nonmyopia_b200.actor import _FlatAdamW
    fx = H.load("baselines_ackley")
    if algo not in fx["algos"]:
        pytest.skip(f"{algo} not in the fixture")
    g = fx["algos"][algo]
    gp = product_gp(fx["gp"])
    smp = acqfs.SobolQMCNormalSampler(64, seed=0)
    smp.base_samples = g["base"].clone()
    acqf = acqfs.qBOAcqf(name=algo, model=gp, lookahead_steps=1 if algo == "qKG" else 0, n_actions=1,
                         n_fantasy_at_design_pts=[64], loss_function_class=acqfs.qLossFunctionTopK,
                         loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5), cost_function_class=acqfs.qCostFunction,
                         cost_func_hypers=dict(fx["cost"], max_noise=0.0), sampler=smp, best_f=fx["best_f"])
    prev_X = g["prev_X"].cuda()
    outs, logged = [], []
    for use_graph in (False, True):
        params = [torch.logit(m.clone().clamp(0.01, 0.99)).cuda().requires_grad_(True) for m in g["maps"]]
        opt = _FlatAdamW(params, lr=1e-2)
        res = sh.free_restart_loop(lambda maps: acqf.forward(prev_X, None, maps), params, opt.step, budget=fx.get("budget", 2.0),
                                   n_iter=20, past_initial=True, log=logged.append, use_graph=use_graph)
        outs.append((params, res))
    # the capture must succeed: a host->device copy inside the objective (e.g. re-uploading the sampler's base samples)
    # would silently drop the loop back to eager launches
    assert not [m for m in logged if "eagerly" in m], logged
    (pa, ra), (pb, rb) = outs
    assert ra["n_iter"] == rb["n_iter"] == 20
    for a, b in zip(pa, pb):
        assert H.scaled_err(a, b) < 1e-9
    assert H.scaled_err(ra["best_loss"], rb["best_loss"]) < 1e-9


@pytest.mark.parametrize("n,d,kernel", [(333, 3, "rbf"), (700, 6, "matern52"), (800, 8, "matern32"), (97, 12, "matern52")])
def test_whiten_large_n_vs_oracle(n, d, kernel):
    """Shared-root whitening at training-set sizes where the kernels fall back to 24 / 16-row tiles and 3-stage rings
    (n up to GPyTorch's max_cholesky_size = 800), odd n (padding to ldl), d > 8: values and gradient vs the oracle."""
    import ctypes as C
    from nonmyopia_b200 import _lib
    from nonmyopia_b200.gp import SurrogateGP
    from nonmyopia_b200.mc_tree import gp_struct
    from oracle import gp_math
    gen = torch.Generator().manual_seed(n)
    X = torch.rand(n, d, dtype=F64, generator=gen)
    Y = torch.randn(n, 1, dtype=F64, generator=gen)
    ls = torch.full((1, d), 0.4, dtype=F64)
    ref = gp_math.DenseGP.fit_caches(X, Y, kind=kernel, lengthscale=ls, outputscale=1.3, noise=1e-2, mean_const=0.1)
    gp = SurrogateGP(X, Y, kernel=kernel, lengthscale=ls.reshape(-1), outputscale=1.3, noise=1e-2, mean_const=0.1)
    g, keep = gp_struct(gp, F64)
    P = 101
    x = torch.rand(P, d, dtype=F64, generator=gen)
    xr = x.clone().requires_grad_(True)
    k = gp_math.kernel_matrix(xr, ref.train_X, ref.kind, ref.lengthscale, ref.outputscale)
    u_ref, m0_ref = k @ ref.R, k @ ref.alpha
    gU = torch.randn(P, g.ldu, dtype=F64, generator=gen)
    gm = torch.randn(P, dtype=F64, generator=gen)
    (gx_ref,) = torch.autograd.grad((u_ref * gU[:, :n]).sum() + (m0_ref * gm).sum(), xr)
    xc, gUc, gmc = x.cuda(), gU.cuda(), gm.cuda()
    U = torch.zeros(P, g.ldu, dtype=F64, device="cuda")
    m0, q = torch.zeros(P, dtype=F64, device="cuda"), torch.zeros(P, dtype=F64, device="cuda")
    gx = torch.zeros(P, d, dtype=F64, device="cuda")
    lib = _lib.lib()
    _lib.check(lib.nmb_gp_whiten_fwd(C.byref(g), 0, _lib.ptr(xc), C.c_int64(P), _lib.ptr(U), _lib.ptr(m0), _lib.ptr(q),
                                     _lib.stream_ptr()))
    _lib.check(lib.nmb_gp_whiten_bwd(C.byref(g), 0, _lib.ptr(xc), C.c_int64(P), _lib.ptr(gUc), _lib.ptr(gmc), _lib.ptr(gx),
                                     _lib.stream_ptr()))
    torch.cuda.synchronize()
    # conditioning of K + s2 I at these sizes: 1e-7 of the largest entry
    assert H.scaled_err(U[:, :n], u_ref) < 1e-7
    assert H.scaled_err(m0, m0_ref) < 1e-7
    assert H.scaled_err(q, (u_ref * u_ref).sum(-1)) < 1e-7
    assert float((gx.cpu() - gx_ref).abs().max()) / float(gx_ref.abs().max()) < 1e-7
