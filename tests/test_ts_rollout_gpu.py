"""GPU parity of the fused HES-TS-AM rollout (forward, backward, weight gradients) against the
golden vectors produced by the reference's acqfs.py / amortized_network.py (oracle/make_golden.py)
and against the CPU oracle at larger sizes."""
import pytest
import torch

from tests import helpers as H
from tests.test_path_gpu import _product_path

pytestmark = pytest.mark.gpu

TS_AM = ["ts_am_ackley_L3", "ts_am_ackley_L1", "ts_am_cosine8_L3", "ts_am_alpine_m32", "ts_am_ackley_rbf_disc"]


def pack_draws(rot, cost_noise, L, N, dtype=torch.float64):
    rq = torch.zeros(L + 1, N, 4, dtype=dtype)
    rd = torch.zeros(L + 1, N, 2, dtype=torch.int32)
    for l, r in enumerate(rot):
        if r is not None:
            rq[l] = r[0].reshape(N, 4).to(dtype)
            rd[l] = r[1].to(torch.int32)
    cn = torch.stack([z.reshape(N) for z in cost_noise]).to(dtype)
    return rq.cuda(), rd.cuda(), cn.cuda()


def run_product(fx, mode, dtype=torch.float64, theta=None):
    from nonmyopia_b200 import ts_engine as te
    d, Hh, L, N = fx["d"], fx["H"], fx["L"], fx["n_r"]
    path = _product_path(fx, dtype)
    eng = te.TSRollout(path, d, Hh, L, N, fx["cost"], dtype=dtype)
    if theta is None:
        theta = te.flatten_params(fx["policy"]).to("cuda", dtype)
    theta_t = eng.transpose(theta)
    g = fx["out"][mode]
    rq, rd, cn = pack_draws(g["rot"], g["cost_noise"], L, N, dtype)
    buf = te.RolloutBuffers(d, Hh, L, N, dtype, "cuda", with_save=True)
    prob = eng.problem(fx["prev_X"].to("cuda", dtype).contiguous(), fx["prev_y"].reshape(-1).to("cuda", dtype).contiguous(),
                       fx["prev_h"].to("cuda", dtype).contiguous())
    ro = eng.rollout_struct(buf, theta, theta_t, noisy_design=(mode == "orig"), rot_q=rq, rot_dims=rd, cost_noise=cn)
    eng.forward(prob, [ro])
    g_loss = torch.full((N,), 1.0 / N, dtype=dtype, device="cuda")
    g_cost = -torch.sign(fx["budget"] - buf.cost) / N
    eng.backward(prob, ro, g_loss, g_cost)
    grad = eng.weight_grad(prob, buf)
    torch.cuda.synchronize()
    return eng, buf, grad


@pytest.mark.parametrize("name", TS_AM)
@pytest.mark.parametrize("mode", ["main", "orig"])
def test_rollout_matches_reference_golden(name, mode):
    from nonmyopia_b200 import ts_engine as te
    fx = H.load(name)
    g = fx["out"][mode]
    eng, buf, grad = run_product(fx, mode)
    L, N, d = fx["L"], fx["n_r"], fx["d"]
    # north_star tolerance: 1e-8 relative in FP64
    assert H.rel_err(buf.loss, g["acqf_loss"], atol=1e-2) < 1e-8
    assert H.rel_err(buf.cost, g["acqf_cost"], atol=1e-3) < 1e-8
    for l in range(L):
        assert H.rel_err(buf.X[l], g["X"][l].reshape(N, d), atol=1e-3) < 1e-8
        assert H.rel_err(buf.HS[l + 1], g["hidden_state"][l], atol=1e-3) < 1e-8
    assert H.rel_err(buf.A, g["actions"].reshape(N, d), atol=1e-3) < 1e-8
    gd = te.unflatten_params(grad.cpu(), d, fx["H"])
    for k, ref in g["grads"].items():
        scale = max(float(ref.abs().max()), 1e-6)
        err = float((gd[k] - ref).abs().max()) / scale
        assert err < 1e-8, (k, err)


@pytest.mark.parametrize("n_r,L,Hd", [(70, 2, 64), (33, 4, 16), (257, 1, 32)])
def test_rollout_vs_oracle_ragged_tiles(n_r, L, Hd):
    """Sizes that do not fill the 32-chain tiles, several hidden sizes; oracle = hes_oracle on CPU."""
    from oracle import hes_oracle as ho
    from nonmyopia_b200 import ts_engine as te
    fx = H.load("ts_am_ackley_L3")
    d = fx["d"]
    gen = torch.Generator().manual_seed(n_r)
    P = ho.init_policy_params(d, Hd, generator=gen)
    fx = dict(fx, H=Hd, L=L, n_r=n_r, policy=P,
              prev_X=fx["prev_X"][:1].expand(n_r, -1).clone(), prev_y=fx["prev_y"][:1].expand(n_r, -1).clone(),
              prev_h=torch.randn(n_r, Hd, dtype=torch.float64, generator=gen))
    spec = ho.HESSpec(L, [1] * L, 1, H.cost_spec(fx["cost"]), enable_ts=True)
    path = H.oracle_path(fx)
    Pq = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    torch.manual_seed(7)
    o = ho.hes_forward(spec, fx["prev_X"], fx["prev_y"], fx["prev_h"], Pq, path=path, enable_noise=True)
    tot = (o["acqf_loss"] + (fx["budget"] - o["acqf_cost"]).abs()).mean()
    gref = dict(zip(Pq.keys(), torch.autograd.grad(tot, list(Pq.values()))))
    fx["out"] = {"orig": dict(rot=o["draws"].rot, cost_noise=o["draws"].cost_noise)}
    eng, buf, grad = run_product(fx, "orig")
    assert H.rel_err(buf.loss, o["acqf_loss"], atol=1e-2) < 1e-8
    assert H.rel_err(buf.cost, o["acqf_cost"], atol=1e-3) < 1e-8
    gd = te.unflatten_params(grad.cpu(), d, Hd)
    for k, ref in gref.items():
        scale = max(float(ref.abs().max()), 1e-6)
        assert float((gd[k] - ref).abs().max()) / scale < 1e-8, k


def test_rollout_fp32_variant_tolerance():
    fx = H.load("ts_am_cosine8_L3")
    g = fx["out"]["main"]
    eng, buf, grad = run_product(fx, "main", dtype=torch.float32)
    # FP32/TF32 variant: 1e-3 relative (north_star), relative to max(|ref|, sqrt(os)=1)
    assert H.rel_err(buf.loss, g["acqf_loss"], atol=1.0) < 1e-3
    assert H.rel_err(buf.cost, g["acqf_cost"], atol=1.0) < 1e-3
    assert H.rel_err(buf.A, g["actions"].reshape(fx["n_r"], fx["d"]), atol=1.0) < 1e-3


@pytest.mark.parametrize("d,Hd,n_r,L,kernel", [(3, 16, 9, 2, "rbf"), (3, 24, 40, 2, "matern52"), (5, 64, 65, 1, "matern32"),
                                               (6, 32, 17, 3, "matern12"), (12, 64, 33, 2, "matern52")])
def test_rollout_other_dims_vs_oracle(d, Hd, n_r, L, kernel):
    """Input dimensions beyond the fixtures' 2 and 8 (odd d takes the non-tensor-pipe backward; d > 8 the 16-wide
    template), all four kernels, hidden sizes that are not 64: a synthetic GP + Thompson path built by the oracle
    (d = 1 is not a valid input of the reference either: its rotation draws a coordinate PAIR)."""
    from oracle import gp_math, hes_oracle as ho
    from nonmyopia_b200 import ts_engine as te
    gen = torch.Generator().manual_seed(100 * d + Hd)
    n = 23
    X = torch.rand(n, d, dtype=torch.float64, generator=gen)
    Y = torch.randn(n, 1, dtype=torch.float64, generator=gen)
    ls = torch.full((1, d), 0.35, dtype=torch.float64)
    gp = gp_math.DenseGP.fit_caches(X, Y, kind=kernel, lengthscale=ls, outputscale=1.2, noise=1e-2, mean_const=0.05)
    path = gp_math.draw_matheron_path(gp, Y, 1, generator=gen)
    P = ho.init_policy_params(d, Hd, generator=gen)
    cost = dict(radius=0.1, k=1.0, max_noise=1e-5, p_norm=2.0, discount=0.0, discount_threshold=-1.0)
    fx = dict(d=d, H=Hd, L=L, n_r=n_r, policy=P, cost=cost, budget=2.0,
              gp=dict(train_X=X, train_Y=Y, kernel=kernel, lengthscale=0.35, outputscale=1.2, noise=1e-2, mean_const=0.05, ard=True),
              path=dict(W=path.W, w=path.w, v=path.v),
              prev_X=X[-1:].expand(n_r, -1).clone(), prev_y=Y[-1:].expand(n_r, -1).clone(),
              prev_h=torch.randn(n_r, Hd, dtype=torch.float64, generator=gen))
    spec = ho.HESSpec(L, [1] * L, 1, H.cost_spec(cost), enable_ts=True)
    Pq = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    torch.manual_seed(11)
    o = ho.hes_forward(spec, fx["prev_X"], fx["prev_y"], fx["prev_h"], Pq, path=path, enable_noise=True)
    tot = (o["acqf_loss"] + (fx["budget"] - o["acqf_cost"]).abs()).mean()
    gref = dict(zip(Pq.keys(), torch.autograd.grad(tot, list(Pq.values()))))
    fx["out"] = {"orig": dict(rot=o["draws"].rot, cost_noise=o["draws"].cost_noise)}
    eng, buf, grad = run_product(fx, "orig")
    assert H.rel_err(buf.loss, o["acqf_loss"], atol=1e-2) < 1e-8
    assert H.rel_err(buf.cost, o["acqf_cost"], atol=1e-3) < 1e-8
    assert H.rel_err(buf.A, o["actions"].reshape(n_r, d), atol=1e-3) < 1e-8
    gd = te.unflatten_params(grad.cpu(), d, Hd)
    for k, ref in gref.items():
        scale = max(float(ref.abs().max()), 1e-6)
        assert float((gd[k] - ref).abs().max()) / scale < 1e-8, k
