"""GPU parity: nmb_path_eval (value + input gradient) vs the oracle's Matheron path."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu

CASES = ["ts_am_ackley_L3", "ts_am_cosine8_L3", "ts_am_alpine_m32", "ts_am_ackley_rbf_disc"]


def _product_path(fx, dtype=torch.float64, device="cuda"):
    from nonmyopia_b200.path import ThompsonPath
    g, p = fx["gp"], fx["path"]
    d = g["train_X"].shape[-1]
    ls = torch.full((d,), g["lengthscale"], dtype=torch.float64)
    return ThompsonPath(p["W"].to(device, dtype), p["w"].to(device, dtype), p["v"].to(device, dtype),
                        g["train_X"].to(device, dtype), ls.to(device, dtype), kernel=g["kernel"],
                        outputscale=g["outputscale"], mean_const=g["mean_const"])


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("n_pts", [1, 7, 64, 3000])
def test_path_eval_fp64(name, n_pts):
    fx = H.load(name)
    ref = H.oracle_path(fx)
    dut = _product_path(fx)
    g = torch.Generator().manual_seed(n_pts)
    d = fx["d"]
    x = torch.rand(n_pts, 1, d, dtype=torch.float64, generator=g)
    x[0, 0] = fx["gp"]["train_X"][-1]          # exactly on a training point (r = 0)
    xr = x.clone().requires_grad_(True)
    yr = ref(xr)
    (gr,) = torch.autograd.grad(yr.sum(), xr)
    xd = x.cuda().requires_grad_(True)
    yd = dut(xd)
    (gd,) = torch.autograd.grad(yd.sum(), xd)
    assert yd.shape == yr.shape
    # tolerance of north_star: 1e-8 relative in FP64 (relative to max(|ref|, atol): f crosses zero)
    assert H.rel_err(yd, yr, atol=1e-2) < 1e-8
    assert H.rel_err(gd, gr, atol=1e-1) < 1e-8


@pytest.mark.parametrize("name", CASES[:2])
def test_path_eval_fp32_variant(name):
    fx = H.load(name)
    ref = H.oracle_path(fx)
    dut = _product_path(fx, torch.float32)
    g = torch.Generator().manual_seed(5)
    x = torch.rand(512, 1, fx["d"], dtype=torch.float64, generator=g)
    yr = ref(x)
    yd = dut(x.cuda().float())
    assert yd.dtype == torch.float32
    # FP32 variant: 1e-3 relative to max(|ref|, sqrt(os)) (f crosses zero)
    assert H.rel_err(yd, yr, atol=1.0) < 1e-3


def test_multi_path_batch_matches_single_paths():
    fx = H.load("ts_am_ackley_L3")
    from nonmyopia_b200.path import ThompsonPath
    g, p = fx["gp"], fx["path"]
    gen = torch.Generator().manual_seed(3)
    S = 5
    w = torch.randn(S, p["w"].shape[-1], dtype=torch.float64, generator=gen)
    v = torch.randn(S, p["v"].shape[-1], dtype=torch.float64, generator=gen)
    ls = torch.full((2,), g["lengthscale"], dtype=torch.float64)
    mk = lambda ww, vv: ThompsonPath(p["W"].cuda(), ww.cuda(), vv.cuda(), g["train_X"].cuda(), ls.cuda(),
                                     kernel=g["kernel"], outputscale=g["outputscale"], mean_const=g["mean_const"])
    x = torch.rand(33, 1, 2, dtype=torch.float64, generator=gen).cuda()
    ys = mk(w, v)(x)
    assert ys.shape == (S, 33, 1)
    for s in range(S):
        assert torch.equal(ys[s], mk(w[s:s + 1], v[s:s + 1])(x))


def test_no_cpu_fallback():
    fx = H.load("ts_am_ackley_L3")
    dut = _product_path(fx, device="cpu")
    from nonmyopia_b200._lib import NmbError
    with pytest.raises(NmbError):
        dut(torch.rand(4, 1, 2, dtype=torch.float64))
