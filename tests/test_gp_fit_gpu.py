"""GPU parity: exact-MLL hyper-parameter fit (reference main.py:472-486 -> BoTorch fit_gpytorch_mll)
on the B200 kernels vs the CPU oracle ``oracle/gp_fit.py`` (torch autograd over the same objective)."""
import numpy as np
import pytest
import torch

from nonmyopia_b200.synthetic import make_problem
from oracle import gp_fit as ofit

pytestmark = pytest.mark.gpu
F64 = torch.float64

CASES = [("Ackley", None, "matern52", True, True, True), ("Ackley", None, "matern32", False, False, True),
         ("Hartmann", 150, "matern52", True, True, True), ("Cosine8", 120, "rbf", True, True, False),
         ("Alpine", None, "matern12", False, False, True)]


def _data(env, n):
    p = make_problem(env, seed=2)
    X, Y = p.train_X, p.train_Y
    return (X[-n:], Y[-n:]) if n else (X, Y)


@pytest.mark.parametrize("env,n,kernel,ard,scaled,priors", CASES)
def test_mll_value_and_gradient_match_oracle_autograd(env, n, kernel, ard, scaled, priors):
    from nonmyopia_b200.gp_fit import ExactMLL
    X, Y = _data(env, n)
    mll = ExactMLL(X, Y, kernel=kernel, ard=ard, scaled=scaled, priors=priors)
    g = torch.Generator().manual_seed(7)
    for trial in range(4):
        raw = (torch.randn(mll.n_raw, dtype=F64, generator=g) * (0.0 if trial == 0 else 0.8))
        raw[-2] -= 2.0 * (trial > 0)                       # smaller noise: worse conditioned
        r = raw.clone().requires_grad_(True)
        f_ref = ofit.objective(r, X.double(), Y.double().reshape(-1), kernel, ard, scaled, priors)
        (g_ref,) = torch.autograd.grad(f_ref, r)
        f, grad = mll(raw.numpy())
        assert abs(f - float(f_ref)) <= 1e-9 * max(1.0, abs(float(f_ref))), (trial, f, float(f_ref))
        scale = max(float(g_ref.abs().max()), 1e-3)
        assert float(np.abs(grad - g_ref.numpy()).max()) / scale < 1e-8, (trial, grad, g_ref)


@pytest.mark.parametrize("env,n,kernel,ard,scaled", [("Ackley", None, "matern52", True, True),
                                                     ("Hartmann", 150, "matern52", True, True),
                                                     ("Ackley", None, "matern32", False, False)])
def test_fit_reaches_the_oracle_optimum(env, n, kernel, ard, scaled):
    from nonmyopia_b200.gp_fit import fit_exact_mll
    X, Y = _data(env, n)
    h_ref, raw_ref, res_ref = ofit.fit(X, Y, kind=kernel, ard=ard, scaled=scaled)
    gp, raw, res = fit_exact_mll(X, Y, kernel=kernel, ard=ard, scaled=scaled)
    assert res.success and abs(res.fun - res_ref.fun) < 1e-8
    assert np.abs(raw - raw_ref.numpy()).max() < 1e-4
    ls_ref = h_ref["lengthscale"].numpy()
    assert np.abs(gp.lengthscale.cpu().numpy() - np.broadcast_to(ls_ref, (X.shape[1],))).max() < 1e-4
    assert abs(gp.noise - float(h_ref["noise"])) < 1e-6 and abs(gp.mean_const - float(h_ref["mean_const"])) < 1e-4
    # warm start (the reference re-uses its likelihood across BO iterations): converges at once from the optimum
    _, _, res2 = fit_exact_mll(X, Y, kernel=kernel, ard=ard, scaled=scaled, raw0=raw)
    assert res2.nit <= 2


def test_fit_refuses_cpu():
    from nonmyopia_b200.gp_fit import ExactMLL
    X, Y = _data("Ackley", None)
    with pytest.raises(Exception):
        ExactMLL(X, Y, device="cpu")
