"""CPU oracle of the GP hyper-parameter fit: what ``fit_gpytorch_mll(ExactMarginalLogLikelihood(
likelihood, SingleTaskGP(X, Y, likelihood, covar_module)))`` minimises in the reference's BO loop (main.py:472-486).

TEST INFRASTRUCTURE ONLY.  BoTorch 0.10.0 / GPyTorch 1.11 are not installable here, so this restates their published
algorithm [3P, parity unpinned against the real packages]:

* model: constant mean ``c``; ``covar_module=None`` => ``ScaleKernel(MaternKernel(2.5, ard_num_dims=d))`` with
  ``lengthscale ~ Gamma(3, 6)`` and ``outputscale ~ Gamma(2, 0.15)`` priors (the reference's ``--kernel RBF`` maps
  to ``None``, main.py:84-85, i.e. to this Matern default); ``--kernel Matern-nu`` is a bare isotropic
  ``MaternKernel(nu)`` without priors; likelihood ``GaussianLikelihood(noise_prior=NormalPrior(0, 1e-2))`` with the
  default ``GreaterThan(1e-4)`` noise constraint (main.py:472-474);
* raw parameters through softplus (``Positive`` / ``GreaterThan``): ``ls = softplus(raw)``, ``os = softplus(raw)``,
  ``noise = 1e-4 + softplus(raw)``; the mean constant is unconstrained; all raw values start at 0 unless warm-started;
* objective: ``-(log N(y; c 1, K + noise I) + sum of log-priors) / n`` (``ExactMarginalLogLikelihood.forward`` adds
  the priors and divides by the number of data), minimised with scipy L-BFGS-B on the raw vector (BoTorch
  ``scipy_minimize`` defaults).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F
from torch import Tensor

from . import gp_math

NOISE_FLOOR = 1e-4


def unpack(raw: Tensor, d: int, ard: bool, scaled: bool) -> Dict[str, Tensor]:
    """raw = [raw_ls (d or 1), raw_os (if scaled), raw_noise, c] -> constrained hyper-parameters."""
    n_ls = d if ard else 1
    o = 0
    ls = F.softplus(raw[o:o + n_ls]); o += n_ls
    os_ = F.softplus(raw[o]) if scaled else None
    o += 1 if scaled else 0
    noise = NOISE_FLOOR + F.softplus(raw[o]); o += 1
    return dict(lengthscale=ls, outputscale=os_, noise=noise, mean_const=raw[o])


def n_raw(d: int, ard: bool, scaled: bool) -> int:
    return (d if ard else 1) + (1 if scaled else 0) + 2


def objective(raw: Tensor, X: Tensor, y: Tensor, kind: str, ard: bool, scaled: bool, priors: bool) -> Tensor:
    n, d = X.shape
    h = unpack(raw, d, ard, scaled)
    K = gp_math.kernel_matrix(X, X, kind, h["lengthscale"].reshape(1, -1), h["outputscale"])
    Khat = K + h["noise"] * torch.eye(n, dtype=X.dtype)
    L = torch.linalg.cholesky(Khat)
    r = (y - h["mean_const"]).unsqueeze(-1)
    a = torch.cholesky_solve(r, L)
    logp = -0.5 * (r * a).sum() - torch.log(torch.diagonal(L)).sum() - 0.5 * n * math.log(2 * math.pi)
    if priors:
        if scaled:
            ls, os_ = h["lengthscale"], h["outputscale"]
            logp = logp + (3.0 * math.log(6.0) - math.lgamma(3.0) + 2.0 * torch.log(ls) - 6.0 * ls).sum()
            logp = logp + (2.0 * math.log(0.15) - math.lgamma(2.0) + torch.log(os_) - 0.15 * os_)
        nz = h["noise"]
        logp = logp + (-0.5 * (nz / 1e-2) ** 2 - math.log(1e-2) - 0.5 * math.log(2 * math.pi))
    return -logp / n


def fit(X: Tensor, y: Tensor, *, kind: str = "matern52", ard: bool = True, scaled: bool = True, priors: bool = True,
        raw0: Optional[Tensor] = None, maxiter: int = 15000):
    """L-BFGS-B over the raw vector; returns (constrained hypers, raw optimum, scipy result)."""
    from scipy.optimize import minimize
    X, y = X.detach().double(), y.detach().double().reshape(-1)
    d = X.shape[1]
    x0 = np.zeros(n_raw(d, ard, scaled)) if raw0 is None else raw0.detach().double().numpy().copy()

    def fun(v):
        raw = torch.tensor(v, dtype=torch.float64, requires_grad=True)
        f = objective(raw, X, y, kind, ard, scaled, priors)
        (g,) = torch.autograd.grad(f, raw)
        return float(f), g.numpy().astype(np.float64)

    res = minimize(fun, x0, jac=True, method="L-BFGS-B", options=dict(maxiter=maxiter))
    raw = torch.tensor(res.x, dtype=torch.float64)
    h = {k: (None if v is None else v.detach()) for k, v in unpack(raw, d, ard, scaled).items()}
    return h, raw, res
