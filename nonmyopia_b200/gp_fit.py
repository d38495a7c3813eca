"""GP hyper-parameter fit on the B200 kernels: the step of the reference's BO loop in front of
the acquisition path, ``fit_gpytorch_mll(ExactMarginalLogLikelihood(likelihood, SingleTaskGP(X, Y, likelihood,
covar_module)))`` (main.py:472-486), restated from BoTorch 0.10.0 / GPyTorch 1.11 [3P]:

* constant mean; ``covar_module=None`` => ``ScaleKernel(MaternKernel(2.5, ard))`` with ``lengthscale ~ Gamma(3, 6)``,
  ``outputscale ~ Gamma(2, 0.15)`` (the reference's ``--kernel RBF`` maps to ``None``, main.py:84-85);
  ``--kernel Matern-nu`` => bare isotropic ``MaternKernel(nu)``; ``GaussianLikelihood(noise_prior=NormalPrior(0, 1e-2))``
  with the ``GreaterThan(1e-4)`` constraint;
* raw parameters through softplus, all zero at the start unless warm-started (the reference re-uses its likelihood
  object across BO iterations, so the noise is warm-started; pass ``raw0``);
* objective ``-(log p(y) + log-priors) / n`` minimised by scipy L-BFGS-B on the raw vector (BoTorch's defaults).

Every evaluation runs on the device: ``nmb_kernel_matrix`` -> ``nmb_cholesky_root`` (GPyTorch's jitter rule) ->
``nmb_gp_mll_value_grad``; the host reads back 5 + d doubles, adds the priors and the softplus chain rule and hands
value + gradient to L-BFGS-B.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import numpy as np
import torch
from torch import Tensor

from . import _lib
from .gp import SurrogateGP, _psd_safe_cholesky, kernel_matrix

NOISE_FLOOR = 1e-4


def _softplus(v):
    return np.logaddexp(0.0, v)


def _sigmoid(v):
    return 1.0 / (1.0 + np.exp(-v))


class ExactMLL:
    """Value and gradient of the scaled negative marginal log-likelihood (+ priors) at a raw parameter vector
    ``[raw_lengthscale (d or 1), raw_outputscale (if scaled), raw_noise, mean_const]``."""

    def __init__(self, train_X: Tensor, train_Y: Tensor, *, kernel: str = "matern52", ard: bool = True,
                 scaled: bool = True, priors: bool = True, device="cuda"):
        if kernel not in _lib.KERNEL_IDS:
            raise NotImplementedError(f"kernel {kernel!r} (supported: {sorted(_lib.KERNEL_IDS)})")
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.NmbError("nonmyopia_b200 ops need CUDA tensors (no CPU fallback)")
        self.X = train_X.detach().to(dev, torch.float64).contiguous()
        self.y = train_Y.detach().to(dev, torch.float64).reshape(-1).contiguous()
        self.n, self.d = self.X.shape
        self.kernel, self.ard, self.scaled, self.priors = kernel, bool(ard), bool(scaled), bool(priors)
        self.n_ls = self.d if ard else 1
        self.n_raw = self.n_ls + (1 if scaled else 0) + 2
        lib = _lib.lib()
        lib.nmb_gp_mll_workspace_elems.restype = C.c_int64
        self.lib = lib
        self.work = torch.empty(int(lib.nmb_gp_mll_workspace_elems(self.n, self.d)), dtype=torch.float64, device=dev)
        self.out = torch.empty(5 + self.d, dtype=torch.float64, device=dev)
        self.n_eval = 0

    def unpack(self, raw: np.ndarray) -> dict:
        o = self.n_ls
        ls = _softplus(raw[:o])
        os_ = float(_softplus(raw[o])) if self.scaled else None
        o += 1 if self.scaled else 0
        return dict(lengthscale=ls, outputscale=os_, noise=NOISE_FLOOR + float(_softplus(raw[o])),
                    mean_const=float(raw[o + 1]))

    def __call__(self, raw: np.ndarray):
        raw = np.asarray(raw, dtype=np.float64)
        h = self.unpack(raw)
        n, d = self.n, self.d
        ls = torch.tensor(np.broadcast_to(h["lengthscale"], (d,)).copy(), dtype=torch.float64, device=self.X.device)
        amp = 1.0 if h["outputscale"] is None else h["outputscale"]
        Khat = kernel_matrix(self.X, self.X, self.kernel, ls, h["outputscale"], h["noise"])
        L, Linv = _psd_safe_cholesky(Khat)
        inv_ls = (1.0 / ls).contiguous()
        dp = lambda t: C.cast(_lib.ptr(t), C.POINTER(C.c_double))
        _lib.check(self.lib.nmb_gp_mll_value_grad(
            _lib.KERNEL_IDS[self.kernel], d, n, dp(self.X), dp(inv_ls), dp(self.y), C.c_double(h["mean_const"]), dp(L),
            C.c_int64(n), dp(Linv), C.c_int64(n), dp(self.work), dp(self.out), _lib.stream_ptr()), "nmb_gp_mll_value_grad")
        o = self.out.cpu().numpy()
        self.n_eval += 1
        quad, logdet, sum_alpha, tr_g, tr_gk = o[0], o[1], o[2], o[3], o[4]
        tr_gl = o[5:5 + d]
        logp = -0.5 * quad - 0.5 * logdet - 0.5 * n * math.log(2 * math.pi)
        # gradient of -log p w.r.t. the constrained parameters
        g_ls = -0.5 * amp * tr_gl
        if not self.ard:
            g_ls = np.array([g_ls.sum()])
        g_os = -0.5 * tr_gk
        g_noise = -0.5 * tr_g
        g_c = -sum_alpha
        if self.priors:
            if self.scaled:
                l_ = h["lengthscale"]
                logp += float((3.0 * math.log(6.0) - math.lgamma(3.0) + 2.0 * np.log(l_) - 6.0 * l_).sum())
                g_ls = g_ls - (2.0 / l_ - 6.0)
                logp += 2.0 * math.log(0.15) - math.lgamma(2.0) + math.log(amp) - 0.15 * amp
                g_os = g_os - (1.0 / amp - 0.15)
            nz = h["noise"]
            logp += -0.5 * (nz / 1e-2) ** 2 - math.log(1e-2) - 0.5 * math.log(2 * math.pi)
            g_noise = g_noise + nz / 1e-4
        # softplus chain rule back to the raw vector
        grad = np.empty(self.n_raw)
        k = self.n_ls
        grad[:k] = g_ls * _sigmoid(raw[:k])
        if self.scaled:
            grad[k] = g_os * _sigmoid(raw[k])
            k += 1
        grad[k] = g_noise * _sigmoid(raw[k])
        grad[k + 1] = g_c
        return -logp / n, grad / n


def fit_exact_mll(train_X: Tensor, train_Y: Tensor, *, kernel: str = "matern52", ard: bool = True, scaled: bool = True,
                  priors: bool = True, raw0: Optional[np.ndarray] = None, maxiter: int = 15000, device="cuda"):
    """Fit and return ``(SurrogateGP, raw optimum, scipy result)``; the raw optimum warm-starts the next BO step."""
    from scipy.optimize import minimize
    mll = ExactMLL(train_X, train_Y, kernel=kernel, ard=ard, scaled=scaled, priors=priors, device=device)
    x0 = np.zeros(mll.n_raw) if raw0 is None else np.asarray(raw0, dtype=np.float64).copy()
    res = minimize(mll, x0, jac=True, method="L-BFGS-B", options=dict(maxiter=maxiter))
    h = mll.unpack(res.x)
    gp = SurrogateGP(train_X, train_Y, kernel=kernel, lengthscale=torch.tensor(h["lengthscale"], dtype=torch.float64),
                     outputscale=h["outputscale"], noise=h["noise"], mean_const=h["mean_const"], device=device)
    res.n_device_evals = mll.n_eval
    return gp, res.x, res
