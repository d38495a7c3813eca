"""Exact-GP surrogate state on the device + Thompson-path draw (host side).

Host-side mirror of the slice of ``SingleTaskGP`` the lookahead path consumes
(reference main.py:479-486 builds the model; acqfs.py:258-266 ``dump_model`` draws the
Thompson path; acqfs.py:364-373,413-415 call posterior / condition_on_observations).
The acquisition kernels never see a BoTorch object: ``SurrogateGP`` extracts training
data and hyper-parameters once per BO step, builds ``K + s2 I`` with the
``nmb_kernel_matrix`` kernel, factorises it once (n <= 800 = GPyTorch's
max_cholesky_size) and keeps the caches the kernels stream:

* ``L``     lower Cholesky root of ``K + s2 I``
* ``Linv``  explicit inverse root (GPyTorch's covar cache is ``R = L^-T``; ``Linv = R^T``)
* ``alpha`` mean cache ``(K + s2 I)^-1 (y - c)``
* ``beta``  whitened targets ``L^-1 (y - c)`` (``mu(x) = c + u(x) . beta`` with ``u = L^-1 k``)

The once-per-BO-step factorisation and the inverse root are one launch of the ``nmb_cholesky_root``
kernel (csrc/chol.cu: one CTA per matrix, blocked, panel in shared memory).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch
from torch import Tensor

from . import _lib
from .path import ThompsonPath

_NU_NAME = {0.5: "matern12", 1.5: "matern32", 2.5: "matern52"}


def kernel_matrix(x1: Tensor, x2: Tensor, kernel: str, lengthscale: Tensor, outputscale: Optional[float],
                  diag_add: float = 0.0) -> Tensor:
    """``amp * kappa(x1, x2)`` [n1, n2] by the ``nmb_kernel_matrix`` kernel."""
    n1, d = x1.shape
    n2 = x2.shape[0]
    x1, x2 = x1.contiguous(), x2.contiguous()
    inv_ls = (1.0 / lengthscale.to(x1)).reshape(-1)
    if inv_ls.numel() == 1:
        inv_ls = inv_ls.expand(d)
    inv_ls = inv_ls.contiguous()
    K = torch.empty(n1, n2, dtype=x1.dtype, device=x1.device)
    amp = 1.0 if outputscale is None else float(outputscale)
    _lib.check(_lib.lib().nmb_kernel_matrix(
        _lib.dtype_code(x1.dtype), _lib.KERNEL_IDS[kernel], d, _lib.ptr(x1), C.c_int64(n1), _lib.ptr(x2),
        C.c_int64(n2), _lib.ptr(inv_ls), C.c_double(amp), C.c_double(diag_add), _lib.ptr(K), C.c_int64(n2),
        _lib.stream_ptr()), "nmb_kernel_matrix")
    return K


def cholesky_root(A: Tensor, inverse: bool = True):
    """Lower Cholesky factor (and its explicit inverse) of ``A [..., n, n]`` by the ``nmb_cholesky_root`` kernel.
    Returns ``(L, Linv or None, info [batch] int32)``; ``A`` is not modified."""
    n = A.shape[-1]
    L = A.detach().reshape(-1, n, n).clone().contiguous()
    batch = L.shape[0]
    Linv = torch.empty_like(L) if inverse else None
    info = torch.zeros(batch, dtype=torch.int32, device=A.device)
    _lib.check(_lib.lib().nmb_cholesky_root(
        _lib.dtype_code(A.dtype), n, batch, _lib.ptr(L), C.c_int64(n), C.c_int64(n * n), _lib.ptr(Linv), C.c_int64(n),
        C.c_int64(n * n), C.c_void_p(info.data_ptr()), _lib.stream_ptr()), "nmb_cholesky_root")
    return L.reshape(A.shape), (Linv.reshape(A.shape) if inverse else None), info


def _psd_safe_cholesky(A: Tensor):
    """GPyTorch's jitter rule (linear_operator psd_safe_cholesky) around ``nmb_cholesky_root``.
    Returns ``(L, L^-1)``."""
    L, Linv, info = cholesky_root(A)
    if not bool(info.any()):
        return L, Linv
    jitter = 1e-6 if A.dtype == torch.float32 else 1e-8
    prev = 0.0
    Ap = A.clone()
    for i in range(3):
        new = jitter * 10**i
        Ap.diagonal().add_(new - prev)
        prev = new
        L, Linv, info = cholesky_root(Ap)
        if not bool(info.any()):
            return L, Linv
    raise RuntimeError("Matrix not positive definite after repeatedly adding jitter")


class SurrogateGP:
    """Exact GP with constant mean and a stationary kernel; hyper-parameters are constants
    of the acquisition optimisation (the reference detaches the caches)."""

    def __init__(self, train_X: Tensor, train_Y: Tensor, *, kernel: str, lengthscale, outputscale: Optional[float],
                 noise: float, mean_const: float = 0.0, device="cuda", dtype=torch.float64):
        if kernel not in _lib.KERNEL_IDS:
            raise NotImplementedError(f"kernel {kernel!r} (supported: {sorted(_lib.KERNEL_IDS)})")
        dev = torch.device(device)
        self.kernel = kernel
        self.train_X = train_X.detach().to(dev, dtype).contiguous()
        self.train_Y = train_Y.detach().to(dev, dtype).reshape(-1).contiguous()
        n, d = self.train_X.shape
        ls = torch.as_tensor(lengthscale, dtype=dtype).reshape(-1).to(dev)
        self.ard = ls.numel() > 1
        self.lengthscale = (ls if ls.numel() == d else ls.expand(d)).contiguous()
        self.outputscale = None if outputscale is None else float(outputscale)
        self.noise = float(noise)
        self.mean_const = float(mean_const)
        Khat = kernel_matrix(self.train_X, self.train_X, kernel, self.lengthscale, self.outputscale, self.noise)
        self.L, self.Linv = _psd_safe_cholesky(Khat)
        resid = (self.train_Y - self.mean_const).unsqueeze(-1)
        # GPyTorch multiplies by the explicit inverse root (k R, R = L^-T): whiten the same way
        self.beta = (self.Linv @ resid).squeeze(-1).contiguous()
        self.alpha = (self.Linv.T @ self.beta.unsqueeze(-1)).squeeze(-1).contiguous()

    # ---- views
    @property
    def n(self) -> int:
        return self.train_X.shape[0]

    @property
    def d(self) -> int:
        return self.train_X.shape[1]

    @property
    def amplitude(self) -> float:
        return 1.0 if self.outputscale is None else self.outputscale

    @property
    def dtype(self):
        return self.train_X.dtype

    @property
    def device(self):
        return self.train_X.device

    def k(self, x1: Tensor, x2: Tensor) -> Tensor:
        return kernel_matrix(x1, x2, self.kernel, self.lengthscale, self.outputscale)

    # ---- construction from a BoTorch-like model (duck typed; BoTorch is not a dependency)
    @classmethod
    def from_model(cls, model, device=None, dtype=None) -> "SurrogateGP":
        if isinstance(model, SurrogateGP):
            return model
        X = model.train_inputs[0]
        y = model.train_targets
        if X.dim() != 2:
            raise NotImplementedError("batched training data is not supported on this path")
        cm = model.covar_module
        outputscale = None
        if hasattr(cm, "base_kernel"):            # ScaleKernel(base)
            outputscale = float(torch.as_tensor(cm.outputscale).reshape(-1)[0])
            cm = cm.base_kernel
        name = type(cm).__name__
        if name == "RBFKernel":
            kernel = "rbf"
        elif name == "MaternKernel":
            kernel = _NU_NAME[float(cm.nu)]
        else:
            raise NotImplementedError(f"covariance module {name} (the reference's --kernel Linear has no "
                                      "pathwise sampler upstream either, SURVEY A.6)")
        ls = torch.as_tensor(cm.lengthscale).detach().reshape(-1)
        noise = float(torch.as_tensor(model.likelihood.noise).reshape(-1)[0])
        mean_mod = getattr(model, "mean_module", None)
        c = getattr(mean_mod, "constant", 0.0) if mean_mod is not None else 0.0
        c = float(torch.as_tensor(c).reshape(-1)[0]) if c is not None else 0.0
        return cls(X, y, kernel=kernel, lengthscale=ls, outputscale=outputscale, noise=noise, mean_const=c,
                   device=device or (X.device if X.is_cuda else "cuda"), dtype=dtype or X.dtype)


def sobol_normal_base_samples(n: int, q: int, seed: int, dtype=torch.float64) -> Tensor:
    """``draw_sobol_normal_samples`` of BoTorch 0.10.0: host-side, [n, q]."""
    engine = torch.quasirandom.SobolEngine(dimension=q, scramble=True, seed=seed)
    u = engine.draw(n, dtype=dtype)
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    return torch.erfinv(2 * v - 1) * math.sqrt(2)


def draw_thompson_path(gp: SurrogateGP, n_paths: int = 1, num_features: int = 1024,
                       generator: Optional[torch.Generator] = None) -> ThompsonPath:
    """``draw_matheron_paths(model, Size([n_paths]))`` (reference acqfs.py:264).

    Random numbers are drawn on the host in BoTorch's order (Sobol scramble seeds from the
    global generator, Gamma mixing variable for Matern spectra, ``randn`` noise sample) and
    the arithmetic - feature matrix at the training points, pathwise update solve - runs on
    the device.
    """
    n, d, dtype, dev = gp.n, gp.d, gp.dtype, gp.device
    Mh = num_features // 2
    seed_w = int(torch.randint(0, 1_000_000, (1,), generator=generator).item())
    W = sobol_normal_base_samples(Mh, d, seed_w, dtype)
    if gp.kernel != "rbf":
        nu = {"matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[gp.kernel]
        g = torch.distributions.Gamma(nu, nu).sample((Mh, 1)).to(dtype)
        W = W / g.sqrt()
    seed_f = int(torch.randint(0, 1_000_000, (1,), generator=generator).item())
    w = sobol_normal_base_samples(n_paths, num_features, seed_f, dtype)
    eps = torch.randn(n_paths, n, dtype=dtype, generator=generator) * math.sqrt(gp.noise)
    W, w, eps = W.to(dev), w.to(dev), eps.to(dev)
    z = (gp.train_X / gp.lengthscale) @ W.T                               # [n, M/2]
    scale = math.sqrt(gp.amplitude * 2.0 / num_features)
    phi = scale * torch.cat([z.sin(), z.cos()], dim=-1)                  # [n, M]
    prior_at_train = gp.mean_const + (phi @ w.T).T                       # [S, n]
    resid = gp.train_Y.unsqueeze(0) - (prior_at_train + eps)
    v = ((resid @ gp.Linv.T) @ gp.Linv).contiguous()                      # [S, n] = (K + s2 I)^-1 resid, row-wise
    ls = gp.lengthscale if gp.ard else gp.lengthscale[:1]
    return ThompsonPath(W, w, v, gp.train_X, ls, kernel=gp.kernel, outputscale=gp.outputscale,
                        mean_const=gp.mean_const)
