"""CPU oracle, part 1: the Gaussian-process arithmetic the reference delegates to
BoTorch 0.10.0 / GPyTorch 1.11 (third-party, NOT vendored in /root/reference and
NOT installable here: no network, not in /opt/wheelhouse).

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
package ``nonmyopia_b200``; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may use it.

Parity status: **unpinned for the third-party arithmetic** (the reference repo
has no golden vectors for this path, SURVEY.md section 4 / 8c, and the pinned
BoTorch/GPyTorch wheels are absent).  The formulas below restate the published
algorithms of those versions:

* kernels: GPyTorch ``MaternKernel.forward`` / ``RBFKernel.forward`` with the
  ``sq_dist`` matmul expansion, mean-centering and ``clamp_min(1e-30).sqrt()``
  (reference call sites: synthetic_experiment/main.py:86-92, 479-484);
* exact posterior with Cholesky caches (mean cache ``alpha = Khat^-1 (y-c)``,
  covar cache ``R = L^-T``), reference call sites acqfs.py:366, 414;
* ``condition_on_observations`` / ``get_fantasy_strategy`` block update,
  reference call site acqfs.py:371-373;
* Sobol-normal base samples (``SobolQMCNormalSampler``), acqfs.py:517-519,
  actor.py:417-419;
* Matheron pathwise samples (``draw_matheron_paths``), acqfs.py:264.

Everything is dense, batched plain PyTorch on the CPU and differentiable with
autograd; it is deliberately the *naive* formulation (dense per-node roots and
inverse roots, as GPyTorch materialises them) so that it is an independent check
of the recursion the CUDA kernels use.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import Tensor

KERNEL_IDS = {"rbf": 0, "matern12": 1, "matern32": 2, "matern52": 3}


# --------------------------------------------------------------------------
# kernels (GPyTorch 1.11 kernels/{kernel,matern_kernel,rbf_kernel}.py restated)
# --------------------------------------------------------------------------
def _sq_dist(x1: Tensor, x2: Tensor, x1_eq_x2: bool) -> Tensor:
    """``gpytorch.kernels.kernel.sq_dist``: ||a||^2 + ||b||^2 - 2ab by one matmul."""
    adjustment = x1.mean(-2, keepdim=True)
    x1 = x1 - adjustment
    x2 = x2 - adjustment
    x1_norm = x1.pow(2).sum(dim=-1, keepdim=True)
    x1_pad = torch.ones_like(x1_norm)
    if x1_eq_x2 and not x1.requires_grad and not x2.requires_grad:
        x2_norm, x2_pad = x1_norm, x1_pad
    else:
        x2_norm = x2.pow(2).sum(dim=-1, keepdim=True)
        x2_pad = torch.ones_like(x2_norm)
    x1_ = torch.cat([-2.0 * x1, x1_norm, x1_pad], dim=-1)
    x2_ = torch.cat([x2, x2_pad, x2_norm], dim=-1)
    res = x1_.matmul(x2_.transpose(-2, -1))
    if x1_eq_x2 and not x1.requires_grad and not x2.requires_grad:
        res = res.clone()
        res.diagonal(dim1=-2, dim2=-1).fill_(0)
    return res.clamp_min(0)


def _dist(x1: Tensor, x2: Tensor, x1_eq_x2: bool) -> Tensor:
    return _sq_dist(x1, x2, x1_eq_x2).clamp_min(1e-30).sqrt()


def kernel_matrix(
    x1: Tensor,
    x2: Tensor,
    kind: str,
    lengthscale: Tensor,
    outputscale: Optional[float | Tensor] = None,
) -> Tensor:
    """k(x1, x2) for x1 [..., a, d], x2 [..., b, d] -> [..., a, b].

    ``lengthscale`` is [1, d] (ARD) or [1, 1]; ``outputscale`` None means a bare
    kernel (unit prior variance), otherwise ScaleKernel's multiplier.
    """
    x1_eq_x2 = x1 is x2 or (x1.shape == x2.shape and torch.equal(x1, x2))
    if kind == "rbf":
        x1_ = x1.div(lengthscale)
        x2_ = x2.div(lengthscale)
        k = torch.exp(-0.5 * _sq_dist(x1_, x2_, x1_eq_x2))
    else:
        nu = {"matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[kind]
        mean = x1.reshape(-1, x1.shape[-1]).mean(0)[(None,) * (x1.dim() - 1)]
        x1_ = (x1 - mean).div(lengthscale)
        x2_ = (x2 - mean).div(lengthscale)
        distance = _dist(x1_, x2_, x1_eq_x2)
        exp_component = torch.exp(-math.sqrt(nu * 2) * distance)
        if nu == 0.5:
            constant_component = 1.0
        elif nu == 1.5:
            constant_component = (math.sqrt(3) * distance).add(1)
        else:
            constant_component = (
                (math.sqrt(5) * distance).add(1).add(5.0 / 3.0 * distance**2)
            )
        k = constant_component * exp_component
    if outputscale is not None:
        k = k * outputscale
    return k


def kernel_diag_value(outputscale: Optional[float | Tensor]) -> float | Tensor:
    """k(x, x) for the stationary kernels above."""
    return 1.0 if outputscale is None else outputscale


# --------------------------------------------------------------------------
# Cholesky with GPyTorch's jitter rule (linear_operator psd_safe_cholesky)
# --------------------------------------------------------------------------
def psd_safe_cholesky(A: Tensor, max_tries: int = 3) -> Tensor:
    L, info = torch.linalg.cholesky_ex(A)
    if not torch.any(info):
        return L
    jitter = 1e-6 if A.dtype == torch.float32 else 1e-8
    Aprime = A.clone()
    jitter_prev = 0.0
    for i in range(max_tries):
        jitter_new = jitter * (10**i)
        diag_add = ((info > 0) * (jitter_new - jitter_prev)).unsqueeze(-1).expand(
            *A.shape[:-1]
        )
        Aprime.diagonal(dim1=-1, dim2=-2).add_(diag_add)
        jitter_prev = jitter_new
        L, info = torch.linalg.cholesky_ex(Aprime)
        if not torch.any(info):
            return L
    raise RuntimeError("Matrix not positive definite after repeatedly adding jitter")


# --------------------------------------------------------------------------
# exact GP with prediction caches (GPyTorch DefaultPredictionStrategy restated)
# --------------------------------------------------------------------------
class DenseGP:
    """Exact GP state: train inputs, Cholesky root ``L`` of ``K + s2 I``, the
    explicit inverse root ``R = L^-T`` (covar cache) and ``alpha`` (mean cache).

    All tensors may carry leading batch dims (fantasy models).  Hyper-parameters
    are constants (the acquisition optimisation never differentiates them).
    """

    def __init__(self, train_X, alpha, L, R, *, kind, lengthscale, outputscale,
                 noise, mean_const):
        self.train_X = train_X          # [*b, m, d]
        self.alpha = alpha              # [*b, m]
        self.L = L                      # [*b, m, m]
        self.R = R                      # [*b, m, m]  (= L^-T)
        self.kind = kind
        self.lengthscale = lengthscale  # [1, d] or [1, 1]
        self.outputscale = outputscale  # None or float
        self.noise = float(noise)
        self.mean_const = float(mean_const)

    # -- construction ------------------------------------------------------
    @classmethod
    def fit_caches(cls, train_X: Tensor, train_Y: Tensor, *, kind, lengthscale,
                   outputscale, noise, mean_const) -> "DenseGP":
        X = train_X.detach()
        y = train_Y.detach().reshape(-1)
        lengthscale = torch.as_tensor(lengthscale, dtype=X.dtype).reshape(1, -1)
        K = kernel_matrix(X, X, kind, lengthscale, outputscale)
        Khat = K + noise * torch.eye(X.shape[0], dtype=X.dtype)
        L = psd_safe_cholesky(Khat)
        alpha = torch.cholesky_solve((y - mean_const).unsqueeze(-1), L).squeeze(-1)
        eye = torch.eye(X.shape[0], dtype=X.dtype)
        R = torch.linalg.solve_triangular(L, eye, upper=False).transpose(-1, -2)
        return cls(X, alpha, L, R.contiguous(), kind=kind, lengthscale=lengthscale,
                   outputscale=outputscale, noise=noise, mean_const=mean_const)

    @property
    def batch_shape(self):
        return self.train_X.shape[:-2]

    def k(self, a: Tensor, b: Tensor) -> Tensor:
        return kernel_matrix(a, b, self.kind, self.lengthscale, self.outputscale)

    # -- posterior -------------------------------------------
    def posterior(self, X: Tensor, observation_noise: bool = False):
        """X [*b', q, d] (batch broadcastable against the model batch) ->
        mean [*B, q], covariance [*B, q, q]."""
        bshape = torch.broadcast_shapes(X.shape[:-2], self.batch_shape)
        Xb = X.expand(*bshape, *X.shape[-2:])
        trX = self.train_X.expand(*bshape, *self.train_X.shape[-2:])
        kstar = self.k(Xb, trX)                                   # [*B, q, m]
        mean = self.mean_const + (kstar * self.alpha.unsqueeze(-2)).sum(-1)
        kR = kstar.matmul(self.R)                                 # [*B, q, m]
        kss = self.k(Xb, Xb)
        cov = kss - kR.matmul(kR.transpose(-1, -2))
        if observation_noise:
            cov = cov + self.noise * torch.eye(cov.shape[-1], dtype=cov.dtype)
        return mean, cov

    # -- fantasy update ---------------------------------------
    def condition_on_observations(self, X: Tensor, Y: Tensor) -> "DenseGP":
        """X [*b, q, d]; Y [nf, *b, q, 1] (one more leading dim than X: several
        fantasised outcomes at the same inputs).  Returns a model with batch
        shape [nf, *b].  Gradients flow through X and Y."""
        bshape = torch.broadcast_shapes(X.shape[:-2], self.batch_shape)
        Xb = X.expand(*bshape, *X.shape[-2:])
        trX = self.train_X.expand(*bshape, *self.train_X.shape[-2:])
        L = self.L.expand(*bshape, *self.L.shape[-2:])
        R = self.R.expand(*bshape, *self.R.shape[-2:])
        alpha = self.alpha.expand(*bshape, self.alpha.shape[-1])
        q = X.shape[-2]
        m = trX.shape[-2]
        B = self.k(Xb, trX)                                       # [*b, q, m]
        C = self.k(Xb, Xb) + self.noise * torch.eye(q, dtype=X.dtype)
        V = torch.cholesky_solve(B.transpose(-1, -2), L)          # [*b, m, q]
        S = C - B.matmul(V)
        lS = psd_safe_cholesky(S)                                 # [*b, q, q]
        fant_mean = self.mean_const + (B * alpha.unsqueeze(-2)).sum(-1)   # [*b, q]
        rhs = Y.squeeze(-1) - fant_mean                           # [nf, *b, q]
        a_lo = torch.cholesky_solve(rhs.unsqueeze(-1), lS).squeeze(-1)    # [nf,*b,q]
        a_up = alpha - (V * a_lo.unsqueeze(-2)).sum(-1)           # [nf, *b, m]
        new_alpha = torch.cat([a_up, a_lo], dim=-1)               # [nf, *b, m+q]
        # new root  L' = [[L, 0], [B L^-T, lS]]
        lower_left = B.matmul(R)                                  # [*b, q, m]
        top = torch.cat([L, L.new_zeros(*bshape, m, q)], dim=-1)
        bot = torch.cat([lower_left, lS], dim=-1)
        newL = torch.cat([top, bot], dim=-2)                      # [*b, m+q, m+q]
        # new inverse root R' = L'^-T, block-wise:
        #   L'^-1 = [[L^-1, 0], [-lS^-1 (B L^-T) L^-1, lS^-1]]
        eye_q = torch.eye(q, dtype=X.dtype).expand(*bshape, q, q)
        lS_inv = torch.linalg.solve_triangular(lS, eye_q, upper=False)
        Linv = R.transpose(-1, -2)
        ll = -lS_inv.matmul(lower_left).matmul(Linv)
        top_i = torch.cat([Linv, L.new_zeros(*bshape, m, q)], dim=-1)
        bot_i = torch.cat([ll, lS_inv], dim=-1)
        newR = torch.cat([top_i, bot_i], dim=-2).transpose(-1, -2)
        new_trX = torch.cat([trX, Xb], dim=-2)                    # [*b, m+q, d]
        return DenseGP(new_trX, new_alpha, newL, newR, kind=self.kind,
                       lengthscale=self.lengthscale, outputscale=self.outputscale,
                       noise=self.noise, mean_const=self.mean_const)


# --------------------------------------------------------------------------
# samplers (BoTorch 0.10.0 sampling/{base,normal}.py + utils/sampling.py)
# --------------------------------------------------------------------------
def sobol_normal_base_samples(n: int, q: int, seed: int, dtype=torch.float64) -> Tensor:
    """``draw_sobol_normal_samples(d=q, n=n, seed=seed)`` -> [n, q]."""
    engine = torch.quasirandom.SobolEngine(dimension=q, scramble=True, seed=seed)
    u = engine.draw(n, dtype=dtype)
    v = 0.5 + (1 - torch.finfo(u.dtype).eps) * (u - 0.5)
    return torch.erfinv(2 * v - 1) * math.sqrt(2)


def sample_from_posterior(mean: Tensor, cov: Tensor, base: Tensor) -> Tensor:
    """``posterior.rsample_from_base_samples``: base [nf, 1.., q] ->
    samples [nf, *B, q, 1] = mean + chol(cov) z."""
    root = psd_safe_cholesky(cov)                                 # [*B, q, q]
    z = base.expand(base.shape[0], *mean.shape)                   # [nf, *B, q]
    out = mean + root.matmul(z.unsqueeze(-1)).squeeze(-1)
    return out.unsqueeze(-1)


# --------------------------------------------------------------------------
# Matheron pathwise posterior sample (BoTorch sampling/pathwise)
# --------------------------------------------------------------------------
class MatheronPath(torch.nn.Module):
    """f(x) = c + phi(x) . w + os * k(x, X_tr) . v  for S jointly drawn paths.

    Buffers: W [M/2, d] spectral frequencies, w [S, M] feature weights,
    v [S, n] update weights.  ``forward(x[..., q, d]) -> [S, ..., q]``; with S=1
    the sample dim stays (the reference squeezes/unsqueezes around it,
    acqfs.py:362, 409-411, relying on broadcasting).
    """

    def __init__(self, W, w, v, train_X, *, kind, lengthscale, outputscale, mean_const):
        super().__init__()
        self.register_buffer("W", W)
        self.register_buffer("w", w)
        self.register_buffer("v", v)
        self.register_buffer("train_X", train_X)
        self.register_buffer("lengthscale", lengthscale)
        self.kind = kind
        self.outputscale = outputscale
        self.mean_const = float(mean_const)

    def features(self, x: Tensor) -> Tensor:
        M = self.w.shape[-1]
        z = (x / self.lengthscale).matmul(self.W.transpose(-1, -2))     # [..., M/2]
        os_ = 1.0 if self.outputscale is None else self.outputscale
        scale = math.sqrt(os_) * math.sqrt(2.0 / M)
        return scale * torch.cat([z.sin(), z.cos()], dim=-1)

    def forward(self, x: Tensor) -> Tensor:
        phi = self.features(x)                                            # [..., q, M]
        S = self.w.shape[0]
        w = self.w.reshape(S, *([1] * (phi.dim() - 1)), -1)
        prior = (phi.unsqueeze(0) * w).sum(-1)                            # [S, ..., q]
        kx = kernel_matrix(x, self.train_X, self.kind, self.lengthscale,
                           self.outputscale)                             # [..., q, n]
        v = self.v.reshape(S, *([1] * (kx.dim() - 1)), -1)
        update = (kx.unsqueeze(0) * v).sum(-1)
        out = self.mean_const + prior + update
        # BoTorch returns sample_shape + batch + [q]; S = 1 keeps a leading 1 that
        # broadcasts away when the caller's batch is [n_r, 1, d] (acqfs.py:362).
        if S == 1:
            return out.squeeze(0)
        return out


def draw_matheron_path(gp: DenseGP, train_Y: Tensor, n_paths: int = 1,
                       num_features: int = 1024, generator=None) -> MatheronPath:
    """``draw_matheron_paths(model, Size([S]))`` restated.

    RNG order (CPU Sobol scrambles for W and w, Gamma for the Matern mixing
    variable, randn for the noise sample) is recalled with low confidence; the
    hot path takes (W, w, v) as *inputs*, so its parity does not depend on it.
    """
    X = gp.train_X
    n, d = X.shape[-2:]
    dtype = X.dtype
    Mh = num_features // 2
    seed_w = int(torch.randint(0, 1_000_000, (1,), generator=generator).item())
    Wn = sobol_normal_base_samples(Mh, d, seed_w, dtype)                   # [M/2, d]
    if gp.kind != "rbf":
        nu = {"matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[gp.kind]
        g = torch.distributions.Gamma(nu, nu).sample((Mh, 1)).to(dtype)
        Wn = Wn / g.sqrt()
    seed_f = int(torch.randint(0, 1_000_000, (1,), generator=generator).item())
    w = sobol_normal_base_samples(n_paths, num_features, seed_f, dtype)   # [S, M]
    path = MatheronPath(Wn, w, torch.zeros(n_paths, n, dtype=dtype), X,
                        kind=gp.kind, lengthscale=gp.lengthscale,
                        outputscale=gp.outputscale, mean_const=gp.mean_const)
    phi = path.features(X)                                                 # [n, M]
    prior_at_train = gp.mean_const + phi.matmul(w.transpose(0, 1)).transpose(0, 1)
    eps = torch.randn(n_paths, n, dtype=dtype, generator=generator) * math.sqrt(gp.noise)
    resid = train_Y.reshape(1, -1) - (prior_at_train + eps)               # [S, n]
    v = torch.cholesky_solve(resid.transpose(0, 1), gp.L).transpose(0, 1)
    path.v.copy_(v)
    return path
