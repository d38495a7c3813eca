"""CPU oracle, part 2: a pure-PyTorch stand-in for the exact BoTorch 0.10.0 /
GPyTorch 1.11 / tensordict / plotting symbols that the reference's
``synthetic_experiment/{acqfs,actor,amortized_network,utils,main}.py`` import
(SURVEY.md Appendix B), so that those reference files can be imported VERBATIM
from ``/root/reference`` in the build container (they are never copied).

TEST INFRASTRUCTURE ONLY (see ``oracle/gp_math.py`` for the rules and for the
"parity unpinned for third-party arithmetic" note).  ``install()`` registers the
fake modules in ``sys.modules``; ``load_reference()`` additionally puts the
reference directory on ``sys.path`` and returns the reference modules - that part
only works where ``/root/reference`` exists (the build container, not the GPU box).
"""
from __future__ import annotations

import contextlib
import copy
import math
import os
import sys
import types
import warnings
from typing import Optional

import torch
from torch import Tensor, nn

from . import gp_math
from .gp_math import DenseGP

REFERENCE_DIR = "/root/reference/synthetic_experiment"


# --------------------------------------------------------------------------
# settings / flags
# --------------------------------------------------------------------------
class _Flag:
    _state = False

    def __init__(self, state: bool = True):
        self.state = state

    def __enter__(self):
        self.prev = type(self)._state
        type(self)._state = self.state

    def __exit__(self, *a):
        type(self)._state = self.prev
        return False

    @classmethod
    def on(cls):
        return cls._state

    @classmethod
    def off(cls):
        return not cls._state


class propagate_grads(_Flag):
    _state = False


class fantasize(_Flag):
    _state = False


# --------------------------------------------------------------------------
# gpytorch pieces
# --------------------------------------------------------------------------
class _Kernel(nn.Module):
    kind = "rbf"

    def __init__(self, ard_num_dims=None, **kw):
        super().__init__()
        nd = 1 if ard_num_dims is None else ard_num_dims
        self.register_buffer("_lengthscale", torch.full((1, nd), math.log(2.0), dtype=torch.float64))

    @property
    def lengthscale(self):
        return self._lengthscale

    @lengthscale.setter
    def lengthscale(self, v):
        v = torch.as_tensor(v, dtype=self._lengthscale.dtype)
        self._lengthscale = v.expand_as(self._lengthscale).clone()


class MaternKernel(_Kernel):
    def __init__(self, nu=2.5, ard_num_dims=None, **kw):
        super().__init__(ard_num_dims=ard_num_dims)
        if nu not in (0.5, 1.5, 2.5):
            raise RuntimeError("nu expected to be 0.5, 1.5, or 2.5")
        self.nu = nu
        self.kind = {0.5: "matern12", 1.5: "matern32", 2.5: "matern52"}[nu]


class RBFKernel(_Kernel):
    kind = "rbf"


class LinearKernel(nn.Module):
    kind = "linear"


class ScaleKernel(nn.Module):
    def __init__(self, base_kernel, **kw):
        super().__init__()
        self.base_kernel = base_kernel
        self.register_buffer("_outputscale", torch.tensor(math.log(2.0), dtype=torch.float64))

    @property
    def outputscale(self):
        return self._outputscale

    @outputscale.setter
    def outputscale(self, v):
        self._outputscale = torch.as_tensor(v, dtype=self._outputscale.dtype).clone()


class NormalPrior:
    def __init__(self, loc, scale):
        self.loc, self.scale = loc, scale


class GaussianLikelihood(nn.Module):
    def __init__(self, noise_prior=None, **kw):
        super().__init__()
        self.noise_prior = noise_prior
        self.register_buffer("_noise", torch.tensor([math.log(2.0) + 1e-4], dtype=torch.float64))

    @property
    def noise(self):
        return self._noise

    @noise.setter
    def noise(self, v):
        self._noise = torch.as_tensor(v, dtype=self._noise.dtype).reshape(1).clone()


class ConstantMean(nn.Module):
    def __init__(self):
        super().__init__()
        self.register_buffer("constant", torch.tensor(0.0, dtype=torch.float64))


class ExactMarginalLogLikelihood(nn.Module):
    def __init__(self, likelihood, model):
        super().__init__()
        self.likelihood, self.model = likelihood, model


class MultivariateNormal:
    """Just enough of gpytorch's MVN: mean/covariance, sample / rsample."""

    def __init__(self, mean: Tensor, covariance_matrix: Tensor):
        self.loc = self.mean = mean
        self.covariance_matrix = covariance_matrix

    @property
    def batch_shape(self):
        return self.mean.shape[:-1]

    @property
    def variance(self):
        return self.covariance_matrix.diagonal(dim1=-1, dim2=-2).clamp_min(1e-9)

    def rsample(self, sample_shape=torch.Size(), base_samples=None):
        root = gp_math.psd_safe_cholesky(self.covariance_matrix)
        if base_samples is None:
            z = torch.randn(*sample_shape, *self.mean.shape, dtype=self.mean.dtype)
        else:
            z = base_samples
        return self.mean + root.matmul(z.unsqueeze(-1)).squeeze(-1)

    def sample(self, sample_shape=torch.Size(), base_samples=None):
        with torch.no_grad():
            return self.rsample(sample_shape, base_samples)


# --------------------------------------------------------------------------
# botorch model + posterior
# --------------------------------------------------------------------------
class GPyTorchPosterior:
    def __init__(self, mvn: MultivariateNormal):
        self.distribution = self.mvn = mvn

    @property
    def mean(self):
        return self.mvn.mean.unsqueeze(-1)

    @property
    def variance(self):
        return self.mvn.variance.unsqueeze(-1)

    @property
    def device(self):
        return self.mvn.mean.device

    @property
    def dtype(self):
        return self.mvn.mean.dtype

    @property
    def base_sample_shape(self):
        return self.mvn.mean.shape          # batch_shape + [q]

    @property
    def batch_range(self):
        return (0, -1)

    def rsample_from_base_samples(self, sample_shape, base_samples):
        if base_samples.shape[: len(sample_shape)] != sample_shape:
            raise RuntimeError("sample_shape disagrees with shape of base_samples.")
        return self.mvn.rsample(sample_shape, base_samples).unsqueeze(-1)

    def rsample(self, sample_shape=torch.Size()):
        return self.mvn.rsample(sample_shape).unsqueeze(-1)


class SingleTaskGP(nn.Module):
    """BoTorch 0.10.0 ``SingleTaskGP`` facade over ``gp_math.DenseGP``.

    ``covar_module=None`` selects the 0.10.0 default
    ``ScaleKernel(MaternKernel(2.5, ard_num_dims=d))``.
    Hyper-parameters are plain buffers set by ``fit_gpytorch_mll`` below or by
    the caller (``model.covar_module.base_kernel.lengthscale = ...``).
    """

    def __init__(self, train_X, train_Y, likelihood=None, covar_module=None, **kw):
        super().__init__()
        self.train_inputs = (train_X,)
        self.train_targets = train_Y.squeeze(-1)
        self.likelihood = likelihood if likelihood is not None else GaussianLikelihood()
        self.mean_module = ConstantMean()
        if covar_module is None:
            covar_module = ScaleKernel(MaternKernel(2.5, ard_num_dims=train_X.shape[-1]))
        self.covar_module = covar_module
        self._dense: Optional[DenseGP] = None

    # hyper-parameter access ------------------------------------------------
    def hypers(self):
        cm = self.covar_module
        if isinstance(cm, ScaleKernel):
            base, os_ = cm.base_kernel, float(cm.outputscale)
        else:
            base, os_ = cm, None
        if base.kind == "linear":
            raise NotImplementedError("LinearKernel is outside the hot path (SURVEY section 0)")
        return dict(kind=base.kind, lengthscale=base.lengthscale.detach().to(torch.float64),
                    outputscale=os_, noise=float(self.likelihood.noise),
                    mean_const=float(self.mean_module.constant))

    def dense(self) -> DenseGP:
        if self._dense is None:
            X = self.train_inputs[0]
            h = self.hypers()
            h["lengthscale"] = h["lengthscale"].to(X.dtype)
            self._dense = DenseGP.fit_caches(X, self.train_targets, **h)
        return self._dense

    def eval(self):
        return super().eval()

    def transform_inputs(self, X):
        return X

    def posterior(self, X, observation_noise=False, **kw):
        mean, cov = self.dense().posterior(X, observation_noise=observation_noise)
        return GPyTorchPosterior(MultivariateNormal(mean, cov))

    def __call__(self, X):
        mean, cov = self.dense().posterior(X)
        return MultivariateNormal(mean, cov)

    def condition_on_observations(self, X, Y, **kw):
        new = copy.copy(self)
        new._modules = dict(self._modules)
        new._dense = self.dense().condition_on_observations(X, Y)
        return new

    def fantasize(self, X, sampler, observation_noise=True, **kw):
        with fantasize(True):
            post = self.posterior(X, observation_noise=observation_noise)
            Y = sampler(post)
            return self.condition_on_observations(X=X, Y=Y)

    def __deepcopy__(self, memo):
        new = copy.copy(self)
        new._modules = dict(self._modules)
        return new


def fit_gpytorch_mll(mll, **kw):
    """Exact-MLL hyper-parameter fit is OUT of the hot path.
    The oracle uses the fixed hyper-parameters SURVEY 8d prescribes
    (lengthscale 0.25, outputscale 1, noise 1e-2, constant mean 0) unless the
    caller has already set some."""
    model = mll.model
    cm = model.covar_module
    base = cm.base_kernel if isinstance(cm, ScaleKernel) else cm
    if not getattr(base, "_user_set", False):
        base.lengthscale = 0.25
    if isinstance(cm, ScaleKernel):
        cm.outputscale = 1.0
    model.likelihood.noise = 1e-2
    model._dense = None
    model.eval()
    return mll


# --------------------------------------------------------------------------
# samplers (BoTorch 0.10.0 sampling/base.py, sampling/normal.py)
# --------------------------------------------------------------------------
class MCSampler(nn.Module):
    def __init__(self, sample_shape, seed=None, **kwargs):
        super().__init__()
        if not isinstance(sample_shape, torch.Size):
            if isinstance(sample_shape, int):
                sample_shape = torch.Size([sample_shape])
            else:
                raise ValueError("sample_shape must be a torch.Size or int")
        for k, v in kwargs.items():
            if k == "resample" and v is False:
                warnings.warn("`resample` is deprecated", DeprecationWarning)
            elif k == "collapse_batch_dims" and v is True:
                warnings.warn("`collapse_batch_dims` is deprecated", DeprecationWarning)
            else:
                raise RuntimeError(f"Received an unknown argument {k}: {v}")
        self.sample_shape = sample_shape
        self.seed = seed if seed is not None else int(torch.randint(0, 1000000, (1,)).item())
        self.register_buffer("base_samples", None)

    def _get_collapsed_shape(self, posterior):
        base_sample_shape = posterior.base_sample_shape
        start, end = posterior.batch_range
        base_sample_shape = (
            base_sample_shape[:start]
            + torch.Size([1 for _ in base_sample_shape[start:end]])
            + base_sample_shape[end:]
        )
        return self.sample_shape + base_sample_shape

    def _get_extended_base_sample_shape(self, posterior):
        return self.sample_shape + posterior.base_sample_shape


class NormalMCSampler(MCSampler):
    def forward(self, posterior):
        self._construct_base_samples(posterior=posterior)
        return posterior.rsample_from_base_samples(
            sample_shape=self.sample_shape,
            base_samples=self.base_samples.expand(
                self._get_extended_base_sample_shape(posterior=posterior)
            ),
        )


class SobolQMCNormalSampler(NormalMCSampler):
    def _construct_base_samples(self, posterior):
        target_shape = self._get_collapsed_shape(posterior=posterior)
        if self.base_samples is None or self.base_samples.shape != target_shape:
            base_collapsed_shape = target_shape[len(self.sample_shape):]
            output_dim = base_collapsed_shape.numel()
            base = gp_math.sobol_normal_base_samples(
                self.sample_shape.numel(), output_dim, self.seed, posterior.dtype
            )
            self.register_buffer("base_samples", base.view(target_shape))
        self.to(dtype=posterior.dtype)


class IIDNormalSampler(NormalMCSampler):
    def _construct_base_samples(self, posterior):
        target_shape = self._get_collapsed_shape(posterior=posterior)
        if self.base_samples is None or self.base_samples.shape != target_shape:
            g = torch.Generator().manual_seed(self.seed)
            self.register_buffer(
                "base_samples", torch.randn(target_shape, generator=g, dtype=posterior.dtype)
            )


# --------------------------------------------------------------------------
# acquisition functions (BoTorch 0.10.0 acquisition/monte_carlo.py)
# --------------------------------------------------------------------------
class MCAcquisitionFunction(nn.Module):
    def __init__(self, model=None, sampler=None, **kw):
        super().__init__()
        self.model = model
        self.sampler = sampler

    def _samples(self, X):
        post = self.model.posterior(X)
        return self.sampler(post).squeeze(-1)           # [S, *batch, q]


class qExpectedImprovement(MCAcquisitionFunction):
    def __init__(self, model, best_f, sampler=None, **kw):
        super().__init__(model, sampler)
        self.best_f = torch.as_tensor(best_f)

    def forward(self, X):
        s = self._samples(X)
        return (s - self.best_f).clamp_min(0).max(dim=-1).values.mean(dim=0)


class qProbabilityOfImprovement(MCAcquisitionFunction):
    def __init__(self, model, best_f, sampler=None, tau=1e-3, **kw):
        super().__init__(model, sampler)
        self.best_f = torch.as_tensor(best_f)
        self.tau = tau

    def forward(self, X):
        s = self._samples(X)
        max_obj = s.max(dim=-1).values
        return torch.sigmoid((max_obj - self.best_f) / self.tau).mean(dim=0)


class qSimpleRegret(MCAcquisitionFunction):
    def forward(self, X):
        return self._samples(X).max(dim=-1).values.mean(dim=0)


class qUpperConfidenceBound(MCAcquisitionFunction):
    def __init__(self, model, beta, sampler=None, **kw):
        super().__init__(model, sampler)
        self.beta_prime = math.sqrt(beta * math.pi / 2)

    def forward(self, X):
        s = self._samples(X)
        mean = s.mean(dim=0)
        ucb = mean + self.beta_prime * (s - mean).abs()
        return ucb.max(dim=-1).values.mean(dim=0)


class qKnowledgeGradient(MCAcquisitionFunction):
    """One-shot KG: X [b, q + num_fantasies, d]; value = mean over fantasies of the
    posterior mean (under the fantasy model) at that fantasy's own inner point."""

    def __init__(self, model, num_fantasies=64, sampler=None, **kw):
        if sampler is None:
            sampler = SobolQMCNormalSampler(torch.Size([num_fantasies]))
        super().__init__(model, sampler)
        self.num_fantasies = num_fantasies

    def extract_candidates(self, X_full):
        return X_full[..., : -self.num_fantasies, :]

    def forward(self, X):
        X_actual = X[..., : -self.num_fantasies, :]                 # [b, q, d]
        X_fant = X[..., -self.num_fantasies:, :]                    # [b, nf, d]
        X_fant = X_fant.permute(-2, *range(X_fant.dim() - 2), -1).unsqueeze(-2)  # [nf,b,1,d]
        fmodel = self.model.fantasize(X=X_actual, sampler=self.sampler,
                                      observation_noise=True)
        with propagate_grads(True):
            mean, _ = fmodel.dense().posterior(X_fant)               # [nf, b, 1]
        values = mean.max(dim=-1).values                             # [nf, b]
        return values.mean(dim=0)


class qMultiStepLookahead(MCAcquisitionFunction):
    """BoTorch 0.10.0 ``qMultiStepLookahead`` as the reference uses it (acqfs.py:94-99): ``batch_sizes = [1] * k``,
    ``num_fantasies = [f_1..f_k]``, default samplers (one ``SobolQMCNormalSampler(Size([f_i]))`` per stage, seed from
    the global generator), default value functions ``[None, ..., None, PosteriorMean]`` (zero stage values, posterior
    mean of the k-times fantasised model at the terminal points), uniform sample weights.  One-shot input
    ``X [b, q', d]``, ``q' = 1 + f_1 + f_1 f_2 + ...``; stage i's points are ``X_i [f_i, .., f_1, b, 1, d]``.
    Fantasies use ``observation_noise=True`` and the fantasy caches keep their gradients (SURVEY A.4, A.8 [3P])."""

    def __init__(self, model, batch_sizes, num_fantasies=None, samplers=None, **kw):
        super().__init__(model, None)
        self.batch_sizes = list(batch_sizes)
        if samplers is None:
            samplers = [SobolQMCNormalSampler(torch.Size([nf])) for nf in num_fantasies]
        self.num_fantasies = [int(smp.sample_shape[0]) for smp in samplers]
        self.samplers = nn.ModuleList(samplers)

    def get_split_shapes(self, X):
        batch_shape, (q_aug, d) = X.shape[:-2], X.shape[-2:]
        rest = sum(math.prod(self.num_fantasies[:i + 1]) * q for i, q in enumerate(self.batch_sizes))
        batch_sizes = [q_aug - rest] + self.batch_sizes
        shapes = [torch.Size(self.num_fantasies[:i][::-1] + [*batch_shape, q_i, d])
                  for i, q_i in enumerate(batch_sizes)]
        sizes = [s[: (-2 - len(batch_shape))].numel() * s[-2] for s in shapes]
        return batch_shape, shapes, sizes

    def get_multi_step_tree_input_representation(self, X):
        batch_shape, shapes, sizes = self.get_split_shapes(X)
        Xsplit = torch.split(X, sizes, dim=-2)
        perm = [-2] + list(range(len(batch_shape))) + [-1]
        X0 = Xsplit[0].reshape(shapes[0])
        return [X0] + [Xi.permute(*perm).reshape(shape) for Xi, shape in zip(Xsplit[1:], shapes[1:])]

    def forward(self, X):
        Xs = self.get_multi_step_tree_input_representation(X)
        model = self.model
        for i, Xi in enumerate(Xs[:-1]):
            model = model.fantasize(X=Xi, sampler=self.samplers[i], observation_noise=True)
        with propagate_grads(True):
            mean, _ = model.dense().posterior(Xs[-1])                # [f_k, .., f_1, b, 1]
        val = mean.squeeze(-1)
        b = X.shape[:-2]
        return val.reshape(-1, *b).mean(dim=0)


class qNegIntegratedPosteriorVariance(MCAcquisitionFunction):
    """The reference constructs it with ``mc_points=0`` (acqfs.py:101-104); BoTorch's constructor hands that to
    ``register_buffer``, which accepts only tensors: the name cannot be instantiated upstream."""

    def __init__(self, model, mc_points, sampler=None, **kw):
        super().__init__(model, sampler)
        self.register_buffer("mc_points", mc_points)


# --------------------------------------------------------------------------
# pathwise sampling
# --------------------------------------------------------------------------
def draw_matheron_paths(model, sample_shape, **kw):
    S = int(torch.Size(sample_shape).numel())
    return gp_math.draw_matheron_path(model.dense(), model.train_targets, n_paths=S)


# --------------------------------------------------------------------------
# tensordict / env / plotting stubs
# --------------------------------------------------------------------------
class TensorDict:
    def __init__(self, source, batch_size=None, device=None):
        self._d = dict(source)
        self.batch_size = batch_size

    def __getitem__(self, key):
        if isinstance(key, str):
            return self._d[key]
        return TensorDict({k: v[key] for k, v in self._d.items()})

    def __setitem__(self, key, value):
        self._d[key] = value

    def keys(self):
        return self._d.keys()


class _TestFunction(nn.Module):
    dim = 2
    _default_bounds = (0.0, 1.0)

    def __init__(self, dim=None, negate=False, noise_std=None, **kw):
        super().__init__()
        if dim is not None:
            self.dim = dim
        self.negate = negate
        self.noise_std = noise_std
        lo, hi = self._default_bounds
        self.register_buffer(
            "bounds", torch.tensor([[lo] * self.dim, [hi] * self.dim], dtype=torch.float64)
        )

    def evaluate_true(self, X):
        raise NotImplementedError

    def forward(self, X, noise=True):
        f = self.evaluate_true(X)
        if noise and self.noise_std:
            f = f + self.noise_std * torch.randn_like(f)
        return -f if self.negate else f


class Ackley(_TestFunction):
    def evaluate_true(self, X):
        a, b, c = 20.0, 0.2, 2 * math.pi
        part1 = -a * torch.exp(-b / math.sqrt(self.dim) * torch.linalg.norm(X, dim=-1))
        part2 = -torch.exp(torch.mean(torch.cos(c * X), dim=-1))
        return part1 + part2 + a + math.e


class Hartmann(_TestFunction):
    dim = 6

    def evaluate_true(self, X):
        ALPHA = torch.tensor([1.0, 1.2, 3.0, 3.2], dtype=X.dtype)
        A = torch.tensor([[10, 3, 17, 3.5, 1.7, 8], [0.05, 10, 17, 0.1, 8, 14],
                          [3, 3.5, 1.7, 10, 17, 8], [17, 8, 0.05, 10, 0.1, 14]], dtype=X.dtype)
        P = 1e-4 * torch.tensor([[1312, 1696, 5569, 124, 8283, 5886],
                                 [2329, 4135, 8307, 3736, 1004, 9991],
                                 [2348, 1451, 3522, 2883, 3047, 6650],
                                 [4047, 8828, 8732, 5743, 1091, 381]], dtype=X.dtype)
        inner = torch.sum(A * (X.unsqueeze(-2) - P) ** 2, dim=-1)
        return -(ALPHA * torch.exp(-inner)).sum(dim=-1)


class Cosine8(_TestFunction):
    dim = 8

    def evaluate_true(self, X):
        return torch.sum(0.1 * torch.cos(5 * math.pi * X) - X**2, dim=-1)


class AlpineN1(_TestFunction):
    def evaluate_true(self, X):
        return -torch.sum(torch.abs(X * torch.sin(X) + 0.1 * X), dim=-1)


class _GenericTestFunction(_TestFunction):
    def evaluate_true(self, X):
        return -(X**2).sum(-1)


class EnvWrapper(nn.Module):
    """Un-vendored ``synthetic_functions.env_wrapper.EnvWrapper`` stand-in: maps the
    unit cube to the function's bounds and standardises with fixed statistics
    estimated on a seeded Sobol design (documented deviation, SURVEY 8d cfg 1)."""

    def __init__(self, name, function):
        super().__init__()
        self.name, self.function = name, function
        g = torch.quasirandom.SobolEngine(function.dim, scramble=True, seed=0).draw(4096).double()
        vals = function.evaluate_true(self._to_bounds(g))
        vals = -vals if function.negate else vals
        self.mu, self.sd = float(vals.mean()), float(vals.std())

    def _to_bounds(self, X):
        lo, hi = self.function.bounds[0].to(X), self.function.bounds[1].to(X)
        return lo + (hi - lo) * X

    def forward(self, X):
        return (self.function(self._to_bounds(X)) - self.mu) / self.sd


# --------------------------------------------------------------------------
# installation
# --------------------------------------------------------------------------
def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []          # behave as a package so sub-imports resolve
    sys.modules[name] = m
    return m


def install() -> None:
    """Register the fake third-party modules (idempotent)."""
    if getattr(sys.modules.get("botorch"), "_is_oracle_shim", False):
        return
    acq = dict(qExpectedImprovement=qExpectedImprovement, qKnowledgeGradient=qKnowledgeGradient,
               qMultiStepLookahead=qMultiStepLookahead, qNegIntegratedPosteriorVariance=qNegIntegratedPosteriorVariance,
               qProbabilityOfImprovement=qProbabilityOfImprovement, qSimpleRegret=qSimpleRegret,
               qUpperConfidenceBound=qUpperConfidenceBound)
    settings = _mod("botorch.settings", propagate_grads=propagate_grads)
    bt = _mod("botorch", settings=settings, fit_gpytorch_model=fit_gpytorch_mll,
              _is_oracle_shim=True)
    bt.acquisition = _mod("botorch.acquisition", **acq)
    _mod("botorch.acquisition.monte_carlo", MCAcquisitionFunction=MCAcquisitionFunction)
    _mod("botorch.models", SingleTaskGP=SingleTaskGP)
    _mod("botorch.models.utils")
    _mod("botorch.models.utils.assorted", fantasize=fantasize)
    _mod("botorch.posteriors", Posterior=GPyTorchPosterior)
    _mod("botorch.sampling")
    _mod("botorch.sampling.base", MCSampler=MCSampler)
    _mod("botorch.sampling.normal", NormalMCSampler=NormalMCSampler,
         SobolQMCNormalSampler=SobolQMCNormalSampler, IIDNormalSampler=IIDNormalSampler)
    _mod("botorch.sampling.pathwise")
    _mod("botorch.sampling.pathwise.posterior_samplers", draw_matheron_paths=draw_matheron_paths)
    _mod("botorch.fit", fit_gpytorch_mll=fit_gpytorch_mll)
    _mod("botorch.test_functions")
    names = ["Beale", "Branin", "EggHolder", "Griewank", "HolderTable", "Levy", "Powell",
             "Rosenbrock", "SixHumpCamel", "StyblinskiTang"]
    _mod("botorch.test_functions.synthetic", Ackley=Ackley, Hartmann=Hartmann, Cosine8=Cosine8,
         **{n: type(n, (_GenericTestFunction,), {}) for n in names})
    kernels = _mod("gpytorch.kernels", MaternKernel=MaternKernel, LinearKernel=LinearKernel,
                   RBFKernel=RBFKernel, ScaleKernel=ScaleKernel)
    lik = _mod("gpytorch.likelihoods", GaussianLikelihood=GaussianLikelihood)
    pri = _mod("gpytorch.priors", NormalPrior=NormalPrior)
    mlls = _mod("gpytorch.mlls", ExactMarginalLogLikelihood=ExactMarginalLogLikelihood)
    _mod("gpytorch", kernels=kernels, likelihoods=lik, priors=pri, mlls=mlls)
    _mod("tensordict", TensorDict=TensorDict)
    _mod("synthetic_functions")
    _mod("synthetic_functions.alpine", AlpineN1=AlpineN1)
    _mod("synthetic_functions.env_wrapper", EnvWrapper=EnvWrapper)
    _mod("synthetic_functions.nightlight", NightLight=_GenericTestFunction)
    _mod("synthetic_functions.syngp", SynGP=_GenericTestFunction)

    class _Anything:
        def __getattr__(self, k):
            return _Anything()

        def __call__(self, *a, **kw):
            return _Anything()

        def update(self, *a, **kw):
            return None

    def _lazy(k):
        if k.startswith("__"):
            raise AttributeError(k)
        return _Anything()

    plt = _mod("matplotlib.pyplot", rcParams=_Anything(), __getattr__=_lazy)
    _mod("matplotlib", pyplot=plt)
    _mod("imageio", __getattr__=_lazy)
    bundles = _mod("tueplots.bundles", iclr2023=lambda *a, **k: {}, iclr2024=lambda *a, **k: {})
    _mod("tueplots", bundles=bundles, figsizes=_Anything())


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "acqfs.py"))


@contextlib.contextmanager
def _ref_path():
    sys.path.insert(0, REFERENCE_DIR)
    try:
        yield
    finally:
        sys.path.remove(REFERENCE_DIR)


def load_reference():
    """Import the reference's acqfs / amortized_network / actor VERBATIM over the
    shim (build container only).  Returns a namespace of the three modules."""
    if not reference_available():
        raise FileNotFoundError(f"{REFERENCE_DIR} not present (GPU box?)")
    install()
    import importlib

    with _ref_path():
        mods = {}
        for name in ("amortized_network", "acqfs", "utils", "actor"):
            # the product package has same-named modules under nonmyopia_b200/,
            # never top-level, so these names are unambiguous here
            mods[name] = importlib.import_module(name)
    return types.SimpleNamespace(**mods)
