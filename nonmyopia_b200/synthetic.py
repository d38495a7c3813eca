"""Synthetic benchmark problems for the lookahead-acquisition path (host side, data only).

Restates how the reference builds its inputs so that benchmarks and parity tests
run on the shapes BASELINE.json names:

* initial design: stratified cells + uniform remainder in the unit cube with the
  env-specific fixed last point (reference synthetic_experiment/main.py:246-307),
  drawn from ``np.random`` seeded like ``utils.set_seed`` (utils.py:42-49);
* objective values: closed-form Ackley / Hartmann-6 / Cosine8 / Alpine-N1 on the
  env bounds (main.py:94-188), negated as ``utils.make_env`` does (utils.py:56-93),
  standardised (the reference's ``EnvWrapper`` is an un-vendored dependency; the
  z-score used here is a documented stand-in) plus N(0, env_noise^2).

Nothing here touches the GPU; tensors are float64 CPU and the caller moves them.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

# name -> (dim, (lo, hi), n_initial_points, fixed last point)   [main.py:94-188, 276-307]
ENVS = {
    "Ackley": (2, (-2.0, 2.0), 50, [0.2, 0.8]),
    "Ackley4D": (4, (-2.0, 2.0), 100, [0.2, 0.8, 0.2, 0.8]),
    "Alpine": (2, (0.0, 10.0), 100, [0.5, 0.5]),
    "Cosine8": (8, (-1.0, 1.0), 200, [0.8625] * 8),
    "Hartmann": (6, (0.0, 1.0), 500, [0.5] * 6),
    "SynGP": (2, (-1.0, 1.0), 25, [0.725, 0.75]),
}
# spotlight radius by input dim [main.py:202-209]
RADIUS = {2: 0.075, 4: 0.1, 6: 0.125, 8: 0.15}


def initial_design(env_name: str, seed: int, n_points: Optional[int] = None) -> torch.Tensor:
    d, _, n_default, last = ENVS[env_name]
    n = n_default if n_points is None else n_points
    rs = np.random.RandomState(seed)
    n_part = int(n ** (1 / d))
    remaining = n - n_part**d
    edges = np.linspace(0.0, 1.0, n_part + 1)
    cells = np.array(np.meshgrid(*([list(range(n_part))] * d))).T.reshape(-1, d)
    lo, hi = edges[cells], edges[cells + 1]
    pts = np.concatenate(
        (rs.uniform(low=lo, high=hi, size=[n_part**d, d]),
         rs.uniform(low=0.0, high=1.0, size=[remaining, d])), axis=0)
    pts[-1] = last
    return torch.tensor(pts, dtype=torch.float64)


def _ackley(X):
    d = X.shape[-1]
    a, b, c = 20.0, 0.2, 2 * math.pi
    return (-a * torch.exp(-b / math.sqrt(d) * torch.linalg.norm(X, dim=-1))
            - torch.exp(torch.cos(c * X).mean(-1)) + a + math.e)


def _hartmann6(X):
    alpha = torch.tensor([1.0, 1.2, 3.0, 3.2], dtype=X.dtype)
    A = torch.tensor([[10, 3, 17, 3.5, 1.7, 8], [0.05, 10, 17, 0.1, 8, 14],
                      [3, 3.5, 1.7, 10, 17, 8], [17, 8, 0.05, 10, 0.1, 14]], dtype=X.dtype)
    P = 1e-4 * torch.tensor([[1312, 1696, 5569, 124, 8283, 5886], [2329, 4135, 8307, 3736, 1004, 9991],
                             [2348, 1451, 3522, 2883, 3047, 6650], [4047, 8828, 8732, 5743, 1091, 381]],
                            dtype=X.dtype)
    return -(alpha * torch.exp(-(A * (X.unsqueeze(-2) - P) ** 2).sum(-1))).sum(-1)


def _cosine8(X):
    return (0.1 * torch.cos(5 * math.pi * X) - X**2).sum(-1)


def _alpine(X):
    return (X * torch.sin(X) + 0.1 * X).abs().sum(-1)


def objective(env_name: str, X_unit: torch.Tensor) -> torch.Tensor:
    """Utility (to be maximised) at unit-cube points, before standardisation."""
    d, (lo, hi), _, _ = ENVS[env_name]
    X = lo + (hi - lo) * X_unit
    if env_name.startswith("Ackley"):
        return -_ackley(X)
    if env_name == "Hartmann":
        return -_hartmann6(X)
    if env_name == "Cosine8":
        return _cosine8(X)
    if env_name == "Alpine":
        return -_alpine(X)
    if env_name == "SynGP":
        # seeded smooth surrogate of the un-vendored SynGP env: a fixed random-feature draw
        g = torch.Generator().manual_seed(1234)
        W = torch.randn(64, d, generator=g, dtype=torch.float64) * 2.0
        b = torch.rand(64, generator=g, dtype=torch.float64) * 2 * math.pi
        a = torch.randn(64, generator=g, dtype=torch.float64)
        return math.sqrt(2 / 64) * (torch.cos(X @ W.T + b) * a).sum(-1)
    raise NotImplementedError(env_name)


@dataclass
class Problem:
    env_name: str
    d: int
    train_X: torch.Tensor        # [n, d] unit cube
    train_Y: torch.Tensor        # [n, 1] standardised utility
    radius: float
    kernel: str = "matern52"     # 0.10.0 default for --kernel RBF
    lengthscale: float = 0.25
    outputscale: Optional[float] = 1.0
    noise: float = 1e-2
    mean_const: float = 0.0


def make_problem(env_name: str, seed: int = 2, n_points: Optional[int] = None,
                 env_noise: float = 0.01, **hypers) -> Problem:
    d = ENVS[env_name][0]
    X = initial_design(env_name, seed, n_points)
    f = objective(env_name, X)
    y = (f - f.mean()) / f.std()
    g = torch.Generator().manual_seed(seed)
    y = y + env_noise * torch.randn(y.shape, generator=g, dtype=torch.float64)
    return Problem(env_name, d, X, y.reshape(-1, 1), RADIUS[d], **hypers)
