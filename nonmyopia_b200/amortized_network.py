"""Amortized policy network - host-side mirror of the reference module
(synthetic_experiment/amortized_network.py:15-139).

Same constructor, same ``forward`` signature and the same ``state_dict`` keys
(``embedding``, ``prepro.2``, ``rnn.*``, ``postpro_A.{0,3,6}``, ``postpro_X.{0,3,6}``), because
``state_dict()`` / ``load_state_dict()`` / ``.to(dtype, device)`` cross the boundary
(reference actor.py:48,525-535,635,710).

Inside the lookahead objective the network never runs as a PyTorch module: the whole
rollout (every policy step, rotation, Thompson path and cost) is one fused kernel that reads
the parameters as a flat buffer (``flat_parameters`` / ``nmb_ts_rollout_fwd``).  ``forward``
below serves the stand-alone calls outside the optimisation loop (the imitation
pre-training of ``Actor.reset_parameters``, reference actor.py:171-220) with device tensor
ops; it consumes the host generator exactly like the reference when ``enable_noise`` is set.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor, nn


def _mlp(n_in: int, hidden: int, n_out: int, p: float) -> nn.Sequential:
    # indices 0, 3, 6 carry the Linear layers (reference postpro_A / postpro_X layout)
    return nn.Sequential(nn.Linear(n_in, hidden), nn.ELU(), nn.Dropout(p=p),
                         nn.Linear(hidden, hidden), nn.ELU(), nn.Dropout(p=p),
                         nn.Linear(hidden, n_out))


def draw_rotation(n: int, dim: int, like: Tensor) -> Tuple[Tensor, Tensor]:
    """Random planar rotations in the reference's RNG order (amortized_network.py:136,142-156):
    ``rand(n)`` float32 angles on the host generator, float32 cos/sin cast to ``like``'s dtype
    afterwards, then ``n`` x ``randperm(dim)[:2]`` coordinate pairs.
    Returns ``Q [n, 2, 2]`` and ``dims [n, 2]`` (int32) on ``like``'s device."""
    if n >= 8 and dim >= 2:
        # the same generator stream walked in bulk by the C side (bit for bit; tests/test_host_rng.py)
        from .host_rng import CpuStream
        cs = CpuStream()
        theta = cs.rand_f32(n) * 2 * torch.pi - torch.pi
        dims = cs.randperm(n, dim)[:, :2]
        cs.commit()
    else:
        theta = torch.rand(n) * 2 * torch.pi - torch.pi
        dims = torch.stack([torch.randperm(dim)[:2] for _ in range(n)]) if n else torch.zeros(0, 2, dtype=torch.int64)
    c, s = torch.cos(theta), torch.sin(theta)
    Q = torch.stack([torch.stack([c, -s], -1), torch.stack([s, c], -1)], -2)
    return Q.to(like), dims.to(device=like.device, dtype=torch.int32)


def rotate_about(out: Tensor, x: Tensor, Q: Tensor, dims: Tensor) -> Tensor:
    """``x + M (out - x)`` with ``M`` the identity except for the 2x2 block ``Q`` on ``dims``."""
    v = out - x
    i, j = dims[:, 0:1].long(), dims[:, 1:2].long()
    vi, vj = v.gather(1, i), v.gather(1, j)
    ri = Q[:, 0, 0:1] * vi + Q[:, 0, 1:2] * vj
    rj = Q[:, 1, 0:1] * vi + Q[:, 1, 1:2] * vj
    return v.scatter(1, i, ri).scatter(1, j, rj) + x


class Project2Range(nn.Module):
    """Kept for attribute compatibility; like the reference it ignores its bounds and
    returns ``sigmoid`` (amortized_network.py:193-194)."""

    def __init__(self, _min, _max) -> None:
        super().__init__()
        self.min, self.max = _min, _max
        self.range = self.max - self.min

    def forward(self, x):
        return torch.sigmoid(x)


class AmortizedNetwork(nn.Module):
    def __init__(self, input_dim: int, output_dim: int, hidden_dim: int, n_actions: int, output_bounds,
                 discrete: bool = False, num_categories: int = 0) -> None:
        super().__init__()
        if discrete:
            raise NotImplementedError("the discrete branch is never enabled by the reference's sweeps "
                                      "(env_discretized: [False]) and is out of scope")
        self.input_dim, self.output_dim, self.hidden_dim = input_dim, output_dim, hidden_dim
        self.n_actions, self.output_bounds = n_actions, output_bounds
        self.discrete, self.num_categories = discrete, num_categories
        self.p = 0.0
        H = hidden_dim
        self.embedding = nn.Linear(input_dim, H)
        self.prepro = nn.Sequential(nn.ELU(), nn.Dropout(p=self.p), nn.Linear(H, H), nn.ELU(), nn.Dropout(p=self.p))
        self.rnn = nn.GRUCell(H, H)
        self.postpro_A = _mlp(2 * H, H, output_dim * n_actions, self.p)
        self.postpro_X = _mlp(2 * H, H, output_dim, self.p)
        self.project_output = Project2Range(output_bounds[..., 0], output_bounds[..., 1])

    # ---- flat views for the fused kernels (state_dict order == nmb layout)
    def flat_parameters(self) -> Tensor:
        """All parameters concatenated in ``state_dict`` order; differentiable w.r.t. each."""
        return torch.cat([p.reshape(-1) for p in self.parameters()])

    def forward(self, x: Tensor, y: Tensor, prev_hid_state: Tensor, return_actions: bool,
                enable_noise: bool = True):
        head = self.postpro_A if return_actions else self.postpro_X
        p = self.prepro(self.embedding(torch.cat([x, y], dim=-1)))
        hidden_state = self.rnn(p, prev_hid_state)
        output = self.project_output(head(torch.cat([p, hidden_state], dim=-1)))
        if enable_noise:
            Q, dims = draw_rotation(output.shape[0], output.shape[-1], output)
            output = rotate_about(output, x, Q, dims)
        return output, hidden_state
