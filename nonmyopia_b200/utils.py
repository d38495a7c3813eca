"""Bayes-action evaluation - the reference's ``utils.eval_func``
(``synthetic_experiment/utils.py:143-299``, called by ``compute_metrics.py:26-87``)
on the B200 kernels.

Same call, same return value: ``eval_func(env, model, parms, buffer, iteration)`` ->
``((u_observed, u_posterior, regret), A_chosen.cpu())``.  The 1000-step AdamW loop over the
``n_restarts`` free actions runs on the device: the posterior mean / sigma and their gradients come
from the shared-root whitening kernels (``nmb_gp_whiten_fwd`` / ``_bwd`` through ``MCTree``), the
spotlight cost from ``nmb_tree_cost_*``, the update from ``nmb_adamw_flat``; after two eager steps one
step is captured as a CUDA graph and replayed, and nothing is read back before the last step.

Differences from the reference, on purpose:

* the non-HES branch of the reference raises ``NameError`` (``qBOAcqf`` / ``SobolQMCNormalSampler`` are
  never imported in utils.py); here it runs what the code plainly intends;
* ``n_actions > 1`` and discrete environments (``embedder``) raise ``NotImplementedError`` - outside
  the hot-path scope like everywhere else in this package (reference default ``n_actions = 1``).

Random draws: ``parms.rng_mode == "host"`` draws on the CPU generator in the reference's order
(bit-comparable with the oracle); the default draws on the device generator inside the graph.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch
from torch import Tensor

from . import _lib
# the helpers the reference's actor.py takes from utils (utils.py:313-365) live beside the Actor here
from .actor import generate_random_points_batch, generate_random_rotation_matrix, rotate_points  # noqa: F401
from .gp import SurrogateGP
from .mc_tree import MCTree

N_EVAL_STEPS = 1000                                                         # utils.py:171,250


def set_seed(seed: int) -> None:
    """utils.py:42-49: seed the host generators (python, numpy, torch) and every CUDA generator."""
    import random

    import numpy as np
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    torch.cuda.manual_seed_all(seed)


def _previous_cost(buffer, iteration: int, parms) -> float:
    """utils.py:146-150."""
    if iteration + 1 > parms.n_initial_points:
        return float(buffer["cost"][: iteration + 1].sum())
    return 0.0


class _Draws:
    """Standard-normal draws: the device generator (graph-safe), the host generator in the reference's
    order, or a replayed record ``dict(z0=..., steps=[[n_steps, ...], ...])`` (tests)."""

    def __init__(self, mode: str, dtype, device, replay: Optional[dict] = None):
        self.mode, self.dtype, self.device, self.replay = mode, dtype, device, replay
        self.step, self.j = -1, 0

    def begin_step(self, step: int) -> None:
        self.step, self.j = step, 0

    def __call__(self, *shape) -> Tensor:
        if self.replay is not None:
            src = self.replay["z0"] if self.step < 0 else self.replay["steps"][self.j][self.step]
            self.j += 1
            return src.reshape(shape).to(self.device, self.dtype)
        if self.mode == "host":
            return torch.randn(*shape, dtype=self.dtype).to(self.device)
        return torch.randn(*shape, dtype=self.dtype, device=self.device)

    @property
    def graph_safe(self) -> bool:
        return self.replay is None and self.mode != "host"


def _run_steps(step_fn: Callable[[], Tensor], draws: _Draws, n_steps: int, device) -> Tensor:
    """``n_steps`` calls of ``step_fn`` (one AdamW step; returns the per-restart ``losses``).  With a
    graph-safe draw source: two eager steps on a side stream, then one captured step replayed."""
    losses = None
    if not draws.graph_safe or n_steps < 3:
        for i in range(n_steps):
            draws.begin_step(i)
            losses = step_fn()
        return losses
    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(2):
            losses = step_fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        losses = step_fn()
    for _ in range(n_steps - 2):           # capture records the step without running it
        graph.replay()
    return losses


def eval_func(env, model, parms, buffer, iteration, embedder=None, *args, n_steps: int = N_EVAL_STEPS,
              _replay_draws: Optional[dict] = None, _z0: Optional[Tensor] = None, **kwargs):
    from .acqfs import SobolQMCNormalSampler, _TreeCostFn, qBOAcqf
    from .actor import _FlatAdamW
    if embedder is not None:
        raise NotImplementedError("discrete environments are outside the B200 hot-path scope")
    if parms.n_actions != 1:
        raise NotImplementedError("eval_func on the B200 path takes n_actions = 1 (reference main.py:57)")
    gp = SurrogateGP.from_model(model)
    dev, dt = gp.device, parms.torch_dtype
    if torch.device(dev).type != "cuda":
        raise _lib.NmbError("nonmyopia_b200 ops need CUDA tensors (no CPU fallback)")
    n_r, d = parms.n_restarts, parms.x_dim
    cost_fn = parms.cost_function_class(**parms.cost_func_hypers)
    previous_cost = _previous_cost(buffer, iteration, parms)
    u_observed = torch.max(buffer["y"][: iteration + 1]).item()
    x_it = buffer["x"][iteration].to(dev, dt)
    draws = _Draws(getattr(parms, "rng_mode", "device"), dt, dev, _replay_draws)

    # utils.py:158-159 / :223-224 - every restart starts at the last query plus 0.01 jitter
    z0 = draws(n_r, d) if _z0 is None else _z0.reshape(n_r, d).to(dev, dt)
    A = (x_it.clone().repeat(n_r, 1) + z0 * 0.01).contiguous()
    A.requires_grad_(True)
    opt = _FlatAdamW([A], lr=parms.acq_opt_lr)
    prev_X = x_it.reshape(1, d).expand(n_r, d).contiguous()

    if parms.algo.startswith("HES"):
        tree = MCTree(gp, [], n_r, [], None, dtype=dt)
        hyp = cost_fn.hypers()

        def step_fn():
            mean, sig = tree.leaf_posterior([A])                           # model(actions): noise-free posterior
            y = mean + sig * draws(n_r)                                    # ppd.rsample(), q = 1
            cost = _TreeCostFn.apply(prev_X, draws(n_r), hyp, previous_cost, A)
            # loss_fn(A, Y) with Y [n_r, 1] averages over dim 0 = the restarts (utils.py:179 passes no
            # sample dim), so the value term is one scalar shared by every restart
            losses = -y.mean() + cost
            opt.step(torch.autograd.grad(losses.mean(), [A]))
            return losses.detach()
    else:
        sampler = SobolQMCNormalSampler(sample_shape=1, seed=0, resample=False)
        acqf = qBOAcqf(name=parms.algo, model=gp, lookahead_steps=0 if parms.algo_lookahead_steps == 0 else 1,
                       n_actions=parms.n_actions, n_fantasy_at_design_pts=[1],
                       loss_function_class=parms.loss_function_class, loss_func_hypers=parms.loss_func_hypers,
                       cost_function_class=parms.cost_function_class, cost_func_hypers=parms.cost_func_hypers,
                       sampler=sampler, best_f=buffer["y"][: iteration + 1].max(),
                       rng_mode=getattr(parms, "rng_mode", "device"))
        lead = [prev_X.clone()] if parms.algo_lookahead_steps > 0 else []
        prev_y = buffer["y"][iteration: iteration + 1].expand(n_r, -1).to(dev, dt)
        n_noise = 2 if parms.algo == "qKG" else 1

        def step_fn():
            if not draws.graph_safe:
                nz = [draws(n_r, 1, 1) for _ in range(n_noise)]
                acqf._given_draws = nz if parms.algo == "qKG" else nz[0].reshape(-1)
            rd = acqf.forward(prev_X=prev_X, prev_y=prev_y, prev_hid_state=None, maps=lead + [A], embedder=None,
                              prev_cost=previous_cost)
            losses = rd["acqf_loss"] + rd["acqf_cost"]
            opt.step(torch.autograd.grad(losses.mean(), [A], allow_unused=True))
            return losses.detach()

    losses = _run_steps(step_fn, draws, n_steps, dev)
    aidx = losses.reshape(-1).argmin()
    A_chosen = A.detach().reshape(n_r, parms.n_actions, d)[aidx]
    if torch.isnan(A_chosen).any():                                        # utils.py:288-290
        print("A_chosen contains nan")
        A_chosen = x_it.clone().detach().reshape(1, d)
    u_posterior = env(A_chosen).item()
    regret = 3 - u_posterior                                               # utils.py:295 (sic)
    return (u_observed, u_posterior, regret), A_chosen.cpu()
