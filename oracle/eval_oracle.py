"""CPU restatement of the reference's Bayes-action evaluation ``utils.eval_func``
(``/root/reference/synthetic_experiment/utils.py:143-299``).

TEST INFRASTRUCTURE ONLY - imported by ``tests/`` and ``oracle/make_golden.py``; the
product (``nonmyopia_b200/utils.py``) never touches it.

Pinned by ``oracle/make_golden.py`` (``case_eval``) against the reference's own
``utils.eval_func`` imported verbatim over ``oracle/botorch_shim.py``:

* HES algorithms (``utils.py:156-194``): 1000 AdamW steps on ``n_restarts x n_actions``
  free actions of ``-mean(rsample of the posterior at A) + cost``; the restart with
  the smallest ``loss + cost`` row is evaluated on the environment.
* the other algorithms (``utils.py:196-276``) call ``qBOAcqf`` / ``SobolQMCNormalSampler``
  which ``utils.py:1-30`` never imports (latent ``NameError``); the pin
  supplies those two names to the reference module and is otherwise verbatim.

The random stream is consumed in the reference's order: the 0.01 initial jitter, then
per step the posterior draw and the cost noise (HES) or the cost noise(s) (others).
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch
from torch import Tensor

from . import gp_math
from . import hes_oracle as ho


def _previous_cost(buffer: Dict[str, Tensor], iteration: int, n_initial_points: int):
    """utils.py:146-150."""
    if iteration + 1 > n_initial_points:
        return buffer["cost"][: iteration + 1].sum()
    return 0.0


def eval_func(env: Callable[[Tensor], Tensor], gp: gp_math.DenseGP, buffer: Dict[str, Tensor], iteration: int, *,
              algo: str, n_restarts: int, n_actions: int, n_initial_points: int, lr: float,
              cost_spec: ho.CostSpec, lookahead_steps: int = 0, dist_weight: float = 1.0,
              dist_threshold: float = 0.5, n_steps: int = 1000, record: Optional[list] = None,
              draws: Optional[dict] = None):
    """Returns ``((u_observed, u_posterior, regret), A_chosen)`` like utils.py:299.

    ``record`` (a list) receives every random draw in consumption order so that a
    device implementation can be replayed on the same stream; ``draws`` (the fixture form
    ``dict(z0=..., steps=[stacked j-th draw of every step])``) replays such a record
    instead of drawing."""
    prev_cost = _previous_cost(buffer, iteration, n_initial_points)
    u_observed = torch.max(buffer["y"][: iteration + 1]).item()
    x_it = buffer["x"][iteration]
    d = x_it.shape[-1]
    rec = (lambda t: record.append(t.clone())) if record is not None else (lambda t: None)
    cursor = {"step": -1, "j": 0}

    def randn(shape):
        if draws is None:
            return torch.randn(*shape, dtype=x_it.dtype)
        if cursor["step"] < 0:
            return draws["z0"].reshape(shape)
        t = draws["steps"][cursor["j"]][cursor["step"]].reshape(shape)
        cursor["j"] += 1
        return t

    if algo.startswith("HES"):
        A = x_it.clone().repeat(n_restarts, n_actions, 1)                      # utils.py:158
        z0 = randn(A.shape)
        rec(z0)
        A = (A + z0 * 0.01).requires_grad_(True)
        opt = torch.optim.AdamW([A], lr=lr)
        for step in range(n_steps):
            cursor.update(step=step, j=0)
            mean, cov = gp.posterior(A)                                        # model(actions), noise-free
            root = gp_math.psd_safe_cholesky(cov)
            z = randn(mean.shape)                    # ppd.rsample()
            rec(z)
            y = mean + root.matmul(z.unsqueeze(-1)).squeeze(-1)
            dist_shape = (n_restarts, n_actions, n_actions)
            nz = randn(dist_shape)                     # cost_fn's randn_like(dist)
            rec(nz)
            losses = ho.topk_loss(A, y, dist_weight, dist_threshold) + ho.spotlight_cost(
                x_it.expand_as(A), A, cost_spec, nz, prev_cost)                # utils.py:179-183
            loss = losses.mean()
            opt.zero_grad()
            loss.backward()
            opt.step()
        aidx = losses.squeeze(-1).argmin()                                     # utils.py:194
    else:
        nfant = 1                                                              # nf_design_pts = [1], utils.py:199
        base = gp_math.sobol_normal_base_samples(1, 1, 0, dtype=x_it.dtype).reshape(-1)   # sample_shape=1, seed=0
        best_f = float(buffer["y"][: iteration + 1].max())
        maps = []
        if lookahead_steps > 0:
            maps.append(x_it.clone().repeat(n_restarts, 1))                    # utils.py:216
        A = x_it.clone().repeat(n_restarts * n_actions, 1)
        z0 = randn(A.shape)
        rec(z0)
        A = (A + z0 * 0.01).requires_grad_(True)
        maps.append(A)
        opt = torch.optim.AdamW([A], lr=lr)
        prev_X = buffer["x"][iteration: iteration + 1].expand(n_restarts, -1)
        for step in range(n_steps):
            cursor.update(step=step, j=0)
            q = (len(maps) - 1) + n_actions                                    # rows of the [n_r, q, d] reshape
            if algo == "qKG":                                                  # acqfs.py:152-164 (two cdist blocks)
                nzs = [randn((n_restarts, q - nfant, q - nfant)), randn((n_restarts, nfant, nfant))]
            else:
                nzs = [randn((n_restarts, q, q))]
            for t in nzs:
                rec(t)
            rd = ho.baseline_forward(algo, gp, prev_X, maps, base=base, cost_spec=cost_spec, best_f=best_f,
                                     prev_cost=prev_cost, cost_noise=nzs, num_fantasies=nfant)
            losses = rd["acqf_loss"] + rd["acqf_cost"]
            loss = losses.mean()
            (g,) = torch.autograd.grad(loss, [A], allow_unused=True)
            A.grad = g
            opt.step()
            opt.zero_grad()
        aidx = losses.squeeze(-1).argmin()
        A = A.reshape(n_restarts, n_actions, d)

    A_chosen = A[aidx].detach()
    if torch.isnan(A_chosen).any():                                            # utils.py:288-290
        A_chosen = x_it.clone().detach()
    u_posterior = env(A_chosen).item()
    regret = 3 - u_posterior                                                   # utils.py:295 (sic)
    return (u_observed, u_posterior, regret), A_chosen.cpu()
