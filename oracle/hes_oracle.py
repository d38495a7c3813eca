"""CPU oracle, part 3: a plain-PyTorch (CPU, autograd) restatement of the
reference's lookahead-acquisition hot path.

TEST INFRASTRUCTURE ONLY - never imported by ``nonmyopia_b200`` (see
``oracle/gp_math.py``).  It restates, function by function, what the reference
computes in

* ``synthetic_experiment/acqfs.py:277-456``   multi-step H-entropy search forward
* ``synthetic_experiment/acqfs.py:109-179``   myopic / KG baselines forward
* ``synthetic_experiment/acqfs.py:548-580``   top-k loss
* ``synthetic_experiment/acqfs.py:604-633``   spotlight cost
* ``synthetic_experiment/amortized_network.py:97-166``  policy network + rotation
* ``synthetic_experiment/actor.py:439-734``   the AdamW optimisation loop

on top of ``oracle/gp_math.py`` (the third-party arithmetic).  The restatement is
pinned in the build container against the reference files themselves, imported
verbatim over ``oracle/botorch_shim.py`` (``oracle/make_golden.py`` asserts exact
agreement on seeded inputs and writes ``tests/golden/*.pt``); third-party
arithmetic stays "parity unpinned" (BoTorch/GPyTorch wheels are unavailable).

Random numbers: the reference draws (i) a rotation angle + coordinate pair per
sample inside every noisy policy call (CPU generator; float32 angles,
amortized_network.py:136,142-156) and (ii) ``randn_like`` noise inside every cost
call (acqfs.py:624).  Every function here takes those draws as explicit tensors;
``draw_*`` helpers consume the global generator in exactly the reference's order.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F
from torch import Tensor

from . import gp_math
from .gp_math import DenseGP, MatheronPath


# --------------------------------------------------------------------------
# cost / loss
# --------------------------------------------------------------------------
@dataclass
class CostSpec:
    """Hyper-parameters of the spotlight cost (acqfs.py:586-602; presets main.py:219-237)."""
    radius: float
    k: float = 1.0
    max_noise: float = 1e-5
    p_norm: float = 2.0
    discount: float = 0.0
    discount_threshold: float = -1.0


def spotlight_cost(prev: Tensor, cur: Tensor, spec: CostSpec, noise: Optional[Tensor],
                   previous_cost=None) -> Tensor:
    """acqfs.py:604-633.  ``prev`` [..., P, d], ``cur`` [..., C, d] -> [..., C, P].
    ``noise`` is the standard-normal draw of that call (None: draw it here)."""
    dist = torch.cdist(cur, prev, p=spec.p_norm)
    if noise is None:
        noise = torch.randn_like(dist)
    # the reference keeps the radius in a default-dtype (float32) buffer
    # (acqfs.py:597), so the FP64 path subtracts the float32-rounded value
    radius = float(torch.tensor(spec.radius, dtype=torch.float32))
    hinge = torch.clamp_min(spec.k * (dist - radius), 0.0) + noise * spec.max_noise
    if spec.discount > 0.0:
        over = (previous_cost + hinge > spec.discount_threshold).to(torch.float32)
        hinge = hinge * (1 - spec.discount * over)
    return hinge


def topk_loss(A: Tensor, Y: Tensor, dist_weight: float = 1.0, dist_threshold: float = 0.5) -> Tensor:
    """acqfs.py:548-580.  A [*t, n_r, n_a, d]; Y [S, *t, n_r, n_a] -> [*t, n_r]."""
    value = Y.sum(dim=-1).mean(dim=0)
    n_a = A.shape[-2]
    diversity = 0.0
    if n_a >= 2:
        pair = torch.triu(torch.cdist(A.contiguous(), A.contiguous(), p=1.0))
        pair = torch.where(pair > dist_threshold, torch.full_like(pair, dist_threshold), pair)
        diversity = pair.sum((-1, -2)) / (n_a * (n_a - 1) / 2.0)
    return -value - dist_weight * diversity


# --------------------------------------------------------------------------
# policy network (amortized_network.py) - functional, parameters by state-dict name
# --------------------------------------------------------------------------
def init_policy_params(d: int, H: int, n_actions: int = 1, dtype=torch.float64,
                       generator=None) -> Dict[str, Tensor]:
    """Random parameters with the reference module's names and shapes
    (amortized_network.py:51-91; torch default init is irrelevant to parity)."""
    def lin(o, i):
        b = 1.0 / math.sqrt(i)
        return ((torch.rand(o, i, dtype=dtype, generator=generator) * 2 - 1) * b,
                (torch.rand(o, dtype=dtype, generator=generator) * 2 - 1) * b)
    P = {}
    P["embedding.weight"], P["embedding.bias"] = lin(H, d + 1)
    P["prepro.2.weight"], P["prepro.2.bias"] = lin(H, H)
    b = 1.0 / math.sqrt(H)
    for nm, shp in (("weight_ih", (3 * H, H)), ("weight_hh", (3 * H, H)),
                    ("bias_ih", (3 * H,)), ("bias_hh", (3 * H,))):
        P["rnn." + nm] = (torch.rand(*shp, dtype=dtype, generator=generator) * 2 - 1) * b
    for head, o in (("postpro_A", d * n_actions), ("postpro_X", d)):
        P[head + ".0.weight"], P[head + ".0.bias"] = lin(H, 2 * H)
        P[head + ".3.weight"], P[head + ".3.bias"] = lin(H, H)
        P[head + ".6.weight"], P[head + ".6.bias"] = lin(o, H)
    return P


def policy_forward(P: Dict[str, Tensor], x: Tensor, y: Tensor, h: Tensor, head: str):
    """amortized_network.py:97-133 (continuous branch, dropout p=0):
    Linear(d+1,H) . ELU . Linear(H,H) . ELU -> GRUCell -> cat -> 3-layer MLP -> sigmoid.
    ``head`` is 'X' (design step) or 'A' (action step).  Returns (output, h')."""
    e = F.linear(torch.cat([x, y], dim=-1), P["embedding.weight"], P["embedding.bias"])
    p = F.elu(F.linear(F.elu(e), P["prepro.2.weight"], P["prepro.2.bias"]))
    H = h.shape[-1]
    gi = F.linear(p, P["rnn.weight_ih"], P["rnn.bias_ih"])
    gh = F.linear(h, P["rnn.weight_hh"], P["rnn.bias_hh"])
    r = torch.sigmoid(gi[..., :H] + gh[..., :H])
    z = torch.sigmoid(gi[..., H:2 * H] + gh[..., H:2 * H])
    n = torch.tanh(gi[..., 2 * H:] + r * gh[..., 2 * H:])
    h_new = (1 - z) * n + z * h
    pre = "postpro_A" if head == "A" else "postpro_X"
    u = F.elu(F.linear(torch.cat([p, h_new], dim=-1), P[pre + ".0.weight"], P[pre + ".0.bias"]))
    u = F.elu(F.linear(u, P[pre + ".3.weight"], P[pre + ".3.bias"]))
    out = torch.sigmoid(F.linear(u, P[pre + ".6.weight"], P[pre + ".6.bias"]))
    return out, h_new


def draw_rotation(N: int, D: int, dtype=torch.float64):
    """The random draws of one noisy policy call, in the reference's order
    (amortized_network.py:136 then :144-156): ``rand(N)`` float32 angles, the 2x2
    blocks built in float32 and only then cast, and N
    ``randperm(D)[:2]`` coordinate pairs.  Returns (Q [N,2,2] dtype, dims [N,2] int64)."""
    theta = torch.rand(N) * 2 * torch.pi - torch.pi
    c, s = torch.cos(theta), torch.sin(theta)
    Q = torch.stack([torch.stack([c, -s], -1), torch.stack([s, c], -1)], -2).to(dtype)
    dims = torch.stack([torch.randperm(D)[:2] for _ in range(N)])
    return Q, dims


def apply_rotation(out: Tensor, x: Tensor, Q: Tensor, dims: Tensor) -> Tensor:
    """amortized_network.py:137,157-166: rotate (out - x) in the plane of the two
    chosen coordinates, leave the others, add x back."""
    v = out - x
    i, j = dims[:, 0:1], dims[:, 1:2]
    vi, vj = v.gather(1, i), v.gather(1, j)
    ri = Q[:, 0, 0:1] * vi + Q[:, 0, 1:2] * vj
    rj = Q[:, 1, 0:1] * vi + Q[:, 1, 1:2] * vj
    # the reference multiplies by a dense D x D matrix (identity off the block);
    # summation order inside that einsum differs from this 2-term form only by
    # exact zeros and ones, so results are bit-identical.
    v = v.scatter(1, i, ri).scatter(1, j, rj)
    return v + x


# --------------------------------------------------------------------------
# the H-entropy-search objective  (acqfs.py:277-456)
# --------------------------------------------------------------------------
@dataclass
class HESSpec:
    lookahead: int
    nf: List[int]                     # fantasy widths per design level (TS: all 1)
    n_actions: int = 1
    cost: CostSpec = None
    dist_weight: float = 1.0
    dist_threshold: float = 0.5
    enable_ts: bool = False


@dataclass
class HESDraws:
    """Every random tensor one objective evaluation consumes, in call order."""
    rot: List[Optional[tuple]] = field(default_factory=list)   # per policy call: (Q, dims) or None
    cost_noise: List[Tensor] = field(default_factory=list)     # per cost call


def hes_forward(spec: HESSpec, prev_X: Tensor, prev_y: Tensor, prev_h: Optional[Tensor],
                maps, *, path: Optional[Callable] = None, gp: Optional[DenseGP] = None,
                design_base: Optional[Sequence[Tensor]] = None, action_base: Optional[Tensor] = None,
                prev_cost=0.0, enable_noise: bool = True,
                draws: Optional[HESDraws] = None) -> Dict[str, object]:
    """One evaluation of the multi-step objective for every restart.

    ``maps`` is either a parameter dict of the policy network (amortised) or a list
    of L+1 free tensors already in the unit cube.  TS: ``path`` is the Thompson
    sample; MC: ``gp`` + Sobol base samples ``design_base[s]`` [nf_s] and
    ``action_base`` [S_a] (zeros reproduce ``PosteriorMeanSampler``).
    ``draws=None`` consumes the global RNG like the reference does; the draws
    actually used are returned under key 'draws'.
    """
    amortised = isinstance(maps, dict)
    L, nf, n_a = spec.lookahead, list(spec.nf), spec.n_actions
    n_r, d = prev_X.shape
    used = HESDraws()
    given = draws
    call_no = 0

    def policy(x, y, h, head, noisy):
        nonlocal call_no
        out, h2 = policy_forward(maps, x, y, h, head)
        rot = None
        if noisy:
            rot = given.rot[call_no] if given is not None else draw_rotation(out.shape[0], out.shape[1], out.dtype)
            out = apply_rotation(out, x, rot[0], rot[1])
        used.rot.append(rot)
        call_no += 1
        return out, h2

    cur_x, cur_y, cur_h = prev_X, prev_y, prev_h
    model = gp
    X_levels, H_levels = [], []
    for s in range(L):
        if amortised:
            X, h = policy(cur_x, cur_y, cur_h, "X", enable_noise)
        else:
            X, h = maps[s], cur_h
        X = X.reshape(*nf[:s][::-1], n_r, 1, d)
        X_levels.append(X)
        H_levels.append(h)
        if spec.enable_ts:
            ys = path(X.reshape(n_r, 1, d)).unsqueeze(0)             # all nf are 1 in TS
        else:
            mean, cov = model.posterior(X)
            base = design_base[s].reshape(nf[s], *([1] * (mean.dim() - 1)), 1)
            ys = gp_math.sample_from_posterior(mean, cov, base)       # [nf_s, *b, 1, 1]
            model = model.condition_on_observations(X, ys)
        cur_x = X.unsqueeze(0).expand(nf[s], *X.shape).reshape(-1, d)
        cur_y = ys.reshape(-1, 1)
        if h is not None:
            cur_h = h.unsqueeze(0).expand(nf[s], *h.shape).reshape(-1, h.shape[-1])

    if amortised:
        # the action call never receives enable_noise => always rotated (acqfs.py:390-395)
        A, h = policy(cur_x, cur_y, cur_h, "A", True)
    else:
        A = maps[L]
    A = A.reshape(*nf[::-1], n_r, n_a, d)
    if spec.enable_ts:
        A_y = path(A.reshape(n_r, n_a, d)).unsqueeze(0)
        A_y = A_y.reshape(1, *A.shape[:-1])
    else:
        mean, cov = model.posterior(A)
        base = action_base.reshape(-1, *([1] * (mean.dim() - 1)), 1)
        if n_a != 1:
            raise NotImplementedError("oracle restates the n_actions == 1 path only")
        A_y = gp_math.sample_from_posterior(mean, cov, base).squeeze(-1)   # [S_a, *t, n_r, n_a]

    loss = topk_loss(A, A_y, spec.dist_weight, spec.dist_threshold)

    def cost_call(prev, cur, acc):
        k = len(used.cost_noise)
        nz = given.cost_noise[k] if given is not None else None
        if nz is None:
            nz = torch.randn(*cur.shape[:-1], prev.shape[-2], dtype=cur.dtype)
        used.cost_noise.append(nz)
        return spotlight_cost(prev, cur, spec.cost, nz, acc)

    cost = cost_call(prev_X[:, None, :], X_levels[0], prev_cost)
    for s in range(1, L):
        cur = X_levels[s]
        cost = cost + cost_call(X_levels[s - 1].unsqueeze(0).expand_as(cur), cur, cost + prev_cost)
    for a in range(n_a):
        cur = A[..., a:a + 1, :]
        cost = cost + cost_call(X_levels[-1].unsqueeze(0).expand_as(cur), cur, cost + prev_cost)
    cost = cost.squeeze(-1).sum(-1)

    while loss.dim() > 1:
        loss = loss.mean(0)
    while cost.dim() > 1:
        cost = cost.mean(0)
    return {"acqf_loss": loss, "acqf_cost": cost, "X": X_levels, "actions": A,
            "hidden_state": H_levels, "action_y": A_y, "draws": used}


# --------------------------------------------------------------------------
# myopic / KG baselines (acqfs.py:109-179 over BoTorch q* classes)
# --------------------------------------------------------------------------
def baseline_forward(name: str, gp: DenseGP, prev_X: Tensor, maps: List[Tensor], *,
                     base: Tensor, cost_spec: CostSpec, best_f: float = 0.0, prev_cost=0.0,
                     cost_noise: Optional[List[Tensor]] = None, num_fantasies: int = 64):
    """``base`` = Sobol-normal base samples [S] (seed 0 in the reference, actor.py:417-419)."""
    n_r, d = prev_X.shape
    actions = torch.cat(list(maps)).reshape(n_r, -1, d)         # row-major scramble for qKG (App. D#6)
    noises = list(cost_noise) if cost_noise is not None else None

    def nz(shape_like):
        if noises is None:
            return None
        return noises.pop(0)

    if name == "qMSL":
        # acqfs.py:94-99,137-145: k-step lookahead tree over the one-shot representation; ``base`` = list of the k
        # stages' Sobol base samples, ``num_fantasies`` = [f_1..f_k]
        nf = list(num_fantasies)
        sizes = [math.prod(nf[:i]) for i in range(len(nf) + 1)]
        assert actions.shape[1] == sum(sizes), (actions.shape, sizes)
        Xs, off = [], 0
        for i, qt in enumerate(sizes):
            Xi = actions[:, off:off + qt, :].permute(1, 0, 2).reshape(*nf[:i][::-1], n_r, 1, d)
            Xs.append(Xi)
            off += qt
        model = gp
        for i, Xi in enumerate(Xs[:-1]):
            mean, cov = model.posterior(Xi, observation_noise=True)
            ys = gp_math.sample_from_posterior(mean, cov, base[i].reshape(nf[i], *([1] * (mean.dim() - 1)), 1))
            model = model.condition_on_observations(Xi, ys)
        fmean, _ = model.posterior(Xs[-1])
        value = fmean.squeeze(-1).reshape(-1, n_r).mean(0)
        loss = -value
        pX, cost = prev_X[:, None, :], 0
        for cX in Xs:
            cost = cost + spotlight_cost(pX.expand_as(cX), cX, cost_spec, nz(cX), cost + prev_cost)
            pX = cX[None, ...]
        cost = cost.squeeze(-1).sum(-1)
        while cost.dim() > 1:
            cost = cost.mean(0)
        return {"acqf_loss": loss, "acqf_cost": cost, "actions": actions, "X": Xs}
    if name == "qKG":
        Xa = actions[:, :-num_fantasies, :]                     # [n_r, 1, d]
        Xf = actions[:, -num_fantasies:, :].permute(1, 0, 2).unsqueeze(-2)   # [nf, n_r, 1, d]
        mean, cov = gp.posterior(Xa, observation_noise=True)
        ys = gp_math.sample_from_posterior(mean, cov, base.reshape(-1, 1, 1))
        fm = gp.condition_on_observations(Xa, ys)
        fmean, _ = fm.posterior(Xf)
        value = fmean.max(-1).values.mean(0)
    else:
        mean, cov = gp.posterior(actions)
        s = gp_math.sample_from_posterior(mean, cov, base.reshape(-1, 1, 1)).squeeze(-1)   # [S,n_r,q]
        if name == "qEI":
            value = (s - best_f).clamp_min(0).max(-1).values.mean(0)
        elif name == "qPI":
            value = torch.sigmoid((s.max(-1).values - best_f) / 1e-3).mean(0)
        elif name == "qSR":
            value = s.max(-1).values.mean(0)
        elif name.startswith("qUCB"):
            beta = 0.1 if name == "qUCB" else float(name.split("-")[-1])
            m = s.mean(0)
            value = (m + math.sqrt(beta * math.pi / 2) * (s - m).abs()).max(-1).values.mean(0)
        else:
            raise NotImplementedError(name)
    loss = -value
    pX = prev_X[:, None, :]
    if name == "qKG":
        cX = actions[:, :-num_fantasies, :]
        cost = spotlight_cost(pX.expand_as(cX), cX, cost_spec, nz(cX), prev_cost)
        pX2, cX2 = cX, actions[:, -num_fantasies:, :]
        cost = cost + spotlight_cost(pX2.expand_as(cX2), cX2, cost_spec, nz(cX2), cost + prev_cost)
    else:
        cost = spotlight_cost(pX.expand_as(actions), actions, cost_spec, nz(actions), prev_cost)
    cost = cost.squeeze(-1).sum(-1)
    while cost.dim() > 1:
        cost = cost.mean(0)
    return {"acqf_loss": loss, "acqf_cost": cost, "actions": actions}


# --------------------------------------------------------------------------
# the optimisation loop (actor.py:439-734), amortised-TS and free-parameter forms
# --------------------------------------------------------------------------
def cosine_distance(a: Tensor, b: Tensor, eps: float = 1e-6) -> Tensor:
    """1 - nn.CosineSimilarity(dim=-1, eps)(a, b), flattened (actor.py:594-609)."""
    return (1 - F.cosine_similarity(a, b, dim=-1, eps=eps)).flatten()


def query_loop(spec: HESSpec, prev_X, prev_y, prev_h, params, *, budget: float, lr: float,
               n_iter: int, past_initial: bool, path=None, gp=None, design_base=None,
               action_base=None, prev_cost=0.0, draws_per_iter=None, record=None):
    """Restates actor.py:489-662 for the HES objective.

    ``params``: policy parameter dict (amortised; optimised in place) or a list of
    logit tensors (free form; sigmoid applied every iteration, actor.py:559-560).
    ``past_initial`` is ``iteration > n_initial_points`` (early-stop gate, :638).
    ``draws_per_iter[ep] = (draws_main, draws_original)`` or None to consume the RNG.
    Returns the reference's bookkeeping: best_* snapshots, argmin index, #iterations.
    """
    amortised = isinstance(params, dict)
    leaves = list(params.values()) if amortised else list(params)
    for p in leaves:
        p.requires_grad_(True)
    opt = torch.optim.AdamW(leaves, lr=lr)
    frozen = {k: v.detach().clone() for k, v in params.items()} if amortised else None
    best = dict(loss=torch.tensor([float("inf")]), cost=torch.tensor([float("inf")]), KL=0)
    snap = {}
    stall = 0
    n_done = 0
    for ep in range(n_iter):
        maps = params if amortised else [torch.sigmoid(p) for p in params]
        dm, do = (draws_per_iter[ep] if draws_per_iter is not None else (None, None))
        out = hes_forward(spec, prev_X, prev_y, prev_h, maps, path=path, gp=gp,
                          design_base=design_base, action_base=action_base,
                          prev_cost=prev_cost, enable_noise=False, draws=dm)
        if amortised:
            with torch.no_grad():
                orig = hes_forward(spec, prev_X, prev_y, prev_h, frozen, path=path, gp=gp,
                                   design_base=design_base, action_base=action_base,
                                   prev_cost=prev_cost, enable_noise=True, draws=do)
            KL = 0
            for s in range(spec.lookahead):
                KL = KL + cosine_distance(orig["X"][s], out["X"][s])
            KL = KL + cosine_distance(orig["actions"], out["actions"])
        else:
            KL = 0
        total = out["acqf_loss"] + (budget - out["acqf_cost"]).abs() + KL
        pick = torch.argmin(total)
        if ep == 0 or total[pick] < (best["loss"] + best["cost"] + best["KL"]).min():
            best = dict(loss=out["acqf_loss"].detach(), cost=out["acqf_cost"].detach(),
                        KL=KL.detach() if amortised else 0)
            snap = dict(X=[x.detach() for x in out["X"]], actions=out["actions"].detach(),
                        hidden=[h.detach() for h in out["hidden_state"]] if amortised else None)
            stall = 0
        elif past_initial or ep >= 200:
            stall += 1
        if record is not None:
            record(ep, out, total)
        grads = torch.autograd.grad(total.mean(), leaves, allow_unused=True)
        for p, g in zip(leaves, grads):
            p.grad = g
        opt.step()
        opt.zero_grad()
        n_done = ep + 1
        if stall > 50:
            break
    idx = torch.argmin(best["loss"] + best["cost"] + best["KL"])
    return dict(best_loss=best["loss"], best_cost=best["cost"], best_KL=best["KL"], idx=idx,
                next_X=snap["X"][0][idx].reshape(1, -1), actions=snap["actions"][..., idx, :, :],
                hidden_state=(snap["hidden"][0][idx:idx + 1] if amortised else None),
                n_iter=n_done, snapshot=snap)


# --------------------------------------------------------------------------
# Amortized pre-training (reference actor.py:106-220)
# --------------------------------------------------------------------------
def ball_points(parents: Tensor, radius: float, n: int) -> Tensor:
    """utils.py:313-335 ``generate_random_points_batch``: ``n`` uniform points in the radius-ball of every parent
    (float32 host draws cast to the parents' dtype); new leading dim ``n``."""
    batch, dim = parents.shape[:-1], parents.shape[-1]
    direction = torch.randn(n, *batch, dim).to(parents)
    direction = direction / torch.norm(direction, dim=-1, keepdim=True)
    r = torch.rand(n, *batch).pow(1 / dim).unsqueeze(-1).to(parents) * radius
    return parents.unsqueeze(0).expand(n, *batch, dim) + r * direction


def pretrain_targets(prev_x: Tensor, L: int, n: int, radius: float) -> Tensor:
    """actor.py:116-162 with every fantasy width 1 (amortized => Thompson sampling): L design levels, each a random
    ball point of its (clamped) parent, then the action level (clamped after the draw).  [L+1, n, d]."""
    d = prev_x.shape[-1]
    X: List[Tensor] = []
    for s in range(L):
        parent, cnt = (prev_x.reshape(1, d), n) if s == 0 else (X[-1], 1)
        X.append(torch.clamp(ball_points(parent, radius, cnt), max=0.99, min=0.01).reshape(-1, d))
    parent, cnt = (prev_x.reshape(1, d), n) if L == 0 else (X[-1], 1)
    X.append(torch.clamp(ball_points(parent, radius, cnt).reshape(-1, d), max=0.99, min=0.01))
    return torch.stack(X, dim=0)


def pretrain_loop(P: Dict[str, Tensor], gp, prev_x: Tensor, prev_y: Tensor, prev_h: Tensor, targets: Tensor, *,
                  n_epochs: int = 300, lr: float = 0.01, patience: int = 50, draws: Optional[List[Tensor]] = None,
                  record: Optional[list] = None):
    """actor.py:164-220: imitation of the random ball designs.  Per epoch: L design calls of the network, each
    followed by ONE draw of the surrogate's marginal posterior at the (detached, through ``sample``'s no_grad)
    output as the next observation - ``model(out.unsqueeze(-2)).sample(Size([1]))``, one ``randn`` of n values on
    the global generator per level - then the action call; L1 loss to the targets; AdamW(lr 0.01); early stop after
    ``patience`` epochs without a new best loss (checked after the step, actor.py:207-219).
    ``prev_h`` already holds ``+ hidden_noise``.  ``draws``: per-epoch [L, n] normals instead of the generator.
    Optimises ``P`` in place; returns the per-epoch losses."""
    L, n = targets.shape[0] - 1, targets.shape[1]
    leaves = list(P.values())
    for p in leaves:
        p.requires_grad_(True)
    opt = torch.optim.AdamW(leaves, lr=lr)
    best, stall, losses = float("inf"), 0, []
    for ep in range(n_epochs):
        outs, h = [], prev_h
        x, y = prev_x.expand(n, -1), prev_y.expand(n, -1)
        zs = []
        for j in range(L):
            out, h = policy_forward(P, x, y, h, "X")
            outs.append(out)
            x = out
            with torch.no_grad():
                mean, cov = gp.posterior(out.unsqueeze(-2))                   # [n, 1], [n, 1, 1]
                z = (torch.randn(1, *mean.shape, dtype=mean.dtype) if draws is None
                     else draws[ep][j].reshape(1, *mean.shape))
                zs.append(z.reshape(-1))
                root = gp_math.psd_safe_cholesky(cov)
                y = (mean + root.matmul(z.unsqueeze(-1)).squeeze(-1)).reshape(-1, 1)
        out, h = policy_forward(P, x, y, h, "A")
        outs.append(out)
        loss = torch.mean(abs(targets - torch.stack(outs, dim=0)))
        losses.append(float(loss))
        if record is not None:
            record.append(torch.stack(zs) if zs else torch.zeros(0, n, dtype=targets.dtype))
        if float(loss) < best:
            best, stall = float(loss), 0
        else:
            stall += 1
        loss.backward()
        opt.step()
        opt.zero_grad()
        if stall >= patience:
            break
    for p in leaves:
        p.requires_grad_(False)
    return losses
