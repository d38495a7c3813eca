"""Lookahead acquisition functions - host-side mirror of the reference's
``synthetic_experiment/acqfs.py`` (same class names, constructor arguments, ``forward``
signatures and return dictionaries), with the arithmetic in the sm_100a
kernels behind ``include/nonmyopia_b200.h``.

* ``qMultiStepHEntropySearch`` (acqfs.py:182-456): multi-step H-entropy search.
  - ``enable_ts`` + amortized map  -> one fused rollout kernel (``nmb_ts_rollout_fwd/bwd`` +
    ``nmb_policy_wgrad``);
  - ``enable_ts`` + free tensors   -> ``nmb_path_eval`` + ``nmb_tree_cost``;
  - Monte-Carlo fantasy tree        -> ``nmb_mc_tree_*`` (shared-root posterior recursion).
* ``qBOAcqf`` (acqfs.py:35-179): qEI / qPI / qSR / qUCB / qKG on the same posterior kernels.
* ``qLossFunctionTopK``, ``qCostFunction`` (acqfs.py:530-633): kept as ``nn.Module`` holders
  of the hyper-parameters (callers construct them by class + kwargs and call
  ``acqf.cost_function`` directly, actor.py:715, utils.py:145,169).

There is no CPU fallback: every ``forward`` needs CUDA tensors and the built library.
"""
from __future__ import annotations

import copy
import ctypes as C
from typing import Dict, List, Optional, Tuple, Type

import torch
import torch.nn as nn
from torch import Tensor

from . import _lib
from . import ts_engine as te
from .amortized_network import AmortizedNetwork, draw_rotation
from .gp import SurrogateGP, draw_thompson_path, sobol_normal_base_samples
from .path import ThompsonPath


# --------------------------------------------------------------------------------------
# loss / cost modules
# --------------------------------------------------------------------------------------
class qLossFunctionTopK(nn.Module):
    """``-mean_samples(sum_actions Y) - w * diversity(A)`` (reference acqfs.py:530-580)."""

    def __init__(self, dist_weight: float, dist_threshold: float) -> None:
        super().__init__()
        self.register_buffer("dist_weight", torch.as_tensor(dist_weight))
        self.register_buffer("dist_threshold", torch.as_tensor(dist_threshold))

    def forward(self, A: Tensor, Y: Tensor) -> Tensor:
        value = Y.sum(dim=-1).mean(dim=0)
        n_a = A.shape[-2]
        reward = 0
        if n_a >= 2:
            pair = torch.triu(torch.cdist(A.contiguous(), A.contiguous(), p=1.0))
            pair = torch.minimum(pair, self.dist_threshold.to(pair))
            reward = pair.sum((-1, -2)) / (n_a * (n_a - 1) / 2.0)
        return -value - self.dist_weight * reward


class qCostFunction(nn.Module):
    """Spotlight cost (reference acqfs.py:583-633).  Inside the acquisition kernels the
    hyper-parameters below are consumed directly; ``forward`` serves the stand-alone calls
    on a handful of points (actor.py:715, utils.py:169)."""

    def __init__(self, radius: float, k: float = 1, max_noise: float = 1e-5, p_norm: float = 2.0,
                 discount: float = 0.0, discount_threshold: float = -1.0) -> None:
        super().__init__()
        self.register_buffer("radius", torch.as_tensor(radius))
        self.k, self.max_noise, self.p_norm = k, max_noise, p_norm
        self.discount, self.discount_threshold = discount, discount_threshold
        # "host": the noise of a stand-alone call comes from torch's global CPU generator, as in the reference run on
        # the CPU (seed-stream parity, actor.py:715); "device": the CUDA generator.  Set by the owning acquisition.
        self.rng_mode = "host"

    def hypers(self) -> dict:
        return dict(radius=float(self.radius), k=self.k, max_noise=self.max_noise, p_norm=self.p_norm,
                    discount=self.discount, discount_threshold=self.discount_threshold)

    def forward(self, prev_X: Tensor, current_X: Tensor, previous_cost=None) -> Tensor:
        dist = torch.cdist(current_X, prev_X, p=self.p_norm)
        if self.rng_mode == "host" and dist.is_cuda:
            noise = torch.randn(dist.shape, dtype=dist.dtype).to(dist.device)
        else:
            noise = torch.randn_like(dist)
        hinge = torch.clamp_min(self.k * (dist - self.radius), 0.0) + noise * self.max_noise
        if self.discount > 0.0:
            over = (previous_cost + hinge > self.discount_threshold).float()
            hinge = hinge * (1 - self.discount * over)
        return hinge


class qCostFunctionEditDistance(nn.Module):
    """Discrete-environment cost (reference acqfs.py:636-680): out of scope - no sweep of the
    reference enables ``env_discretized``."""

    def __init__(self, radius: float, discount: float = 0.0, discount_threshold: float = -1.0,
                 previous_loss=None) -> None:
        super().__init__()
        self.register_buffer("radius", torch.as_tensor(radius))

    def forward(self, prev_X, current_X, previous_cost=None):
        raise NotImplementedError("discrete environments are outside the B200 hot-path scope")


# --------------------------------------------------------------------------------------
# samplers (reference acqfs.py:459-527 over BoTorch's NormalMCSampler)
# --------------------------------------------------------------------------------------
class _BaseSampler(nn.Module):
    """Holds the base samples ``z [n]`` of a q = 1 posterior: shared by every restart and
    every parent node, fixed across Adam iterations (collapse_batch_dims semantics)."""

    def __init__(self, sample_shape, seed: Optional[int] = None, **_deprecated) -> None:
        super().__init__()
        n = int(sample_shape[0] if isinstance(sample_shape, (tuple, list, torch.Size)) else sample_shape)
        self.sample_shape = torch.Size([n])
        # BoTorch draws the seed from the global host generator at construction
        self.seed = int(seed) if seed is not None else int(torch.randint(0, 1_000_000, (1,)).item())
        self.base_samples = None
        self._on_device = {}

    def base(self, dtype=torch.float64, device="cuda") -> Tensor:
        if self.base_samples is None:
            self.base_samples = self._draw(dtype)
        if self._on_device.get("src") is not self.base_samples:      # (re)assigned from outside: drop the copies
            self._on_device = {"src": self.base_samples}
        key = (str(device), dtype)
        if key not in self._on_device:          # one upload; later calls (CUDA-graph capture included) reuse it
            self._on_device[key] = self.base_samples.to(device=device, dtype=dtype).reshape(-1).contiguous()
        return self._on_device[key]


class SobolQMCNormalSampler(_BaseSampler):
    def _draw(self, dtype):
        return sobol_normal_base_samples(self.sample_shape[0], 1, self.seed, dtype).reshape(-1)


class PosteriorMeanSampler(_BaseSampler):
    """Zero base samples => the "sample" is the posterior mean (reference acqfs.py:459-490)."""

    def _draw(self, dtype):
        return torch.zeros(self.sample_shape[0], dtype=dtype)


def set_sampler_and_n_fantasy(sampler, n_fantasy: Optional[int]) -> Tuple[_BaseSampler, int]:
    """reference acqfs.py:493-527: n = 1 => mean sampler, else Sobol-normal."""
    if sampler is None:
        if n_fantasy is None:
            raise ValueError("Must specify `n_fantasy` if no `sampler` is provided.")
        sampler = PosteriorMeanSampler(n_fantasy) if n_fantasy == 1 else SobolQMCNormalSampler(n_fantasy)
    elif n_fantasy is not None:
        if sampler.sample_shape != torch.Size([n_fantasy]):
            raise ValueError("The sampler shape must match {n_fantasy}.")
    else:
        n_fantasy = sampler.sample_shape[0]
    return sampler, n_fantasy


def _host_cost_noise(sizes, like: Tensor) -> Tensor:
    """``cat([randn(n) for n in sizes])`` on torch's global CPU generator - the ``randn_like`` of one cost call per
    tree level (acqfs.py:421-441,624) - walked by the C side (``nmb_host_rng_randn_f64``, bit for bit) and uploaded."""
    if like.dtype != torch.float64:
        return torch.cat([torch.randn(int(n), dtype=like.dtype) for n in sizes]).to(like.device)
    from .host_rng import CpuStream
    cs = CpuStream()
    z = cs.randn_f64(sizes)
    cs.commit()
    return z.to(like.device)


# --------------------------------------------------------------------------------------
# fused TS-AM rollout as an autograd node
# --------------------------------------------------------------------------------------
class _TSAMRolloutFn(torch.autograd.Function):
    """(theta, prev_x, prev_y, prev_h) -> (loss, cost, X, A, HS); the backward pass runs the
    reverse rollout kernel and the split-K weight-gradient kernel."""

    @staticmethod
    def forward(ctx, theta, prev_x, prev_y, prev_h, eng: te.TSRollout, noisy_design, draws, seed, prev_cost):
        need_grad = any(t.requires_grad for t in (theta, prev_x, prev_y, prev_h))
        theta = theta.detach().contiguous()
        theta_t = eng.transpose(theta)
        buf = te.RolloutBuffers(eng.d, eng.H, eng.L, eng.N, eng.dtype, eng.device, with_save=need_grad)
        eng.seed = int(seed)
        prob = eng.problem(prev_x.detach().contiguous(), prev_y.detach().reshape(-1).contiguous(),
                           prev_h.detach().contiguous(), prev_cost=prev_cost)
        rq, rd, cn = draws if draws is not None else (None, None, None)
        ro = eng.rollout_struct(buf, theta, theta_t, noisy_design, rq, rd, cn)
        eng.forward(prob, [ro])
        ctx.eng, ctx.prob, ctx.ro, ctx.buf = eng, prob, ro, buf
        ctx.keep = (theta, theta_t, rq, rd, cn)
        ctx.want_prev = any(t.requires_grad for t in (prev_x, prev_y, prev_h))
        hs = buf.HS if buf.HS is not None else torch.empty(0, dtype=eng.dtype, device=eng.device)
        ctx.mark_non_differentiable(hs)
        return buf.loss, buf.cost, buf.X, buf.A, hs

    @staticmethod
    def backward(ctx, g_loss, g_cost, g_X, g_A, _g_hs):
        eng, buf = ctx.eng, ctx.buf
        z = lambda g, ref: (torch.zeros_like(ref) if g is None else g.contiguous())
        g_prev = (torch.empty(eng.N, eng.d + 1 + eng.H, dtype=eng.dtype, device=eng.device)
                  if ctx.want_prev else None)
        eng.backward(ctx.prob, ctx.ro, z(g_loss, buf.loss), z(g_cost, buf.cost),
                     None if g_X is None else g_X.contiguous(), None if g_A is None else g_A.contiguous(), g_prev)
        g_theta = eng.weight_grad(ctx.prob, buf)
        gx = gy = gh = None
        if g_prev is not None:
            d = eng.d
            gx, gy, gh = g_prev[:, :d], g_prev[:, d:d + 1], g_prev[:, d + 1:]
        return g_theta, gx, gy, gh, None, None, None, None, None


class _TreeCostFn(torch.autograd.Function):
    """Accumulated spotlight cost of a (possibly branching) tree of moves, per leaf.
    levels: list of [cnt_s, d] tensors, cnt_s = n_r * prod(nf[:s]) with the restart index
    fastest (reference layout [*nf_rev[:s], n_r, ., d]); node i of level s has parent
    i % cnt_{s-1}.  Returns acc [cnt_last]."""

    @staticmethod
    def forward(ctx, root, noise, hypers, prev_cost, *levels):
        lib = _lib.lib()
        n_lv = len(levels)
        cnt = [int(l.shape[0]) for l in levels]
        d = levels[0].shape[1]
        dt = levels[0].dtype
        dev = levels[0].device
        pts = torch.cat([l.detach().reshape(-1, d) for l in levels]).contiguous()
        counts = (C.c_int64 * n_lv)(*cnt)
        acc = torch.empty(cnt[-1], dtype=dt, device=dev)
        mult = torch.empty(sum(cnt), dtype=dt, device=dev)
        cs = te.cost_struct(hypers, prev_cost)
        root = root.detach().contiguous()
        _lib.check(lib.nmb_tree_cost_fwd(_lib.dtype_code(dt), d, n_lv, counts, C.byref(cs), _lib.ptr(root),
                                         C.c_int64(root.shape[0]), _lib.ptr(pts), _lib.ptr(noise), _lib.ptr(acc),
                                         _lib.ptr(mult), _lib.stream_ptr()), "nmb_tree_cost_fwd")
        ctx.save_for_backward(root, pts, mult)
        ctx.meta = (cnt, d, cs)
        return acc

    @staticmethod
    def backward(ctx, g_acc):
        root, pts, mult = ctx.saved_tensors
        cnt, d, cs = ctx.meta
        n_lv = len(cnt)
        counts = (C.c_int64 * n_lv)(*cnt)
        g_pts = torch.zeros_like(pts)
        g_root = torch.zeros_like(root)
        _lib.check(_lib.lib().nmb_tree_cost_bwd(_lib.dtype_code(pts.dtype), d, n_lv, counts, C.byref(cs),
                                                _lib.ptr(root), C.c_int64(root.shape[0]), _lib.ptr(pts),
                                                _lib.ptr(mult), _lib.ptr(g_acc.contiguous()), _lib.ptr(g_pts),
                                                _lib.ptr(g_root), _lib.stream_ptr()), "nmb_tree_cost_bwd")
        outs, o = [], 0
        for c in cnt:
            outs.append(g_pts[o:o + c])
            o += c
        return (g_root, None, None, None, *outs)


# --------------------------------------------------------------------------------------
# H-entropy search
# --------------------------------------------------------------------------------------
class qMultiStepHEntropySearch(nn.Module):
    """Multi-step H-entropy search with one-shot optimisation (reference acqfs.py:182-456)."""

    def __init__(self, model, loss_function_class: Type[nn.Module], loss_func_hypers: Dict,
                 cost_function_class: Type[nn.Module], cost_func_hypers: Dict, lookahead_steps: int,
                 n_actions: int, n_fantasy_at_design_pts: Optional[List[int]] = 64,
                 n_fantasy_at_action_pts: Optional[int] = 64, design_samplers=None, action_sampler=None,
                 enable_ts: Optional[bool] = False, **kwargs) -> None:
        super().__init__()
        self.model = model
        self._model = None
        self.f = None
        self.algo_lookahead_steps = lookahead_steps
        self.n_actions = n_actions
        self.cost_function = cost_function_class(**cost_func_hypers)
        self.loss_function = loss_function_class(**loss_func_hypers)
        self.enable_ts = enable_ts
        self.design_samplers, self.n_fantasy_at_design_pts = [], []
        for i in range(lookahead_steps):
            s = design_samplers[i] if design_samplers is not None else None
            s, nf = set_sampler_and_n_fantasy(sampler=s, n_fantasy=n_fantasy_at_design_pts[i])
            self.design_samplers.append(s)
            self.n_fantasy_at_design_pts.append(nf)
        self.action_sampler, self.n_fantasy_at_action_pts = set_sampler_and_n_fantasy(
            sampler=action_sampler, n_fantasy=n_fantasy_at_action_pts)
        # "host" (default): rotation angles / cost noise come from torch's global CPU generator in the reference's
        # order (seed-stream parity; Actor.query feeds them through a device ring at CUDA-graph speed);
        # "device": drawn in-kernel (Philox keyed by rng_seed, iteration, call and global chain index)
        self.rng_mode = kwargs.get("rng_mode", "host")
        self.cost_function.rng_mode = self.rng_mode
        self.rng_seed = None
        self._calls = 0
        self._gp = None
        self._engines = {}
        self._mc = None
        self._given_draws = None      # tests: explicit draws of the next forward
        self._fed_noise = None        # Actor.query, host stream: static device buffer refilled before every iteration

    # ---- surrogate state
    def gp(self) -> SurrogateGP:
        if self._gp is None:
            self._gp = SurrogateGP.from_model(self.model)
        return self._gp

    def dump_model(self, f: Optional[nn.Module] = None):
        """reference acqfs.py:258-266: freeze the surrogate for this BO step; TS: draw f ~ p(f|D)."""
        self._model = self.model
        self._gp = None
        self._engines = {}
        self._mc = None
        if self.enable_ts:
            self.f = draw_thompson_path(self.gp(), 1) if f is None else f
        # the in-kernel RNG needs a seed; in "host" mode nothing the reference does not draw may touch the generator
        self.rng_seed = int(torch.randint(0, 2**31 - 1, (1,)).item()) if self.rng_mode == "device" else 0
        self._calls = 0

    def clean_dump_model(self):
        self._model = None
        self._engines = {}
        self._mc = None
        if self.enable_ts:
            self.f = None

    def cost_noise_sizes(self, n_r: int) -> List[int]:
        """Element counts of the ``randn_like`` calls of one free-parameter ``forward`` (one per tree level)."""
        sizes, w = [], n_r
        for s in range(self.algo_lookahead_steps + 1):
            sizes.append(w)
            if s < self.algo_lookahead_steps and not self.enable_ts:
                w *= self.n_fantasy_at_design_pts[s]
        return sizes

    # ---- forward
    def forward(self, prev_X: Tensor, prev_y: Tensor, prev_hid_state: Tensor, maps, embedder: nn.Module = None,
                prev_cost: float = 0.0, enable_noise: bool = True) -> Dict[str, Tensor]:
        if embedder is not None:
            raise NotImplementedError("discrete environments are outside the B200 hot-path scope")
        if self.n_actions != 1:
            raise NotImplementedError("the B200 kernels implement n_actions == 1 (reference main.py:57)")
        if not prev_X.is_cuda:
            raise _lib.NmbError("nonmyopia_b200 ops need CUDA tensors (no CPU fallback)")
        prev_cost = float(prev_cost)
        if self.rng_seed is None:
            self.rng_seed = int(torch.randint(0, 2**31 - 1, (1,)).item()) if self.rng_mode == "device" else 0
        self._calls += 1
        if isinstance(maps, nn.Module):
            if not self.enable_ts:
                raise NotImplementedError("an amortized map needs enable_ts (the reference's KL term only "
                                          "broadcasts when every fantasy width is 1, SURVEY App. D#7)")
            return self._forward_ts_amortized(prev_X, prev_y, prev_hid_state, maps, prev_cost, enable_noise)
        if self.enable_ts:
            return self._forward_ts_free(prev_X, prev_y, prev_hid_state, maps, prev_cost)
        return self._forward_mc(prev_X, prev_y, prev_hid_state, maps, prev_cost)

    # ---- TS + amortized network: one fused kernel per direction
    def _ts_engine(self, path: ThompsonPath, d, H, N, dtype) -> te.TSRollout:
        key = (id(path), d, H, self.algo_lookahead_steps, N, dtype)
        eng = self._engines.get(key)
        if eng is None:
            S = path.n_paths
            if N % S:
                raise ValueError(f"{N} restarts cannot be split over {S} Thompson paths")
            eng = te.TSRollout(path, d, H, self.algo_lookahead_steps, N, self.cost_function.hypers(), dtype=dtype,
                               device=path.W.device, chains_per_path=N // S)
            self._engines = {key: eng}
        return eng

    def _host_draws(self, L, N, d, like, noisy_design):
        """Rotation blocks and cost noise of ONE forward in the reference's order, all on the global host generator
        (the reference run on the CPU, which the golden fixtures pin): one rotation draw per noisy policy call
        (amortized_network.py:136,156), then one ``randn_like`` per cost call (acqfs.py:624)."""
        rq = torch.zeros(L + 1, N, 4, dtype=like.dtype, device=like.device)
        rd = torch.zeros(L + 1, N, 2, dtype=torch.int32, device=like.device)
        for l in range(L + 1):
            if l == L or noisy_design:
                Q, dims = draw_rotation(N, d, like)
                rq[l], rd[l] = Q.reshape(N, 4), dims
        cn = _host_cost_noise([N] * (L + 1), like).reshape(L + 1, N)
        return rq.contiguous(), rd.contiguous(), cn.contiguous()

    def _forward_ts_amortized(self, prev_X, prev_y, prev_h, maps: AmortizedNetwork, prev_cost, enable_noise):
        if self.f is None:
            raise RuntimeError("dump_model() must be called before forward (no Thompson path drawn)")
        L = self.algo_lookahead_steps
        if L < 1:
            raise IndexError("HES needs lookahead_steps >= 1 (reference acqfs.py:437 indexes X_returned[-1])")
        N, d = prev_X.shape
        H = prev_h.shape[-1]
        eng = self._ts_engine(self.f, d, H, N, prev_X.dtype)
        theta = maps.flat_parameters()
        draws = self._given_draws
        self._given_draws = None
        if draws is None and self.rng_mode == "host":
            draws = self._host_draws(L, N, d, prev_X, enable_noise)
        seed = (self.rng_seed + 0x9E3779B1 * self._calls) & 0x7FFFFFFFFFFFFFFF
        fn = _TSAMRolloutFn.apply
        if not torch.is_grad_enabled():
            theta = theta.detach()
        loss, cost, X, A, HS = fn(theta, prev_X, prev_y, prev_h, eng, bool(enable_noise), draws, seed, prev_cost)
        X_ret = [X[s].reshape(*([1] * s), N, 1, d) for s in range(L)]
        hid = [HS[s + 1] for s in range(L)] if HS.numel() else [None] * L
        return {"acqf_loss": loss, "acqf_cost": cost, "X": X_ret,
                "actions": A.reshape(*([1] * L), N, self.n_actions, d), "hidden_state": hid}

    # ---- TS + free tensors: only the action's path value and the chain of moves matter
    def _forward_ts_free(self, prev_X, prev_y, prev_h, maps, prev_cost):
        if self.f is None:
            raise RuntimeError("dump_model() must be called before forward (no Thompson path drawn)")
        L = self.algo_lookahead_steps
        N, d = prev_X.shape
        levels = [maps[s].reshape(N, d) for s in range(L + 1)]
        draws = self._given_draws
        self._given_draws = None
        if draws is not None:
            noise = draws
        elif self._fed_noise is not None:
            noise = self._fed_noise.reshape(L + 1, N)
        elif self.rng_mode == "host":      # one randn_like per cost call on the host generator (acqfs.py:624)
            noise = _host_cost_noise([N] * (L + 1), prev_X)
        else:
            noise = torch.randn(L + 1, N, dtype=prev_X.dtype, device=prev_X.device)
        y_a = self.f.eval_points(levels[L].contiguous(), N // self.f.n_paths)
        acc = _TreeCostFn.apply(prev_X, noise.contiguous(), self.cost_function.hypers(), prev_cost, *levels)
        X_ret = [levels[s].reshape(*([1] * s), N, 1, d) for s in range(L)]
        return {"acqf_loss": -y_a, "acqf_cost": acc, "X": X_ret,
                "actions": levels[L].reshape(*([1] * L), N, self.n_actions, d),
                "hidden_state": [prev_h] * L}

    # ---- Monte-Carlo fantasy tree
    def _forward_mc(self, prev_X, prev_y, prev_h, maps, prev_cost):
        from .mc_tree import MCTree
        L = self.algo_lookahead_steps
        n_r, d = prev_X.shape
        nf = list(self.n_fantasy_at_design_pts)
        key = (n_r, d, prev_X.dtype)
        if self._mc is None or self._mc_key != key:
            bases = [s.base(prev_X.dtype, prev_X.device) for s in self.design_samplers]
            abase = self.action_sampler.base(prev_X.dtype, prev_X.device)
            self._mc = MCTree(self.gp(), nf, n_r, bases, abase)
            self._mc_key = key
        cnt = self._mc.counts
        levels = [maps[s].reshape(cnt[s], d) for s in range(L + 1)]
        draws = self._given_draws
        self._given_draws = None
        if draws is not None:
            noise = draws
        elif self._fed_noise is not None:
            noise = self._fed_noise
        elif self.rng_mode == "host":
            noise = _host_cost_noise(cnt, prev_X)
        else:
            noise = torch.randn(sum(cnt), dtype=prev_X.dtype, device=prev_X.device)
        leaf_loss = self._mc.apply(levels)                               # [cnt_L]  (-mean_samples y)
        acc = _TreeCostFn.apply(prev_X, noise.contiguous(), self.cost_function.hypers(), prev_cost, *levels)
        nf_rev = lambda s: nf[:s][::-1]
        X_ret = [levels[s].reshape(*nf_rev(s), n_r, 1, d) for s in range(L)]
        return {"acqf_loss": self._mc.restart_mean(leaf_loss), "acqf_cost": self._mc.restart_mean(acc), "X": X_ret,
                "actions": levels[L].reshape(*nf[::-1], n_r, self.n_actions, d),
                "hidden_state": [prev_h] * L}


# --------------------------------------------------------------------------------------
# myopic / KG baselines
# --------------------------------------------------------------------------------------
class qBOAcqf(nn.Module):
    """qEI / qPI / qSR / qUCB[-beta] / qKG over the shared-root posterior kernels
    (reference acqfs.py:35-179; BoTorch classes restated in SURVEY A.7-8)."""

    NAMES = ("qKG", "qEI", "qPI", "qSR", "qUCB", "qMSL")

    def __init__(self, name, model, loss_function_class, loss_func_hypers, cost_function_class, cost_func_hypers,
                 lookahead_steps: int, n_actions: int, n_fantasy_at_design_pts: Optional[List[int]] = [64],
                 sampler=None, best_f=None, **kwargs) -> None:
        super().__init__()
        self.name, self.model = name, model
        self.algo_lookahead_steps, self.n_actions = lookahead_steps, n_actions
        self.cost_function = cost_function_class(**cost_func_hypers)
        self.loss_function = loss_function_class(**loss_func_hypers)
        self.sampler = sampler
        self.best_f = None if best_f is None else float(best_f)
        self.num_fantasies = n_fantasy_at_design_pts[0] if n_fantasy_at_design_pts else 64
        if name == "qUCB":
            self.beta = 0.1
        elif name.startswith("qUCB"):
            self.beta = float(name.split("-")[-1])
        elif name == "qMSL":
            # BoTorch qMultiStepLookahead(batch_sizes=[1] * k, num_fantasies) with its defaults (reference
            # acqfs.py:94-99): one Sobol sampler per stage, seeded from the global host generator at construction;
            # zero stage values, posterior mean at the terminal points, uniform weights
            self.batch_sizes = [1] * lookahead_steps
            self.msl_fantasies = [int(f) for f in n_fantasy_at_design_pts][:lookahead_steps]
            if len(self.msl_fantasies) != lookahead_steps:
                raise ValueError("qMSL needs one fantasy width per lookahead step")
            self.msl_samplers = [SobolQMCNormalSampler(f) for f in self.msl_fantasies]
        elif name == "qNIPV":
            # the reference hands mc_points=0 to BoTorch's constructor (acqfs.py:101-104), whose register_buffer
            # accepts tensors only: the name cannot be instantiated upstream either - same error here
            raise TypeError("cannot assign 'int' object to buffer 'mc_points' (torch Tensor or None required)")
        elif name not in self.NAMES:
            raise NotImplementedError
        self._gp = None
        self._given_draws = None
        self._fed_noise = None
        self.rng_mode = kwargs.get("rng_mode", "host")
        self.cost_function.rng_mode = self.rng_mode

    def gp(self) -> SurrogateGP:
        if self._gp is None:
            self._gp = SurrogateGP.from_model(self.model)
        return self._gp

    def cost_noise_sizes(self, n_r: int) -> List[int]:
        """Element counts of the ``randn_like`` calls of one ``forward`` (acqfs.py:137-161)."""
        nf = self.num_fantasies
        if self.name == "qMSL":
            import math
            return [n_r * math.prod(self.msl_fantasies[:i]) for i in range(len(self.msl_fantasies) + 1)]
        return [n_r, n_r * nf * nf] if self.name == "qKG" else [n_r]

    def forward(self, prev_X: Tensor, prev_y: Tensor, maps, embedder: nn.Module = None, prev_cost: float = 0.0,
                **kwargs) -> Dict[str, Tensor]:
        from .mc_tree import baseline_forward
        if embedder is not None:
            raise NotImplementedError("discrete environments are outside the B200 hot-path scope")
        draws = self._given_draws
        self._given_draws = None
        return baseline_forward(self, prev_X, maps, float(prev_cost), draws)
