"""Host driver of the fantasy-tree posterior kernels (``nmb_gp_whiten_*``, ``nmb_mc_level_*``,
``nmb_mc_gather_gu``): HES-MC leaf losses and the q = 1 posterior of the myopic / KG baselines,
each as one autograd node over the free design tensors (reference acqfs.py:363-373, 413-415 and
109-179).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence

import torch
from torch import Tensor

from . import _lib
from .gp import SurrogateGP

MAX_LEVELS = 16


class GpStruct(C.Structure):
    """``nmb_gp_t``"""
    _fields_ = [("d", C.c_int32), ("n", C.c_int32), ("kernel_id", C.c_int32), ("ldu", C.c_int32),
                ("ldl", C.c_int32), ("_pad", C.c_int32),
                ("amplitude", C.c_double), ("noise", C.c_double), ("mean_const", C.c_double),
                ("train_x", C.c_void_p), ("inv_ls", C.c_void_p), ("linv", C.c_void_p), ("linv_t", C.c_void_p),
                ("alpha", C.c_void_p), ("linv_tiles", C.c_void_p), ("linv_t_tiles", C.c_void_p)]


class TreeStruct(C.Structure):
    """``nmb_tree_t``"""
    _fields_ = [("n_levels", C.c_int32), ("obs_noise_fantasy", C.c_int32), ("counts", C.c_int64 * MAX_LEVELS),
                ("base", C.c_void_p * MAX_LEVELS), ("action_base_mean", C.c_double)]


class TreeState(C.Structure):
    """``nmb_tree_state_t``"""
    _names = ["pts", "U", "m0", "q", "W", "var", "mean", "cst", "sig", "bv", "leaf_loss",
              "Gw", "Gdot", "Gc", "Gb", "Gq", "Gm0", "gU", "g_pts", "Gscr"]
    _fields_ = [(n, C.c_void_p) for n in _names]


def _col_tiles(M: Tensor, width: int = 64, stride: int = 68) -> Tensor:
    """[ldl][ldl] -> [ceil(ldl/width)][ldl][stride]: 64-column blocks padded to the 68-double row stride of the
    shared-memory stage, so one contiguous bulk copy moves a whole operand chunk (csrc/whiten_mma.cuh)."""
    ldl = M.shape[0]
    nb = (ldl + width - 1) // width
    out = torch.zeros(nb, ldl, stride, dtype=M.dtype, device=M.device)
    for b in range(nb):
        w = min(width, ldl - b * width)
        out[b, :, :w] = M[:, b * width:b * width + w]
    return out


def gp_struct(gp: SurrogateGP, dtype) -> tuple:
    """``nmb_gp_t`` + the tensors it points to (kept alive by the caller)."""
    n, d = gp.n, gp.d
    ldl = (n + 7) // 8 * 8
    # zero-padded [ldl][ldl] triangular operands: 16-byte aligned rows for the bulk copies of whiten_mma.cuh
    linv = torch.zeros(ldl, ldl, dtype=dtype, device=gp.device)
    linv[:n, :n] = gp.Linv.to(dtype)
    keep = dict(train_x=gp.train_X.to(dtype).contiguous(), inv_ls=(1.0 / gp.lengthscale).to(dtype).contiguous(),
                linv=linv, linv_t=linv.T.contiguous(), alpha=gp.alpha.to(dtype).contiguous())
    keep["linv_tiles"], keep["linv_t_tiles"] = _col_tiles(keep["linv"]), _col_tiles(keep["linv_t"])
    ldu = (n + 3) // 4 * 4
    st = GpStruct(d=d, n=n, kernel_id=_lib.KERNEL_IDS[gp.kernel], ldu=ldu, ldl=ldl, amplitude=gp.amplitude, noise=gp.noise,
                  mean_const=gp.mean_const, **{k: v.data_ptr() for k, v in keep.items()})
    return st, keep


class MCTree:
    """Posterior over one fantasy tree.  ``nf`` = fantasy widths of the design levels, level s has
    ``n_r * prod(nf[:s])`` nodes with the restart index fastest (reference layout)."""

    def __init__(self, gp: SurrogateGP, nf: Sequence[int], n_r: int, design_bases: Sequence[Tensor],
                 action_base: Optional[Tensor], obs_noise_fantasy: bool = False, dtype=None):
        self.gp, self.nf, self.n_r = gp, list(nf), int(n_r)
        self.dtype = dtype or gp.dtype
        self.device = gp.device
        L = len(self.nf)
        if L + 1 > MAX_LEVELS:
            raise _lib.NmbError(f"fantasy trees deeper than {MAX_LEVELS - 1} levels are not supported")
        self.counts = [self.n_r * math.prod(self.nf[:s]) for s in range(L + 1)]
        self.offsets = [sum(self.counts[:s]) for s in range(L + 2)]
        self.total = self.offsets[-1]
        self.gp_struct, self._gp_keep = gp_struct(gp, self.dtype)
        self.bases = [b.to(self.device, self.dtype).reshape(-1).contiguous() for b in design_bases]
        for s, b in enumerate(self.bases):
            if b.numel() != self.nf[s]:
                raise ValueError(f"level {s}: {b.numel()} base samples for {self.nf[s]} fantasies")
        zbar = 0.0 if action_base is None else float(action_base.to(torch.float64).mean())
        self.tree = TreeStruct(n_levels=L + 1, obs_noise_fantasy=int(obs_noise_fantasy), action_base_mean=zbar)
        for s in range(L + 1):
            self.tree.counts[s] = self.counts[s]
        for s, b in enumerate(self.bases):
            self.tree.base[s] = b.data_ptr()
        self.lib = _lib.lib()

    # ---- workspace
    def _alloc(self, with_grad: bool) -> dict:
        z = lambda *shape: torch.empty(*shape, dtype=self.dtype, device=self.device)
        T, NL, ldu = self.total, len(self.counts), self.gp_struct.ldu
        ws = dict(U=z(T, ldu), m0=z(T), q=z(T), W=z(T, NL), var=z(T), mean=z(T), cst=z(T), sig=z(T), bv=z(T),
                  leaf_loss=z(self.counts[-1]))
        return ws

    def _scratch_elems(self, d: int) -> int:
        self.lib.nmb_mc_level_bwd_scratch_elems.restype = C.c_int64
        return int(self.lib.nmb_mc_level_bwd_scratch_elems(C.byref(self.tree), int(d)))

    def _state(self, pts: Tensor, ws: dict) -> TreeState:
        st = TreeState(pts=pts.data_ptr())
        for k, v in ws.items():
            setattr(st, k, v.data_ptr())
        return st

    def forward(self, pts: Tensor) -> dict:
        """``pts`` [total, d] (all levels back to back).  Returns the workspace (posterior mean / sigma of
        every node, leaf losses)."""
        lib, dt = self.lib, _lib.dtype_code(self.dtype)
        ws = self._alloc(False)
        sp = _lib.stream_ptr()
        _lib.check(lib.nmb_gp_whiten_fwd(C.byref(self.gp_struct), dt, _lib.ptr(pts), C.c_int64(self.total),
                                         _lib.ptr(ws["U"]), _lib.ptr(ws["m0"]), _lib.ptr(ws["q"]), sp),
                   "nmb_gp_whiten_fwd")
        st = self._state(pts, ws)
        for s in range(len(self.counts)):
            _lib.check(lib.nmb_mc_level_fwd(C.byref(self.gp_struct), C.byref(self.tree), C.byref(st), dt, s, sp),
                       "nmb_mc_level_fwd")
        return ws

    def backward(self, pts: Tensor, ws: dict, g_leaf: Tensor, g_leaf_sig: Optional[Tensor] = None) -> Tensor:
        lib, dt = self.lib, _lib.dtype_code(self.dtype)
        z = lambda *shape: torch.zeros(*shape, dtype=self.dtype, device=self.device)
        T, NL, ldu = self.total, len(self.counts), self.gp_struct.ldu
        ws = dict(ws)
        ws.update(Gw=z(T, NL), Gdot=z(T, NL), Gc=z(T), Gb=z(T), Gq=z(T), Gm0=z(T),
                  gU=torch.empty(T, ldu, dtype=self.dtype, device=self.device), g_pts=z(T, pts.shape[1]),
                  Gscr=torch.empty(self._scratch_elems(pts.shape[1]), dtype=self.dtype, device=self.device))
        st = self._state(pts, ws)
        sp = _lib.stream_ptr()
        for s in reversed(range(NL)):
            _lib.check(lib.nmb_mc_level_bwd(C.byref(self.gp_struct), C.byref(self.tree), C.byref(st), dt, s,
                                            _lib.ptr(g_leaf), _lib.ptr(g_leaf_sig), sp), "nmb_mc_level_bwd")
        _lib.check(lib.nmb_mc_gather_gu(C.byref(self.gp_struct), C.byref(self.tree), C.byref(st), dt, sp),
                   "nmb_mc_gather_gu")
        _lib.check(lib.nmb_gp_whiten_bwd(C.byref(self.gp_struct), dt, _lib.ptr(pts), C.c_int64(T), _lib.ptr(ws["gU"]),
                                         _lib.ptr(ws["Gm0"]), _lib.ptr(ws["g_pts"]), sp), "nmb_gp_whiten_bwd")
        return ws["g_pts"]

    def split(self, flat: Tensor) -> List[Tensor]:
        return [flat[self.offsets[s]:self.offsets[s + 1]] for s in range(len(self.counts))]

    # ---- autograd entry points
    def apply(self, levels: Sequence[Tensor]) -> Tensor:
        """HES-MC: per-leaf loss ``-(mean + sigma * mean(z_action))`` [counts[-1]]."""
        return _LeafLossFn.apply(self, *levels)

    def leaf_posterior(self, levels: Sequence[Tensor]):
        """(mean, sigma) of the leaf nodes, differentiable w.r.t. every level's points."""
        return _LeafPosteriorFn.apply(self, *levels)

    def restart_mean(self, leaf_vals: Tensor) -> Tensor:
        return _RestartMeanFn.apply(leaf_vals, self.n_r)


def _cat_levels(tree: MCTree, levels) -> Tensor:
    d = levels[0].shape[-1]
    for s, l in enumerate(levels):
        if l.numel() != tree.counts[s] * d:
            raise ValueError(f"level {s}: expected {tree.counts[s]} points, got {tuple(l.shape)}")
    return torch.cat([l.detach().reshape(-1, d) for l in levels]).to(tree.dtype).contiguous()


class _LeafLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tree: MCTree, *levels):
        pts = _cat_levels(tree, levels)
        ws = tree.forward(pts)
        ctx.tree, ctx.pts, ctx.ws = tree, pts, ws
        ctx.shapes = [l.shape for l in levels]
        return ws["leaf_loss"]

    @staticmethod
    def backward(ctx, g_leaf):
        tree = ctx.tree
        g = tree.backward(ctx.pts, ctx.ws, g_leaf.contiguous())
        return (None, *[x.reshape(shp) for x, shp in zip(tree.split(g), ctx.shapes)])


class _LeafPosteriorFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tree: MCTree, *levels):
        pts = _cat_levels(tree, levels)
        ws = tree.forward(pts)
        ctx.tree, ctx.pts, ctx.ws = tree, pts, ws
        ctx.shapes = [l.shape for l in levels]
        o = tree.offsets[-2]
        return ws["mean"][o:].clone(), ws["sig"][o:].clone()

    @staticmethod
    def backward(ctx, g_mean, g_sig):
        tree = ctx.tree
        zeros = lambda: torch.zeros(tree.counts[-1], dtype=tree.dtype, device=tree.device)
        g = tree.backward(ctx.pts, ctx.ws, zeros() if g_mean is None else g_mean.contiguous(),
                          zeros() if g_sig is None else g_sig.contiguous())
        return (None, *[x.reshape(shp) for x, shp in zip(tree.split(g), ctx.shapes)])


class _RestartMeanFn(torch.autograd.Function):
    """[*tree dims, n_r] -> [n_r] mean over the tree dims (reference acqfs.py:443-447)."""

    @staticmethod
    def forward(ctx, leaf: Tensor, n_r: int):
        leaf = leaf.contiguous()
        out = torch.empty(n_r, dtype=leaf.dtype, device=leaf.device)
        _lib.check(_lib.lib().nmb_tree_restart_mean(_lib.dtype_code(leaf.dtype), _lib.ptr(leaf), C.c_int64(leaf.numel()),
                                                    C.c_int64(n_r), _lib.ptr(out), _lib.stream_ptr()),
                   "nmb_tree_restart_mean")
        ctx.per = leaf.numel() // n_r
        ctx.n_r = n_r
        return out

    @staticmethod
    def backward(ctx, g):
        return (g / ctx.per).repeat(ctx.per), None


# --------------------------------------------------------------------------------------
# myopic / KG baselines (reference acqfs.py:109-179; BoTorch q* classes restated in SURVEY A.7-8)
# --------------------------------------------------------------------------------------
_UTILITY = {"qEI": 0, "qPI": 1, "qSR": 2, "qUCB": 3}


class _UtilityFn(torch.autograd.Function):
    """value = mean_s u(mean + sigma z_s) by the ``nmb_mc_utility`` kernel (value and both partials in one pass)."""

    @staticmethod
    def forward(ctx, mean, sig, base, which, best_f, beta):
        n = mean.numel()
        value, dm, ds = torch.empty_like(mean), torch.empty_like(mean), torch.empty_like(mean)
        _lib.check(_lib.lib().nmb_mc_utility(_lib.dtype_code(mean.dtype), which, C.c_double(best_f), C.c_double(beta),
                                             _lib.ptr(mean.contiguous()), _lib.ptr(sig.contiguous()), C.c_int64(n),
                                             _lib.ptr(base), int(base.numel()), _lib.ptr(value), _lib.ptr(dm),
                                             _lib.ptr(ds), _lib.stream_ptr()), "nmb_mc_utility")
        ctx.save_for_backward(dm, ds)
        return value

    @staticmethod
    def backward(ctx, g):
        dm, ds = ctx.saved_tensors
        return g * dm, g * ds, None, None, None, None


def _cost_pairwise(cost_module, prev: Tensor, cur: Tensor, noise: Tensor, previous_cost) -> Tensor:
    """qCostFunction.forward with the N(0,1) draw supplied (device tensor ops; only used for the [n_r, 64, 64]
    block the reference's qKG branch produces, acqfs.py:147-160)."""
    dist = torch.cdist(cur, prev, p=cost_module.p_norm)
    radius = float(cost_module.radius)
    hinge = torch.clamp_min(cost_module.k * (dist - radius), 0.0) + noise * cost_module.max_noise
    if cost_module.discount > 0.0:
        hinge = hinge * (1 - cost_module.discount * (previous_cost + hinge > cost_module.discount_threshold).float())
    return hinge


def baseline_forward(acqf, prev_X: Tensor, maps, prev_cost: float, draws=None):
    from .acqfs import _TreeCostFn
    if not prev_X.is_cuda:
        raise _lib.NmbError("nonmyopia_b200 ops need CUDA tensors (no CPU fallback)")
    n_r, d = prev_X.shape
    dt, dev = prev_X.dtype, prev_X.device
    gp = acqf.gp()
    actions = torch.concat(list(maps)).reshape(n_r, -1, d)      # row-major scramble for qKG
    hyp = acqf.cost_function.hypers()
    if acqf.name == "qMSL":
        return _msl_forward(acqf, prev_X, actions, prev_cost, draws)
    base = acqf.sampler.base(dt, dev)
    if acqf.name == "qKG":
        nfant = acqf.num_fantasies
        Xa = actions[:, :-nfant, :]
        Xf = actions[:, -nfant:, :]
        if Xa.shape[1] != 1:
            raise NotImplementedError("qKG on the B200 path takes one candidate per restart (reference actor.py:94-95)")
        key = ("kg", n_r, dt)
        if getattr(acqf, "_tree_key", None) != key:
            acqf._tree = MCTree(gp, [nfant], n_r, [base], None, obs_noise_fantasy=True, dtype=dt)
            acqf._tree_key = key
        tree = acqf._tree
        lv = [Xa.reshape(n_r, d), Xf.permute(1, 0, 2).reshape(nfant * n_r, d)]
        mean_leaf, _ = tree.leaf_posterior(lv)
        acqf_loss = -tree.restart_mean(mean_leaf)
        if draws is not None:
            nz = draws
        elif acqf._fed_noise is not None or acqf.rng_mode == "host":
            from .acqfs import _host_cost_noise
            flat = acqf._fed_noise if acqf._fed_noise is not None else _host_cost_noise(acqf.cost_noise_sizes(n_r), prev_X)
            nz = [flat[:n_r].reshape(n_r, 1, 1), flat[n_r:].reshape(n_r, nfant, nfant)]
        else:
            nz = [torch.randn(n_r, 1, 1, dtype=dt, device=dev), torch.randn(n_r, nfant, nfant, dtype=dt, device=dev)]
        pX = prev_X[:, None, :]
        cost = _cost_pairwise(acqf.cost_function, pX.expand_as(Xa), Xa, nz[0], prev_cost)
        cost = cost + _cost_pairwise(acqf.cost_function, Xa.expand_as(Xf), Xf, nz[1], cost + prev_cost)
        acqf_cost = cost.squeeze(dim=-1).sum(dim=-1)
        while acqf_cost.dim() > 1:
            acqf_cost = acqf_cost.mean(dim=0)
        X_ret = [Xa]
    else:
        if actions.shape[1] != 1:
            raise NotImplementedError("the B200 baselines implement q = n_actions = 1 (reference main.py:57)")
        key = ("q1", n_r, dt)
        if getattr(acqf, "_tree_key", None) != key:
            acqf._tree = MCTree(gp, [], n_r, [], None, dtype=dt)
            acqf._tree_key = key
        tree = acqf._tree
        pts = actions.reshape(n_r, d)
        mean, sig = tree.leaf_posterior([pts])
        which = _UTILITY["qUCB" if acqf.name.startswith("qUCB") else acqf.name]
        value = _UtilityFn.apply(mean, sig, base, which, acqf.best_f if acqf.best_f is not None else 0.0,
                                 getattr(acqf, "beta", 0.0))
        acqf_loss = -value
        if draws is not None:
            noise = draws
        elif acqf._fed_noise is not None:
            noise = acqf._fed_noise
        elif acqf.rng_mode == "host":
            from .acqfs import _host_cost_noise
            noise = _host_cost_noise([n_r], prev_X)
        else:
            noise = torch.randn(n_r, dtype=dt, device=dev)
        acqf_cost = _TreeCostFn.apply(prev_X, noise.reshape(-1).contiguous(), hyp, prev_cost, pts)
        X_ret = [actions]
    return {"acqf_loss": acqf_loss, "acqf_cost": acqf_cost, "X": X_ret, "actions": actions, "hidden_state": None}


def _msl_forward(acqf, prev_X: Tensor, actions: Tensor, prev_cost: float, draws=None):
    """qMSL (reference acqfs.py:94-99,137-145,166-167): BoTorch's k-step lookahead over the one-shot representation
    ``actions [b, q', d]``, q' = 1 + f_1 + f_1 f_2 + ...; stage i's points are ``X_i [f_i, .., f_1, b, 1, d]`` - the
    same restart-fastest tree layout as HES-MC - fantasies with observation noise, value = posterior mean of the
    k-times fantasised model at the terminal points averaged over the leaves: the ``nmb_mc_level_*`` kernels with a
    zero action base sample.  The reference runs this name in float32 (main.py:73-77): the dtype follows the inputs."""
    from .acqfs import _TreeCostFn, _host_cost_noise
    n_r, d = prev_X.shape
    dt, dev = prev_X.dtype, prev_X.device
    nf = acqf.msl_fantasies
    sizes = [math.prod(nf[:i]) for i in range(len(nf) + 1)]
    if actions.shape[1] != sum(sizes):
        raise ValueError(f"qMSL: {actions.shape[1]} points per restart, the tree {nf} needs {sum(sizes)}")
    key = ("msl", n_r, dt)
    if getattr(acqf, "_tree_key", None) != key:
        bases = [smp.base(dt, dev) for smp in acqf.msl_samplers]
        acqf._tree = MCTree(acqf.gp(), nf, n_r, bases, None, obs_noise_fantasy=True, dtype=dt)
        acqf._tree_key = key
    tree = acqf._tree
    levels, off = [], 0
    for qt in sizes:
        levels.append(actions[:, off:off + qt, :].permute(1, 0, 2).reshape(qt * n_r, d))
        off += qt
    mean_leaf, _ = tree.leaf_posterior(levels)
    acqf_loss = -tree.restart_mean(mean_leaf)
    cnt = tree.counts
    if draws is not None:
        noise = draws
    elif acqf._fed_noise is not None:
        noise = acqf._fed_noise
    elif acqf.rng_mode == "host":
        noise = _host_cost_noise(cnt, prev_X)
    else:
        noise = torch.randn(sum(cnt), dtype=dt, device=dev)
    acc = _TreeCostFn.apply(prev_X, noise.reshape(-1).contiguous(), acqf.cost_function.hypers(), prev_cost, *levels)
    X_ret = [levels[i].reshape(*nf[:i][::-1], n_r, 1, d) for i in range(len(sizes))]
    return {"acqf_loss": acqf_loss, "acqf_cost": tree.restart_mean(acc), "X": X_ret, "actions": actions,
            "hidden_state": None}
