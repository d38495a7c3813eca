"""Thompson-sample path ``f ~ p(f | D)`` evaluated by the ``nmb_path_eval`` kernel.

Host-side mirror of the object the reference stores in ``acqf.f``
(``draw_matheron_paths(model, Size([1]))``, reference acqfs.py:264): a picklable,
callable ``nn.Module`` (actor.py:508-516 pickles it, draw_trajectory_gif.py:123-125
calls it on 40,000 points) that is differentiable in its input.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch
from torch import Tensor, nn

from . import _lib


class _PathEvalFn(torch.autograd.Function):
    """y = f(x); one kernel launch returns the value and df/dx, backward is gy * df/dx."""

    @staticmethod
    def forward(ctx, x2d: Tensor, path: "ThompsonPath", points_per_path: int):
        N, d = x2d.shape
        y = torch.empty(N, dtype=x2d.dtype, device=x2d.device)
        need_grad = x2d.requires_grad
        dydx = torch.empty(N, d, dtype=x2d.dtype, device=x2d.device) if need_grad else None
        st = path.c_struct(x2d.dtype)
        rc = _lib.lib().nmb_path_eval(C.byref(st), _lib.dtype_code(x2d.dtype), _lib.ptr(x2d),
                                      C.c_int64(N), C.c_int64(points_per_path), _lib.ptr(y),
                                      _lib.ptr(dydx), _lib.stream_ptr())
        _lib.check(rc, "nmb_path_eval")
        if need_grad:
            ctx.save_for_backward(dydx)
        return y

    @staticmethod
    def backward(ctx, gy: Tensor):
        (dydx,) = ctx.saved_tensors
        return gy.unsqueeze(-1) * dydx, None, None


class ThompsonPath(nn.Module):
    """``f(x) = c + phi(x) . w + k(x, X_tr) . v`` for ``S`` jointly drawn paths.

    Buffers follow BoTorch's pathwise sample: ``W [M/2, d]``,
    ``w [S, M]``, ``v [S, n]``, ``train_X [n, d]``, ``lengthscale [d]``.
    ``forward(x[..., q, d])`` returns ``[..., q]`` for one path, ``[S, ..., q]`` for
    ``S > 1`` evaluated at the same points, like the reference object.
    """

    def __init__(self, W: Tensor, w: Tensor, v: Tensor, train_X: Tensor, lengthscale: Tensor, *,
                 kernel: str, outputscale: Optional[float], mean_const: float):
        super().__init__()
        d = train_X.shape[-1]
        self.kernel = kernel
        self.outputscale = None if outputscale is None else float(outputscale)
        self.mean_const = float(mean_const)
        ls = torch.as_tensor(lengthscale, dtype=train_X.dtype, device=train_X.device).reshape(-1)
        if ls.numel() == 1:
            ls = ls.expand(d)
        self.register_buffer("W", W.contiguous())
        self.register_buffer("w", w.reshape(-1, w.shape[-1]).contiguous())
        self.register_buffer("v", v.reshape(-1, v.shape[-1]).contiguous())
        self.register_buffer("train_X", train_X.contiguous())
        self.register_buffer("lengthscale", ls.contiguous())
        self._derived = {}

    # ---- derived device constants (phase matrix in half turns, 1/lengthscale), per dtype
    def _consts(self, dtype):
        key = (dtype, self.W.device, self.W.data_ptr(), self.lengthscale.data_ptr())
        c = self._derived.get("c")
        if c is None or c[0] != key:
            inv_ls = (1.0 / self.lengthscale.double())
            w_phase = (self.W.double() * inv_ls / math.pi)
            tens = dict(w_phase=w_phase.to(dtype).contiguous(), inv_ls=inv_ls.to(dtype).contiguous(),
                        w_feat=self.w.to(dtype).contiguous(), train_x=self.train_X.to(dtype).contiguous(),
                        v_update=self.v.to(dtype).contiguous())
            c = (key, tens)
            self._derived["c"] = c
        return c[1]

    def c_struct(self, dtype) -> _lib.PathStruct:
        t = self._consts(dtype)
        M = self.w.shape[-1]
        os_ = 1.0 if self.outputscale is None else self.outputscale
        return _lib.PathStruct(
            d=self.train_X.shape[-1], n_freq=M // 2, n_train=self.train_X.shape[0],
            n_paths=self.w.shape[0], kernel_id=_lib.KERNEL_IDS[self.kernel], _pad=0,
            feat_scale=math.sqrt(os_ * 2.0 / M), amplitude=os_, mean_const=self.mean_const,
            w_phase=t["w_phase"].data_ptr(), w_feat=t["w_feat"].data_ptr(), inv_ls=t["inv_ls"].data_ptr(),
            train_x=t["train_x"].data_ptr(), v_update=t["v_update"].data_ptr())

    @property
    def n_paths(self) -> int:
        return self.w.shape[0]

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_derived"] = {}
        return st

    def eval_points(self, x2d: Tensor, points_per_path: Optional[int] = None) -> Tensor:
        """Evaluate at a flat list of points; point i uses path i // points_per_path."""
        ppp = x2d.shape[0] if points_per_path is None else points_per_path
        return _PathEvalFn.apply(x2d.contiguous(), self, max(int(ppp), 1))

    def forward(self, x: Tensor) -> Tensor:
        d = x.shape[-1]
        lead = x.shape[:-1]
        x2d = x.reshape(-1, d)
        S = self.n_paths
        if S == 1:
            return self.eval_points(x2d).reshape(lead)
        n = x2d.shape[0]
        xs = x2d.unsqueeze(0).expand(S, n, d).reshape(S * n, d)
        return self.eval_points(xs, n).reshape(S, *lead)
