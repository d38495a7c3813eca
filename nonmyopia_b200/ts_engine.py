"""Host driver of the fused HES-TS-AM kernels (C-ABI calls via ctypes).

``TSRollout`` owns the device buffers of one lookahead problem - N chains
(restarts x paths), L design steps + the action step - and exposes the three
device passes the objective needs:

* ``forward``   : policy -> rotation -> Thompson path -> spotlight cost for all levels
* ``backward``  : BPTT through the same chain, leaves per-layer adjoints in the workspace
* ``weight_grad``: split-K GEMMs over all (level, chain) nodes -> gradient of the flat
  parameter vector

It is the engine behind ``acqfs.qMultiStepHEntropySearch.forward`` (autograd) and
``actor.Actor.query`` (device-resident AdamW loop).  Reference semantics:
acqfs.py:277-456, amortized_network.py:97-166, actor.py:558-656.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch
from torch import Tensor

from . import _lib
from .path import ThompsonPath

# reference state_dict order of AmortizedNetwork (amortized_network.py:46-91)
PARAM_NAMES = [
    "embedding.weight", "embedding.bias", "prepro.2.weight", "prepro.2.bias",
    "rnn.weight_ih", "rnn.weight_hh", "rnn.bias_ih", "rnn.bias_hh",
    "postpro_A.0.weight", "postpro_A.0.bias", "postpro_A.3.weight", "postpro_A.3.bias",
    "postpro_A.6.weight", "postpro_A.6.bias",
    "postpro_X.0.weight", "postpro_X.0.bias", "postpro_X.3.weight", "postpro_X.3.bias",
    "postpro_X.6.weight", "postpro_X.6.bias",
]


def param_shapes(d: int, H: int, n_actions: int = 1):
    return [(H, d + 1), (H,), (H, H), (H,), (3 * H, H), (3 * H, H), (3 * H,), (3 * H,),
            (H, 2 * H), (H,), (H, H), (H,), (d * n_actions, H), (d * n_actions,),
            (H, 2 * H), (H,), (H, H), (H,), (d, H), (d,)]


def flatten_params(params: Dict[str, Tensor]) -> Tensor:
    return torch.cat([params[k].reshape(-1) for k in PARAM_NAMES])


def unflatten_params(flat: Tensor, d: int, H: int, n_actions: int = 1) -> Dict[str, Tensor]:
    out, o = {}, 0
    for name, shp in zip(PARAM_NAMES, param_shapes(d, H, n_actions)):
        n = 1
        for s in shp:
            n *= s
        out[name] = flat[o:o + n].view(*shp)
        o += n
    return out


class CostStruct(C.Structure):
    _fields_ = [("radius", C.c_double), ("k", C.c_double), ("max_noise", C.c_double), ("p_norm", C.c_double),
                ("discount", C.c_double), ("discount_threshold", C.c_double), ("prev_cost", C.c_double)]


class ProblemStruct(C.Structure):
    _fields_ = [("dtype", C.c_int32), ("d", C.c_int32), ("H", C.c_int32), ("L", C.c_int32),
                ("n_chains", C.c_int64), ("chains_per_path", C.c_int64),
                ("path", C.POINTER(_lib.PathStruct)), ("cost", CostStruct),
                ("prev_x", C.c_void_p), ("prev_y", C.c_void_p), ("prev_h", C.c_void_p),
                ("seed", C.c_uint64), ("step_ptr", C.c_void_p), ("done_flag", C.c_void_p),
                ("chain_offset", C.c_int64)]


class RolloutStruct(C.Structure):
    _fields_ = [("theta", C.c_void_p), ("theta_t", C.c_void_p),
                ("noisy_design", C.c_int32), ("noisy_action", C.c_int32),
                ("rot_q", C.c_void_p), ("rot_dims", C.c_void_p), ("cost_noise", C.c_void_p),
                ("rng_stream", C.c_uint32), ("_pad", C.c_uint32),
                ("X", C.c_void_p), ("A", C.c_void_p), ("Y", C.c_void_p), ("loss", C.c_void_p),
                ("cost", C.c_void_p), ("save", C.c_void_p),
                ("ring_cs", C.c_void_p), ("ring_dims", C.c_void_p), ("ring_noise", C.c_void_p),
                ("ring_len", C.c_int32), ("_pad2", C.c_int32),
                ("y_obs", C.c_void_p), ("level_begin", C.c_int32), ("level_end", C.c_int32)]


class LoopState(C.Structure):
    """``nmb_loop_state_t`` (device resident; this mirror is for host reads)."""
    _fields_ = [("step", C.c_int64), ("done", C.c_int32), ("done_next", C.c_int32), ("stall", C.c_int32),
                ("improved", C.c_int32), ("chosen_idx", C.c_int32), ("n_snapshots", C.c_int32),
                ("best_min", C.c_double), ("loss_mean", C.c_double), ("cost_mean", C.c_double),
                ("cand_total", C.c_double), ("cand_lck", C.c_double)]


LOOP_STATE_BYTES = 4096      # NMB_LOOP_STATE_BYTES: struct + scratch of the objective reduction


class AdamWStruct(C.Structure):
    _fields_ = [("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double),
                ("weight_decay", C.c_double)]


def cost_struct(hypers: dict, prev_cost: float = 0.0) -> CostStruct:
    """``qCostFunction`` hyper-parameters -> ``nmb_cost_t``.  The reference keeps the radius in a
    default-dtype (float32) buffer (acqfs.py:597): round it the same way."""
    radius32 = float(torch.tensor(float(hypers["radius"]), dtype=torch.float32))
    return CostStruct(radius=radius32, k=float(hypers.get("k", 1.0)), max_noise=float(hypers.get("max_noise", 1e-5)),
                      p_norm=float(hypers.get("p_norm", 2.0)), discount=float(hypers.get("discount", 0.0)),
                      discount_threshold=float(hypers.get("discount_threshold", -1.0)), prev_cost=float(prev_cost))


class RolloutBuffers:
    """Outputs (+ optional backward workspace) of one rollout."""

    def __init__(self, d, H, L, N, dtype, device, with_save: bool):
        self.X = torch.empty(max(L, 1), N, d, dtype=dtype, device=device)
        self.A = torch.empty(N, d, dtype=dtype, device=device)
        self.Y = torch.empty(L + 1, N, dtype=dtype, device=device)
        self.loss_cost = torch.empty(2, N, dtype=dtype, device=device)     # one buffer: one D2H copy reads both
        self.loss, self.cost = self.loss_cost[0], self.loss_cost[1]
        self.save = None
        self.HS = None
        if with_save:
            nbytes = _lib.lib().nmb_ts_save_bytes(_lib.dtype_code(dtype), d, H, L, C.c_int64(N))
            esz = 8 if dtype == torch.float64 else 4
            self.save = torch.empty(nbytes // esz, dtype=dtype, device=device)
            off = _lib.lib().nmb_ts_save_hidden_offset(d, H, L, C.c_int64(N))
            self.HS = self.save[off:off + (L + 2) * N * H].view(L + 2, N, H)


class TSRollout:
    def __init__(self, path: ThompsonPath, d: int, H: int, L: int, n_chains: int, cost_hypers: dict,
                 dtype=torch.float64, device="cuda", chains_per_path: Optional[int] = None, seed: int = 0):
        lib = _lib.lib()
        lib.nmb_ts_save_bytes.restype = C.c_int64
        lib.nmb_ts_save_hidden_offset.restype = C.c_int64
        lib.nmb_policy_param_count.restype = C.c_int64
        lib.nmb_policy_wgrad_max_splits.restype = C.c_int64
        self.lib = lib
        self.path, self.d, self.H, self.L, self.N = path, d, H, L, int(n_chains)
        self.dtype, self.device = dtype, torch.device(device)
        self.cpp = int(chains_per_path) if chains_per_path else self.N
        self.cost_hypers = dict(cost_hypers)
        self.seed = int(seed)
        self.n_params = int(lib.nmb_policy_param_count(d, H, 1))
        self.dt = _lib.dtype_code(dtype)
        self._path_struct = path.c_struct(dtype)
        self.max_splits = int(lib.nmb_policy_wgrad_max_splits(L, C.c_int64(self.N)))
        self.partials = None
        self.seg_splits = (C.c_int32 * 20)()
        self.step_ptr = None
        self.done_ptr = None
        self.chain_offset = 0       # global index of this rank's first chain (restarts sharded over GPUs)

    # ---- struct builders
    def problem(self, prev_x: Tensor, prev_y: Tensor, prev_h: Tensor, prev_cost: float = 0.0) -> ProblemStruct:
        self._keep = (prev_x, prev_y, prev_h)
        return ProblemStruct(
            dtype=self.dt, d=self.d, H=self.H, L=self.L, n_chains=self.N, chains_per_path=self.cpp,
            path=C.pointer(self._path_struct), cost=cost_struct(self.cost_hypers, prev_cost),
            prev_x=prev_x.data_ptr(), prev_y=prev_y.data_ptr(), prev_h=prev_h.data_ptr(), seed=self.seed,
            step_ptr=self.step_ptr or 0, done_flag=self.done_ptr or 0, chain_offset=int(self.chain_offset))

    def rollout_struct(self, buf: RolloutBuffers, theta: Optional[Tensor], theta_t: Tensor, noisy_design: bool,
                       rot_q: Optional[Tensor] = None, rot_dims: Optional[Tensor] = None,
                       cost_noise: Optional[Tensor] = None, rng_stream: int = 0, ring=None) -> RolloutStruct:
        """``ring`` = (cs [R, L+1, N, 2] float32, dims [R, L+1, N, 2] int32, noise [R, L+1, N]) device tensors of the
        host-stream ring (see ``nmb_ts_rollout_t``); iteration ``*step_ptr`` reads slot ``step % R``."""
        rc, rd, rn = ring if ring is not None else (None, None, None)
        return RolloutStruct(
            ring_cs=rc.data_ptr() if rc is not None else 0, ring_dims=rd.data_ptr() if rd is not None else 0,
            ring_noise=rn.data_ptr() if rn is not None else 0, ring_len=int(rc.shape[0]) if rc is not None else 0,
            theta=theta.data_ptr() if theta is not None else 0, theta_t=theta_t.data_ptr(),
            noisy_design=int(noisy_design), noisy_action=1,
            rot_q=rot_q.data_ptr() if rot_q is not None else 0,
            rot_dims=rot_dims.data_ptr() if rot_dims is not None else 0,
            cost_noise=cost_noise.data_ptr() if cost_noise is not None else 0,
            rng_stream=rng_stream, _pad=0, X=buf.X.data_ptr(), A=buf.A.data_ptr(), Y=buf.Y.data_ptr(),
            loss=buf.loss.data_ptr(), cost=buf.cost.data_ptr(),
            save=buf.save.data_ptr() if buf.save is not None else 0)

    # ---- device passes
    def transpose(self, theta: Tensor, theta_t: Optional[Tensor] = None) -> Tensor:
        if theta_t is None:
            theta_t = torch.empty_like(theta)
        _lib.check(self.lib.nmb_policy_transpose(self.dt, self.d, self.H, 1, _lib.ptr(theta), _lib.ptr(theta_t),
                                                 _lib.stream_ptr()), "nmb_policy_transpose")
        return theta_t

    def forward(self, prob: ProblemStruct, rollouts: List[RolloutStruct]) -> None:
        arr = (RolloutStruct * len(rollouts))(*rollouts)
        _lib.check(self.lib.nmb_ts_rollout_fwd(C.byref(prob), arr, len(rollouts), _lib.stream_ptr()),
                   "nmb_ts_rollout_fwd")

    def backward(self, prob: ProblemStruct, ro: RolloutStruct, g_loss: Tensor, g_cost: Tensor,
                 g_X: Optional[Tensor] = None, g_A: Optional[Tensor] = None, g_prev: Optional[Tensor] = None) -> None:
        _lib.check(self.lib.nmb_ts_rollout_bwd(C.byref(prob), C.byref(ro), _lib.ptr(g_loss), _lib.ptr(g_cost),
                                               _lib.ptr(g_X), _lib.ptr(g_A), _lib.ptr(g_prev), _lib.stream_ptr()),
                   "nmb_ts_rollout_bwd")

    def weight_grad_partials(self, prob: ProblemStruct, buf: RolloutBuffers) -> Tensor:
        if self.partials is None:
            self.partials = torch.empty(self.max_splits, self.n_params, dtype=self.dtype, device=self.device)
        _lib.check(self.lib.nmb_policy_wgrad(C.byref(prob), _lib.ptr(buf.save), _lib.ptr(self.partials),
                                             self.seg_splits, _lib.stream_ptr()), "nmb_policy_wgrad")
        return self.partials

    def weight_grad(self, prob: ProblemStruct, buf: RolloutBuffers) -> Tensor:
        partials = self.weight_grad_partials(prob, buf)
        grad = torch.empty(self.n_params, dtype=self.dtype, device=self.device)
        _lib.check(self.lib.nmb_policy_reduce_partials(self.dt, self.d, self.H, 1, _lib.ptr(partials),
                                                       self.seg_splits, _lib.ptr(grad), _lib.stream_ptr()),
                   "nmb_policy_reduce_partials")
        return grad

    def used_segments(self) -> List[bool]:
        return [self.seg_splits[i] > 0 for i in range(20)]
