"""Amortized pre-training of ``Actor.reset_parameters`` (reference actor.py:106-220) on the rollout kernels.

One epoch is a fixed launch sequence on one stream

    for every level l:   nmb_ts_rollout_fwd (imitation mode, level l)
                         [l < L]  nmb_gp_whiten_fwd on the level's outputs -> nmb_gp_marginal_sample -> y_obs[l]
    nmb_imitation_objective   (L1 loss to the random ball designs, adjoint seeds, early-stop bookkeeping)
    nmb_ts_rollout_bwd -> nmb_policy_wgrad -> nmb_policy_adamw_step (+ state advance)

captured once as a CUDA graph and replayed; the loss record, the best loss and the early-stop counter live in HBM and
the host polls the ``done`` flag once per chunk of epochs (the reference reads the loss on the host every epoch).
No ``torch.optim``, no ``@``: the network runs as the fused FP64 tensor-pipe policy step of the query loop.

Random numbers of the surrogate draws (``model(out.unsqueeze(-2)).sample()``, one ``randn`` of n values per level and
epoch): ``rng="host"`` walks torch's CPU generator in the reference's order in chunks (``nmb_host_rng_randn_f64``) and
feeds them through a device ring indexed by the device epoch counter; after the loop the generator is left exactly
where the reference's loop leaves it.  ``rng="device"``: in-kernel Philox.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional

import torch
from torch import Tensor

from . import _lib
from . import ts_engine as te
from .gp import SurrogateGP
from .host_rng import CpuStream
from .mc_tree import gp_struct
from .path import ThompsonPath


def supported(d: int, H: int, dtype, device) -> bool:
    """The imitation mode lives in the FP64 tensor-pipe rollout kernel (float64, even d, H a multiple of 4, >= 8)."""
    return (dtype == torch.float64 and d % 2 == 0 and H >= 8 and H % 4 == 0 and torch.device(device).type == "cuda")


class _NormalRing:
    """Double-buffered device ring [2K][L][N] of the per-epoch normals in the reference's generator order."""

    def __init__(self, L: int, N: int, chunk: int, device, generator=None):
        self.L, self.N, self.K = L, N, int(chunk)
        self.dev = torch.zeros(2 * self.K, max(L, 1), N, dtype=torch.float64, device=device)
        self.stage = [torch.zeros(self.K, max(L, 1), N, dtype=torch.float64).pin_memory() for _ in range(2)]
        self.cpu = CpuStream(generator)
        self.start_states = {0: self.cpu.state.copy()}
        self.copy_stream = torch.cuda.Stream(device=device)
        self.ready, self.free = [None, None], [None, None]
        self._thread, self._generated = None, -1

    def _generate(self, c: int) -> None:
        if self.L:
            self.cpu.randn_f64([self.N] * (self.L * self.K), out=self.stage[c & 1].view(-1))
        self.start_states[c + 1] = self.cpu.state.copy()
        self._generated = c

    def feed(self, ep: int) -> None:
        if ep % self.K:
            return
        c, h, K = ep // self.K, (ep // self.K) & 1, self.K
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        if self._generated < c:
            if self.ready[h] is not None:
                self.ready[h].synchronize()
            self._generate(c)
        main = torch.cuda.current_stream()
        with torch.cuda.stream(self.copy_stream):
            if self.free[h] is not None:
                self.copy_stream.wait_event(self.free[h])
            self.dev[h * K:(h + 1) * K].copy_(self.stage[h], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
            self.ready[h] = ev
        nh = (c + 1) & 1
        if self.ready[nh] is not None:
            self.ready[nh].synchronize()         # the other staging buffer has been copied out
        self._thread = threading.Thread(target=self._generate, args=(c + 1,), daemon=True)
        self._thread.start()
        main.wait_event(ev)

    def chunk_enqueued(self, ep_last: int) -> None:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.free[(ep_last // self.K) & 1] = ev

    def finish(self, n_done: int) -> None:
        """Leave the torch generator where the reference's loop leaves it after ``n_done`` epochs."""
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        c0 = n_done // self.K
        while c0 not in self.start_states:
            self._generate(max(self.start_states))
        tail = CpuStream(self.cpu.generator)
        tail.state = self.start_states[c0].copy()
        rem = n_done - c0 * self.K
        if rem and self.L:
            tail.randn_f64([self.N] * (self.L * rem))
        tail.commit()


class PretrainLoop:
    def __init__(self, gp: SurrogateGP, theta0: Tensor, prev_x: Tensor, prev_y: Tensor, prev_h: Tensor, targets: Tensor, *,
                 H: int, lr: float = 0.01, max_epochs: int = 300, patience: int = 50, rng: str = "host", seed: int = 0,
                 betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2, chunk: int = 25, generator=None):
        if rng not in ("host", "device"):
            raise ValueError(f"rng={rng!r}")
        dev, dt = gp.device, torch.float64
        L, N, d = targets.shape[0] - 1, targets.shape[1], targets.shape[2]
        if not supported(d, H, dt, dev):
            raise _lib.NmbError("pre-training kernels need float64, even d, hidden_dim >= 8 and a multiple of 4")
        self.gp, self.L, self.N, self.d, self.H = gp, L, N, d, H
        self.max_epochs, self.patience = int(max_epochs), int(patience)
        # the rollout structs want a Thompson path; the imitation mode never evaluates it
        z = lambda *s: torch.zeros(*s, dtype=dt, device=dev)
        dummy = ThompsonPath(z(1, d), z(1, 2), z(1, gp.n), gp.train_X, gp.lengthscale, kernel=gp.kernel,
                             outputscale=gp.outputscale, mean_const=gp.mean_const)
        self.eng = eng = te.TSRollout(dummy, d, H, L, N, dict(radius=0.0, max_noise=0.0), dtype=dt, device=dev, seed=seed)
        self.hp = te.AdamWStruct(lr=lr, beta1=betas[0], beta2=betas[1], eps=eps, weight_decay=weight_decay)
        self.theta = theta0.detach().to(dev, dt).clone().contiguous()
        self.theta_t = eng.transpose(self.theta)
        self.m, self.v = torch.zeros_like(self.theta), torch.zeros_like(self.theta)
        self.state = torch.zeros(te.LOOP_STATE_BYTES, dtype=torch.uint8, device=dev)
        eng.step_ptr = self.state.data_ptr()
        eng.done_ptr = self.state.data_ptr() + 8
        self.prev = (prev_x.to(dev, dt).contiguous(), prev_y.reshape(-1).to(dev, dt).contiguous(),
                     prev_h.to(dev, dt).contiguous())
        self.prob = eng.problem(*self.prev)
        self.buf = te.RolloutBuffers(d, H, L, N, dt, dev, with_save=True)
        self.targets = targets.to(dev, dt).contiguous()
        self.y_obs = z(max(L, 1), N)
        self.g_loss, self.g_cost, self.g_X, self.g_A = z(N), z(N), z(max(L, 1), N, d), z(N, d)
        self.hist = z(self.max_epochs)
        # shared-root whitening of the level outputs (mu, sigma^2 of the marginal posterior)
        self.gp_struct, self._gp_keep = gp_struct(gp, dt)
        self.U, self.m0, self.q = z(N, self.gp_struct.ldu), z(N), z(N)
        self.seed = int(seed)
        self.ring = _NormalRing(L, N, chunk, dev, generator) if rng == "host" else None
        self.chunk = int(chunk)
        self.graph = None
        self.n_launches_per_epoch = (L + 1) + 2 * L + 1 + 3 + 1

    def _iteration(self) -> None:
        eng, lib = self.eng, self.eng.lib
        st = _lib.stream_ptr()
        state_p = C.c_void_p(self.state.data_ptr())
        L, N = self.L, self.N
        ro = eng.rollout_struct(self.buf, self.theta, self.theta_t, noisy_design=False)
        ro.noisy_action = 0
        ro.y_obs = self.y_obs.data_ptr()
        for l in range(L + 1):
            ro.level_begin, ro.level_end = l, l + 1
            eng.forward(self.prob, [ro])
            if l == L:
                break
            _lib.check(lib.nmb_gp_whiten_fwd(C.byref(self.gp_struct), eng.dt, _lib.ptr(self.buf.X[l]), C.c_int64(N),
                                             _lib.ptr(self.U), _lib.ptr(self.m0), _lib.ptr(self.q), st),
                       "nmb_gp_whiten_fwd")
            ring = self.ring
            _lib.check(lib.nmb_gp_marginal_sample(
                eng.dt, C.c_double(self.gp.amplitude), C.c_double(self.gp.mean_const), _lib.ptr(self.m0), _lib.ptr(self.q),
                C.c_int64(N), C.c_void_p(0), _lib.ptr(ring.dev) if ring is not None else C.c_void_p(0),
                2 * ring.K if ring is not None else 0, max(L, 1), l, C.c_uint64(self.seed), C.c_int64(0),
                C.c_void_p(self.state.data_ptr()), C.c_void_p(self.state.data_ptr() + 8), _lib.ptr(self.y_obs[l]), st),
                "nmb_gp_marginal_sample")
        _lib.check(lib.nmb_imitation_objective(
            eng.dt, self.d, L, C.c_int64(N), _lib.ptr(self.buf.X), _lib.ptr(self.buf.A), _lib.ptr(self.targets),
            _lib.ptr(self.g_X), _lib.ptr(self.g_A), _lib.ptr(self.g_loss), _lib.ptr(self.g_cost), _lib.ptr(self.hist),
            C.c_int64(self.max_epochs), self.patience, state_p, st), "nmb_imitation_objective")
        eng.backward(self.prob, ro, self.g_loss, self.g_cost, self.g_X, self.g_A)
        partials = eng.weight_grad_partials(self.prob, self.buf)
        _lib.check(lib.nmb_policy_adamw_step(eng.dt, eng.d, eng.H, 1, C.byref(self.hp), _lib.ptr(self.theta),
                                             _lib.ptr(self.theta_t), _lib.ptr(self.m), _lib.ptr(self.v),
                                             _lib.ptr(partials), eng.seg_splits, state_p, 1, st),
                   "nmb_policy_adamw_step")

    def read_state(self) -> te.LoopState:
        raw = self.state[:C.sizeof(te.LoopState)].cpu().numpy().tobytes()
        return te.LoopState.from_buffer_copy(raw)

    def capture(self) -> None:
        if self.graph is not None:
            return
        # dry pass (sets kernel attributes, fills seg_splits): done = 1 turns the state-changing kernels into no-ops
        self.state[8:12] = torch.tensor([1, 0, 0, 0], dtype=torch.uint8, device=self.state.device)
        self._iteration()
        torch.cuda.synchronize()
        self.state.zero_()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._iteration()
        self.graph = g

    def run(self, use_graph: bool = True) -> te.LoopState:
        if use_graph:
            self.capture()
        K = self.ring.K if self.ring is not None else self.chunk
        ep = 0
        while ep < self.max_epochs:
            k = min(K, self.max_epochs - ep)
            for i in range(k):
                if self.ring is not None:
                    self.ring.feed(ep + i)
                if use_graph:
                    self.graph.replay()
                else:
                    self._iteration()
            ep += k
            if self.ring is not None:
                self.ring.chunk_enqueued(ep - 1)
            if self.read_state().done:
                break
        st = self.read_state()
        if self.ring is not None:
            self.ring.finish(int(st.step))
        return st

    def losses(self, n_done: int) -> Tensor:
        return self.hist[:n_done].cpu()
