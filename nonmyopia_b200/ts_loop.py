"""Device-resident AdamW loop for HES-TS-AM (reference actor.py:489-662).

One Adam iteration is a fixed sequence of launches on one stream

    rollout_fwd(main + frozen "original" rollout, one launch)
    -> objective (KL, total, backward seeds, argmin, snapshot decision, early-stop counter)
    -> snapshot_if_improved (+ best_min refresh)
    -> rollout_bwd -> policy_wgrad -> adamw_step (+ state advance)

with every piece of loop state (iteration counter, early-stop counter, best-restart
snapshot, Adam moments) resident in HBM, so the sequence is captured once in a CUDA
graph and replayed with no host synchronisation; the host only polls the ``done`` flag
every few replays.  The reference synchronises four times per iteration
(actor.py:580-581, 618-619).

Random numbers.  The reference consumes the CPU generator inside every noisy policy call
(per-sample ``randperm``) and the device generator inside every cost call.  ``draws=None``
(default, fast) draws statistically identical numbers in-kernel (Philox keyed by seed,
iteration, call, chain); passing per-iteration draw tensors reproduces a recorded stream
bit for bit (parity tests) at the price of one eager launch sequence per iteration.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
from torch import Tensor

from . import _lib
from . import ts_engine as te


class TSAMLoop:
    def __init__(self, eng: te.TSRollout, theta0: Tensor, prev_x: Tensor, prev_y: Tensor, prev_h: Tensor, *,
                 budget: float, lr: float, past_initial: bool, prev_cost: float = 0.0, n_global: Optional[int] = None,
                 max_iter: int = 500, record_history: bool = False, record_trajectory: bool = False,
                 patience: int = 50, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, group=None, exchange: str = "auto"):
        self.eng = eng
        d, H, L, N = eng.d, eng.H, eng.L, eng.N
        dt, dev = eng.dtype, eng.device
        self.budget, self.past_initial, self.max_iter = float(budget), bool(past_initial), int(max_iter)
        self.n_global = int(n_global) if n_global else N
        self.patience = int(patience)
        self.hp = te.AdamWStruct(lr=lr, beta1=betas[0], beta2=betas[1], eps=eps, weight_decay=weight_decay)
        # parameters: live copy, frozen "original maps" (actor.py:525-535), Adam moments
        self.theta = theta0.detach().to(dev, dt).clone().contiguous()
        self.theta_t = eng.transpose(self.theta)
        self.theta0 = self.theta.clone()
        self.theta0_t = self.theta_t.clone()
        self.m = torch.zeros_like(self.theta)
        self.v = torch.zeros_like(self.theta)
        # device loop state
        self.state = torch.zeros(te.LOOP_STATE_BYTES, dtype=torch.uint8, device=dev)
        eng.step_ptr = self.state.data_ptr()            # offsetof(step) == 0
        eng.done_ptr = self.state.data_ptr() + 8        # offsetof(done) == 8
        self.prev = (prev_x.to(dev, dt).contiguous(), prev_y.reshape(-1).to(dev, dt).contiguous(),
                     prev_h.to(dev, dt).contiguous())
        self.prob = eng.problem(*self.prev, prev_cost=prev_cost)
        self._prev_cost = prev_cost
        self._alt = None          # second input buffer set + graph (step_host with prefetch)
        self.main = te.RolloutBuffers(d, H, L, N, dt, dev, with_save=True)
        self.orig = te.RolloutBuffers(d, H, L, N, dt, dev, with_save=False)
        z = lambda *s: torch.zeros(*s, dtype=dt, device=dev)
        self.KL, self.total = z(N), z(N)
        self.g_loss, self.g_cost, self.g_X, self.g_A = z(N), z(N), z(max(L, 1), N, d), z(N, d)
        self.best_loss, self.best_cost, self.best_KL = z(N), z(N), z(N)
        self.best_X, self.best_A, self.best_HS = z(max(L, 1), N, d), z(N, d), z(L + 2, N, H)
        self.hist_loss = z(self.max_iter, N) if record_history else None
        self.hist_cost = z(self.max_iter, N) if record_history else None
        self.hist_X = z(self.max_iter, max(L, 1), N, d) if record_trajectory else None
        srcs = [self.main.loss, self.main.cost, self.KL, self.main.X, self.main.A, self.main.HS]
        dsts = [self.best_loss, self.best_cost, self.best_KL, self.best_X, self.best_A, self.best_HS]
        self._src = (C.c_void_p * 6)(*[t.data_ptr() for t in srcs])
        self._dst = (C.c_void_p * 6)(*[t.data_ptr() for t in dsts])
        self._cnt = (C.c_int64 * 6)(*[t.numel() for t in srcs])
        self.graph = None
        # process group over which the restarts are sharded (None: single GPU)
        self.group = group
        self.px, self.exchange = None, "none"
        if group is not None:
            import torch.distributed as dist
            from .peer import PeerExchange
            self._dist = dist
            self._cand = self.state[56:72].view(torch.float64)
            self.grad = torch.zeros_like(self.theta)
            self._one_split = (C.c_int32 * 20)()
            # per-iteration exchange (gradient SUM + decision MIN): own kernels over NVLink peer memory, fused with
            # the AdamW update ("peer"), or two NCCL all-reduces ("nccl"); "auto" falls back to NCCL with the reason
            if exchange not in ("auto", "peer", "nccl"):
                raise ValueError(f"exchange={exchange!r}")
            self.exchange, self.exchange_note = "nccl", None
            if exchange != "nccl":
                self.px, why = PeerExchange.try_get(group, self.theta.numel(), dt, dev)
                if self.px is not None:
                    self.exchange = "peer"
                elif exchange == "peer":
                    raise _lib.NmbError(f"peer exchange unavailable: {why}")
                else:
                    self.exchange_note = why
        # rollout_fwd, objective, snapshot, rollout_bwd, wgrad, adamw, advance; sharded: + decide, reduce (NCCL) or
        # + publish, wait/decide (peer memory; the advance is the exchange's own)
        self.n_launches_per_iter = 7 if group is None else 9

    # ------------------------------------------------------------------ one iteration
    def _iteration(self, draws=None, mark=None) -> None:
        mark = mark or (lambda name: None)
        eng, lib = self.eng, self.eng.lib
        st = _lib.stream_ptr()
        if draws is None:
            rm = eng.rollout_struct(self.main, self.theta, self.theta_t, noisy_design=False, rng_stream=0)
            ro = eng.rollout_struct(self.orig, self.theta0, self.theta0_t, noisy_design=True, rng_stream=1)
        else:
            mq, md, mn, oq, od, on = draws
            rm = eng.rollout_struct(self.main, self.theta, self.theta_t, False, mq, md, mn, 0)
            ro = eng.rollout_struct(self.orig, self.theta0, self.theta0_t, True, oq, od, on, 1)
        eng.forward(self.prob, [rm, ro])
        mark("ts_rollout_fwd")
        state_p = C.c_void_p(self.state.data_ptr())
        _lib.check(lib.nmb_ts_objective(
            eng.dt, eng.d, eng.L, C.c_int64(eng.N), C.c_int64(self.n_global), C.c_double(self.budget),
            int(self.past_initial), self.patience, _lib.ptr(self.main.loss), _lib.ptr(self.main.cost), _lib.ptr(self.main.X),
            _lib.ptr(self.main.A), _lib.ptr(self.orig.X), _lib.ptr(self.orig.A), _lib.ptr(self.KL),
            _lib.ptr(self.total), _lib.ptr(self.g_loss), _lib.ptr(self.g_cost), _lib.ptr(self.g_X),
            _lib.ptr(self.g_A), _lib.ptr(self.hist_loss), _lib.ptr(self.hist_cost), _lib.ptr(self.hist_X),
            C.c_int64(self.max_iter), int(self.group is not None), state_p, st), "nmb_ts_objective")
        if self.px is not None:
            return self._finish_iteration_peer(rm, state_p, st, mark)
        if self.group is not None:
            # restarts sharded over ranks: one MIN all-reduce of (min total, min loss+cost+KL) keeps the
            # snapshot / early-stop decision identical on every rank
            self._dist.all_reduce(self._cand, op=self._dist.ReduceOp.MIN, group=self.group)
            _lib.check(lib.nmb_loop_decide(state_p, int(self.past_initial), self.patience, st), "nmb_loop_decide")
        _lib.check(lib.nmb_snapshot_if_improved(eng.dt, 6, self._src, self._dst, self._cnt, state_p, st),
                   "nmb_snapshot_if_improved")
        mark("objective+snapshot")
        eng.backward(self.prob, rm, self.g_loss, self.g_cost, self.g_X, self.g_A)
        mark("ts_rollout_bwd")
        partials = eng.weight_grad_partials(self.prob, self.main)
        mark("policy_wgrad")
        if self.group is not None:
            # data-parallel policy gradient: reduce the split-K partials locally, SUM over ranks (C2)
            _lib.check(lib.nmb_policy_reduce_partials(eng.dt, eng.d, eng.H, 1, _lib.ptr(partials), eng.seg_splits,
                                                      _lib.ptr(self.grad), st), "nmb_policy_reduce_partials")
            self._dist.all_reduce(self.grad, op=self._dist.ReduceOp.SUM, group=self.group)
            partials = self.grad
            for i in range(20):
                self._one_split[i] = 1 if eng.seg_splits[i] > 0 else 0
            splits = self._one_split
        else:
            splits = eng.seg_splits
        _lib.check(lib.nmb_policy_adamw_step(eng.dt, eng.d, eng.H, 1, C.byref(self.hp), _lib.ptr(self.theta),
                                             _lib.ptr(self.theta_t), _lib.ptr(self.m), _lib.ptr(self.v),
                                             _lib.ptr(partials), splits, state_p, 1, st),
                   "nmb_policy_adamw_step")
        mark("adamw")

    def _finish_iteration_peer(self, rm, state_p, st, mark) -> None:
        """Rest of a sharded iteration with the exchange over NVLink peer memory: the decision needs the other
        ranks' minima and the update their gradients, so both wait for ONE rendezvous after the local backward."""
        eng, lib = self.eng, self.eng.lib
        px = self.px.byref()
        mark("objective+snapshot")
        eng.backward(self.prob, rm, self.g_loss, self.g_cost, self.g_X, self.g_A)
        mark("ts_rollout_bwd")
        partials = eng.weight_grad_partials(self.prob, self.main)
        mark("policy_wgrad")
        _lib.check(lib.nmb_peer_publish(px, eng.dt, _lib.ptr(partials), eng.d, eng.H, 1, eng.seg_splits, state_p, st),
                   "nmb_peer_publish")
        _lib.check(lib.nmb_peer_wait_decide(px, eng.dt, state_p, int(self.past_initial), self.patience, st),
                   "nmb_peer_wait_decide")
        _lib.check(lib.nmb_snapshot_if_improved(eng.dt, 6, self._src, self._dst, self._cnt, state_p, st),
                   "nmb_snapshot_if_improved")
        _lib.check(lib.nmb_peer_adamw_step(px, eng.dt, eng.d, eng.H, 1, C.byref(self.hp), _lib.ptr(self.theta),
                                           _lib.ptr(self.theta_t), _lib.ptr(self.m), _lib.ptr(self.v), state_p, st),
                   "nmb_peer_adamw_step")
        mark("adamw")

    def read_state(self) -> te.LoopState:
        raw = self.state[:C.sizeof(te.LoopState)].cpu().numpy().tobytes()
        st = te.LoopState.from_buffer_copy(raw)
        if st.stall == -777:
            raise _lib.NmbError("peer exchange timed out: a rank did not publish within ~2 s (loop stopped)")
        return st

    def capture(self) -> None:
        """Capture one iteration in a CUDA graph (after a dry pass that sets kernel attributes)."""
        if self.graph is not None:
            return
        # dry pass: done = 1 makes every kernel return at once; restores done = 0 afterwards
        self.state[8:12] = torch.tensor([1, 0, 0, 0], dtype=torch.uint8, device=self.state.device)
        self._iteration()
        torch.cuda.synchronize()
        self.state.zero_()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._iteration()
        self.graph = g

    def run(self, n_iter: Optional[int] = None, draws_per_iter: Optional[List] = None, poll_every: int = 16,
            use_graph: bool = True) -> te.LoopState:
        n_iter = self.max_iter if n_iter is None else n_iter
        if draws_per_iter is not None:
            for ep in range(min(n_iter, len(draws_per_iter))):
                self._iteration(draws_per_iter[ep])
                if self.read_state().done:
                    break
            return self.read_state()
        if use_graph:
            self.capture()
        ep = 0
        while ep < n_iter:
            k = min(poll_every, n_iter - ep)
            for _ in range(k):
                if use_graph:
                    self.graph.replay()
                else:
                    self._iteration()
            ep += k
            if self.read_state().done:
                break
        return self.read_state()

    def _enable_alt_inputs(self) -> None:
        """Second set of input buffers with its own captured graph, so the host can copy the NEXT step's inputs
        (side stream) while this step computes."""
        if self._alt is not None:
            return
        if self.graph is None:
            self.capture()
        prev2 = tuple(torch.empty_like(t) for t in self.prev)
        prob2 = self.eng.problem(*prev2, prev_cost=self._prev_cost)
        self.eng._keep = self.prev                      # the engine keeps the primary set alive; prev2 is kept here
        prob1, self.prob = self.prob, prob2
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._iteration()
        self.prob = prob1
        self._alt = dict(prev=prev2, prob=prob2, graph=g, side=torch.cuda.Stream(device=self.state.device),
                         ready=torch.cuda.Event(), which=0, staged=False)

    def step_host(self, host_x: Tensor, host_y: Tensor, host_h: Tensor, host_out: Tensor, next_inputs=None) -> None:
        """One Adam iteration driven from HOST buffers (the reference reads loss / cost on the host every
        iteration, actor.py:580-581,618-619): pinned ``prev_X / prev_y / prev_hid_state`` -> HBM, one graph
        replay (or eager launch sequence), this iteration's per-restart loss and cost -> pinned ``host_out[2, N]``,
        then a stream synchronise.  The graph reads the inputs through fixed device buffers, so the copies
        land in place.  ``next_inputs = (host_x, host_y, host_h)`` of the FOLLOWING call may be passed: they are copied
        on a side stream into the other input-buffer set while this iteration runs (every step's copy still happens,
        it just no longer sits in front of the kernels)."""
        main = torch.cuda.current_stream()
        alt = self._alt
        if next_inputs is not None and alt is None and self.graph is not None:
            self._enable_alt_inputs()
            alt = self._alt
        if alt is not None and alt["staged"]:
            main.wait_event(alt["ready"])               # this step's inputs were prefetched into set `which`
            which = alt["which"]
        else:
            which = alt["which"] if alt is not None else 0
            px, py, ph = self.prev if which == 0 else alt["prev"]
            px.copy_(host_x.view_as(px), non_blocking=True)
            py.copy_(host_y.view_as(py), non_blocking=True)
            ph.copy_(host_h.view_as(ph), non_blocking=True)
        if self.graph is not None:
            (self.graph if which == 0 else alt["graph"]).replay()
        else:
            self._iteration()
        if alt is not None:
            alt["staged"] = False
            if next_inputs is not None:
                # the other set was last read by the previous call, which ended with a synchronise
                nx, ny, nh = next_inputs
                dx, dy, dh = alt["prev"] if which == 0 else self.prev
                with torch.cuda.stream(alt["side"]):
                    dx.copy_(nx.view_as(dx), non_blocking=True)
                    dy.copy_(ny.view_as(dy), non_blocking=True)
                    dh.copy_(nh.view_as(dh), non_blocking=True)
                    alt["ready"].record(alt["side"])
                alt["which"], alt["staged"] = 1 - which, True
        host_out[0].copy_(self.main.loss, non_blocking=True)
        host_out[1].copy_(self.main.cost, non_blocking=True)
        main.synchronize()

    # ------------------------------------------------------------------ results (actor.py:661-733)
    def result(self):
        tot = self.best_loss + self.best_cost + self.best_KL
        idx = torch.argmin(tot)
        return dict(idx=idx, best_loss=self.best_loss, best_cost=self.best_cost, best_KL=self.best_KL,
                    next_X=self.best_X[0][idx].reshape(1, -1), actions=self.best_A[idx].reshape(1, -1),
                    hidden_state=self.best_HS[1][idx:idx + 1], theta=self.theta)
