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

Random numbers.  The reference consumes the global CPU generator inside every noisy policy
call (``rand`` + per-sample ``randperm``, amortized_network.py:136,156) and inside every cost call
(``randn_like``, acqfs.py:624).  Three modes:

* ``rng="host"``: the reference's exact stream at CUDA-graph speed.  ``HostStreamRing`` walks the
  generator ahead of the loop in chunks (``nmb_host_rng_ts_am_stream``: torch's mt19937 stream bit
  for bit), uploads each chunk into one half of a device ring while the previous chunk computes,
  and the captured graph indexes the ring by the device step counter.  After the loop the torch
  generator is left exactly where the reference's loop would have left it (early stop included).
* ``rng="device"``: statistically equivalent draws in-kernel (Philox keyed by seed, iteration,
  call, GLOBAL chain index), no host work at all.
* explicit per-iteration draw tensors (``run(draws_per_iter=...)``): recorded streams of the
  golden fixtures, one eager launch sequence per iteration.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch
from torch import Tensor

from . import _lib
from . import ts_engine as te
from .host_rng import CpuStream


class HostStreamRing:
    """Double-buffered device ring of the per-iteration draws of HES-TS-AM in the reference's generator order.

    Chunk ``c`` = iterations ``[c K, (c + 1) K)`` lives in ring slots ``(c & 1) K ...``; it is generated on a host
    thread while chunk ``c - 1`` computes, staged in pinned memory and copied on a side stream.  ``n_global`` restarts
    are drawn (every rank walks the same stream) and the rank's slice ``own = (lo, hi)`` is uploaded."""

    def __init__(self, eng: te.TSRollout, n_global: int, own, chunk: int = 16, generator=None):
        import threading
        self._threading = threading
        self.L, self.d, self.N, self.Ng = eng.L, eng.d, eng.N, int(n_global)
        self.lo, self.hi = own
        assert self.hi - self.lo == self.N, (own, self.N)
        self.K = K = int(chunk)
        self.dtype, dev, c, N = eng.dtype, eng.device, eng.L + 1, eng.N
        ring = lambda: (torch.zeros(2 * K, c, N, 2, dtype=torch.float32, device=dev),
                        torch.zeros(2 * K, c, N, 2, dtype=torch.int32, device=dev),
                        torch.zeros(2 * K, c, N, dtype=eng.dtype, device=dev))
        self.main, self.orig = ring(), ring()
        pin = lambda *shape, dt: torch.zeros(*shape, dtype=dt).pin_memory()
        self.stage = [dict(main=(pin(K, c, N, 2, dt=torch.float32), pin(K, c, N, 2, dt=torch.int32),
                                 pin(K, c, N, dt=eng.dtype)),
                           orig=(pin(K, c, N, 2, dt=torch.float32), pin(K, c, N, 2, dt=torch.int32),
                                 pin(K, c, N, dt=eng.dtype)), raw={}) for _ in range(2)]
        self.cpu = CpuStream(generator)
        self.start_states = {0: self.cpu.state.copy()}      # chunk index -> generator state before the chunk
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.ready = [None, None]       # upload of the chunk in half h has landed
        self.free = [None, None]        # the replays that read half h have run
        self._thread, self._generated = None, -1
        self.bytes_per_iter = c * N * (2 * (8 + 8) + 2 * (8 if eng.dtype == torch.float64 else 4))

    # ---- host side
    def _generate(self, c: int) -> None:
        st = self.stage[c & 1]
        raw = self.cpu.ts_am_stream(self.K, self.L, self.Ng, self.d, out=st["raw"])
        self.start_states[c + 1] = self.cpu.state.copy()
        lo, hi, L = self.lo, self.hi, self.L
        # float32 angle, cos and sin in the reference's op order (amortized_network.py:136,144-153)
        th = (raw["u_main"][:, lo:hi] * 2 * torch.pi - torch.pi).contiguous()
        mcs, mdim, mnz = st["main"]
        mcs[:, L, :, 0], mcs[:, L, :, 1] = torch.cos(th), torch.sin(th)
        mdim[:, L] = raw["dims_main"][:, lo:hi]
        mnz.copy_(raw["nz_main"][:, :, lo:hi])
        th = (raw["u_orig"][:, :, lo:hi] * 2 * torch.pi - torch.pi).contiguous()
        ocs, odim, onz = st["orig"]
        ocs[..., 0], ocs[..., 1] = torch.cos(th), torch.sin(th)
        odim.copy_(raw["dims_orig"][:, :, lo:hi])
        onz.copy_(raw["nz_orig"][:, :, lo:hi])
        self._generated = c

    def _start_generate(self, c: int) -> None:
        h = c & 1
        if self.ready[h] is not None:
            self.ready[h].synchronize()      # the staging set of chunk c - 2 has been copied out
        self._thread = self._threading.Thread(target=self._generate, args=(c,), daemon=True)
        self._thread.start()

    def feed(self, ep: int) -> None:
        """Call before enqueueing iteration ``ep``: at a chunk boundary make the chunk's draws resident (generated
        ahead on the host thread), start generating the next chunk, and order the main stream behind the upload."""
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
            for ring, key in ((self.main, "main"), (self.orig, "orig")):
                for dst, src in zip(ring, self.stage[h][key]):
                    dst[h * K:(h + 1) * K].copy_(src, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
            self.ready[h] = ev
        self._start_generate(c + 1)
        main.wait_event(ev)

    def chunk_enqueued(self, ep_last: int) -> None:
        """Call after enqueueing the last iteration of a chunk (or of the run): its ring half may be refilled once
        the work enqueued so far has run."""
        h = (ep_last // self.K) & 1
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.free[h] = ev

    def finish(self, n_done: int) -> None:
        """Leave the torch generator where the reference's loop leaves it after ``n_done`` iterations."""
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        c0 = n_done // self.K
        while c0 not in self.start_states:      # loop ended before this chunk was ever needed
            self._generate(max(self.start_states) )
        tail = CpuStream(self.cpu.generator)
        tail.state = self.start_states[c0].copy()
        rem = n_done - c0 * self.K
        if rem:
            tail.ts_am_stream(rem, self.L, self.Ng, self.d)
        tail.commit()


class TSAMLoop:
    def __init__(self, eng: te.TSRollout, theta0: Tensor, prev_x: Tensor, prev_y: Tensor, prev_h: Tensor, *,
                 budget: float, lr: float, past_initial: bool, prev_cost: float = 0.0, n_global: Optional[int] = None,
                 max_iter: int = 500, record_history: bool = False, record_trajectory: bool = False,
                 patience: int = 50, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, group=None, exchange: str = "auto", rng: str = "device",
                 own=None, chunk: int = 16, generator=None):
        if rng not in ("device", "host"):
            raise ValueError(f"rng={rng!r}")
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
        self._host_done = None    # event behind a step_host call's result copies
        self._host_stage = None   # device staging buffers of step_host's loss / cost read-back
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
        # host-stream ring (the reference's generator order at graph speed); own = this rank's restarts of n_global
        self.ring = None
        if rng == "host":
            self.ring = HostStreamRing(eng, self.n_global, own if own is not None else (0, N), chunk, generator)
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
        # rollout_fwd, objective, snapshot, rollout_bwd, wgrad, adamw (the last CTA advances the state); sharded:
        # + decide, reduce (NCCL) or + publish, wait/decide (peer memory)
        # the peer exchange as one kernel (NMB_PEER_UNFUSED=1: publish / wait_decide / adamw_step as three launches)
        self.fused_exchange = self.px is not None and os.environ.get("NMB_PEER_UNFUSED") is None
        self.n_launches_per_iter = 6 if (group is None or self.fused_exchange) else 8

    # ------------------------------------------------------------------ one iteration
    def _iteration(self, draws=None, mark=None) -> None:
        mark = mark or (lambda name: None)
        eng, lib = self.eng, self.eng.lib
        st = _lib.stream_ptr()
        if draws is None and self.ring is not None:
            rm = eng.rollout_struct(self.main, self.theta, self.theta_t, noisy_design=False, ring=self.ring.main)
            ro = eng.rollout_struct(self.orig, self.theta0, self.theta0_t, noisy_design=True, ring=self.ring.orig)
        elif draws is None:
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
        # (measured: the snapshot on a forked stream beside the reverse rollout is not faster - the fork / join of the
        # captured graph costs what the 3 us copy costs: 0.4119 vs 0.4103 ms per step)
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
        if self.fused_exchange:
            # publish -> per-slice rendezvous -> rank-ordered sum + AdamW -> decision + advance: one launch
            _lib.check(lib.nmb_peer_exchange_adamw_step(
                px, eng.dt, _lib.ptr(partials), eng.d, eng.H, 1, eng.seg_splits, C.byref(self.hp), _lib.ptr(self.theta),
                _lib.ptr(self.theta_t), _lib.ptr(self.m), _lib.ptr(self.v), state_p, int(self.past_initial),
                self.patience, st), "nmb_peer_exchange_adamw_step")
            _lib.check(lib.nmb_snapshot_if_improved(eng.dt, 6, self._src, self._dst, self._cnt, state_p, st),
                       "nmb_snapshot_if_improved")
            mark("adamw")
            return
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
            # the ranks' epochs / flags are out of lock step now: drop the cached exchange buffer so that the next
            # loop rendezvouses on a fresh, zeroed one
            if self.px is not None:
                self.px.evict()
            raise _lib.NmbError("peer exchange timed out: a rank did not publish within NMB_PEER_TIMEOUT_MS (default 10 s; loop stopped)")
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
        if self.group is not None:
            # capture / instantiate skew between ranks must not eat into the exchange's spin timeout
            torch.cuda.synchronize()
            self._dist.barrier(self.group)

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
        if self.ring is not None:
            poll_every = self.ring.K        # one poll per ring chunk
        ep = 0
        while ep < n_iter:
            k = min(poll_every, n_iter - ep)
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
                         ready=torch.cuda.Event(), which=0, staged=False, ahead=False, staged_ptrs=None)

    def step_host(self, host_x: Tensor, host_y: Tensor, host_h: Tensor, host_out: Tensor, next_inputs=None) -> None:
        """One Adam iteration driven from HOST buffers (the reference reads loss / cost on the host every
        iteration, actor.py:580-581,618-619): pinned ``prev_X / prev_y / prev_hid_state`` -> HBM, one graph
        replay (or eager launch sequence), this iteration's per-restart loss and cost -> pinned ``host_out[2, N]``,
        then a stream synchronise.  The graph reads the inputs through fixed device buffers, so the copies
        land in place.  ``next_inputs = (host_x, host_y, host_h)`` of the FOLLOWING call may be passed: they are copied
        on a side stream into the other input-buffer set while this iteration runs (every step's copy still happens,
        it just no longer sits in front of the kernels), and - with the in-kernel random stream - the following iteration
        is launched right behind this step's result copies, so the GPU does not idle while the host returns and calls
        again.  Passing ``next_inputs`` therefore COMMITS the caller to that next call (with those very tensors); the
        last call of a sequence passes ``None``."""
        main = torch.cuda.current_stream()
        if self.ring is not None:
            ep = self._host_steps = getattr(self, "_host_steps", -1) + 1
            self.ring.feed(ep)
        alt = self._alt
        if next_inputs is not None and alt is None and self.graph is not None:
            self._enable_alt_inputs()
            alt = self._alt
        if alt is not None and alt.get("ahead"):
            # this step's iteration is already in the stream: the previous call launched it behind its own result
            # copies, as soon as the prefetched inputs had landed (the GPU never waits for the host between steps)
            if alt["staged_ptrs"] != (host_x.data_ptr(), host_y.data_ptr(), host_h.data_ptr()):
                raise _lib.NmbError("step_host: the inputs differ from the next_inputs promised by the previous call "
                                    "(their iteration has already been launched)")
            which = alt["which"]
            alt["ahead"] = False
        else:
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
        # this step's per-restart loss and cost -> host (in stream order behind the iteration)
        # Staged through a device buffer (one small device-to-device copy on this stream) and read back on the COPY
        # stream: a device-to-host copy in the compute stream would put two compute <-> copy-engine hand-overs (~18 us
        # measured) between this iteration and the next one.  Two staging buffers: buffer `which` is rewritten two
        # calls later, after this call has waited for its read-back.
        if self._host_stage is None:
            self._host_stage = [torch.empty_like(self.main.loss_cost) for _ in range(2)]
            self._host_iter_done = torch.cuda.Event()
            self._host_done = torch.cuda.Event()
            self._host_copy_stream = torch.cuda.Stream(device=self.state.device)
        stage = self._host_stage[which]
        stage.copy_(self.main.loss_cost, non_blocking=True)
        self._host_iter_done.record(main)
        with torch.cuda.stream(self._host_copy_stream):
            self._host_copy_stream.wait_event(self._host_iter_done)
            if host_out.is_contiguous() and tuple(host_out.shape) == tuple(stage.shape):
                host_out.copy_(stage, non_blocking=True)
            else:
                host_out[0].copy_(stage[0], non_blocking=True)
                host_out[1].copy_(stage[1], non_blocking=True)
            self._host_done.record(self._host_copy_stream)
        if alt is not None:
            alt["staged"] = False
            if next_inputs is not None:
                # the other set was last read by the previous step's iteration, whose results the previous call waited for
                nx, ny, nh = next_inputs
                dx, dy, dh = alt["prev"] if which == 0 else self.prev
                with torch.cuda.stream(alt["side"]):
                    dx.copy_(nx.view_as(dx), non_blocking=True)
                    dy.copy_(ny.view_as(dy), non_blocking=True)
                    dh.copy_(nh.view_as(dh), non_blocking=True)
                    alt["ready"].record(alt["side"])
                alt["which"], alt["staged"] = 1 - which, True
                alt["staged_ptrs"] = (nx.data_ptr(), ny.data_ptr(), nh.data_ptr())
                if self.ring is None and self.graph is not None:
                    # launch the FOLLOWING iteration now (behind the result copies above and the input prefetch):
                    # passing next_inputs commits the caller to that step
                    main.wait_event(alt["ready"])
                    (alt["graph"] if which == 0 else self.graph).replay()
                    alt["ahead"] = True
        if self.ring is not None and (self._host_steps + 1) % self.ring.K == 0:
            self.ring.chunk_enqueued(self._host_steps)
        self._host_done.synchronize()                       # results of THIS step are on the host

    # ------------------------------------------------------------------ results (actor.py:661-733)
    def result(self):
        tot = self.best_loss + self.best_cost + self.best_KL
        idx = torch.argmin(tot)
        return dict(idx=idx, best_loss=self.best_loss, best_cost=self.best_cost, best_KL=self.best_KL,
                    next_X=self.best_X[0][idx].reshape(1, -1), actions=self.best_A[idx].reshape(1, -1),
                    hidden_state=self.best_HS[1][idx:idx + 1], theta=self.theta)
