"""torch's CPU generator streams reproduced by the C-ABI library (csrc/host_rng.cu), bit for bit.

The reference consumes the global CPU generator inside every Adam iteration of ``Actor.query``
(``torch.rand`` + N x ``torch.randperm`` per noisy policy call, amortized_network.py:136,156; ``randn_like`` per
cost call, acqfs.py:624).  Drawing those with Python loops costs more than the whole GPU iteration, so the C side
walks the same mt19937 stream in bulk.  ``CpuStream`` wraps a copy of the generator state; ``commit`` writes a
state back into torch so that everything drawn afterwards continues exactly where the reference would.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np
import torch

from . import _lib

STATE_BYTES = 5056


def _u8p(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


class CpuStream:
    """A private copy of a torch CPU generator state that the C routines advance."""

    def __init__(self, generator: Optional[torch.Generator] = None):
        self.generator = generator
        st = torch.get_rng_state() if generator is None else generator.get_state()
        if st.numel() != STATE_BYTES:
            raise _lib.NmbError(f"unexpected torch CPU generator state size {st.numel()} (expected {STATE_BYTES})")
        self.state = st.numpy().copy()
        lib = _lib.lib()
        lib.nmb_host_rng_state_bytes.restype = C.c_int64
        assert lib.nmb_host_rng_state_bytes() == STATE_BYTES
        self.lib = lib

    def commit(self, state: Optional[np.ndarray] = None) -> None:
        """Install ``state`` (default: the current private state) as torch's generator state."""
        t = torch.from_numpy(np.ascontiguousarray(self.state if state is None else state).copy())
        if self.generator is None:
            torch.set_rng_state(t)
        else:
            self.generator.set_state(t)

    # ---- primitives (tests, free-parameter algorithms)
    def rand_f32(self, n: int) -> torch.Tensor:
        out = np.empty(n, dtype=np.float32)
        _lib.check(self.lib.nmb_host_rng_rand_f32(_u8p(self.state), C.c_int64(n),
                                                  out.ctypes.data_as(C.POINTER(C.c_float))), "nmb_host_rng_rand_f32")
        return torch.from_numpy(out)

    def randperm(self, n_perms: int, d: int) -> torch.Tensor:
        out = np.empty((n_perms, d), dtype=np.int64)
        _lib.check(self.lib.nmb_host_rng_randperm(_u8p(self.state), n_perms, d,
                                                  out.ctypes.data_as(C.POINTER(C.c_int64))), "nmb_host_rng_randperm")
        return torch.from_numpy(out)

    def randn_f64(self, sizes, out: Optional[torch.Tensor] = None, threads: int = 0) -> torch.Tensor:
        """``torch.cat([torch.randn(s, dtype=float64) for s in sizes])`` (optionally into a host tensor ``out``)."""
        sizes = [int(s) for s in sizes]
        arr = (C.c_int64 * len(sizes))(*sizes)
        if out is None:
            out = torch.empty(sum(sizes), dtype=torch.float64)
        if out.numel() != sum(sizes) or out.dtype != torch.float64 or out.is_cuda or not out.is_contiguous():
            raise ValueError("randn_f64: out must be a contiguous float64 host tensor of sum(sizes) elements")
        _lib.check(self.lib.nmb_host_rng_randn_f64(
            _u8p(self.state), len(sizes), arr, C.cast(C.c_void_p(out.data_ptr()), C.POINTER(C.c_double)),
            int(threads) if threads else min(16, os.cpu_count() or 1)), "nmb_host_rng_randn_f64")
        return out

    # ---- HES-TS-AM iteration stream
    def ts_am_stream(self, n_iters: int, L: int, N: int, d: int, out: Optional[dict] = None, threads: int = 0,
                     want_states: bool = False) -> dict:
        """Draws of ``n_iters`` Adam iterations (see ``nmb_host_rng_ts_am_stream``).  ``out`` may hold preallocated
        (pinned) tensors of the right shapes.  ``want_states``: also return the generator state after every
        iteration (``states`` [n_iters, 5056], numpy; serial path)."""
        c = L + 1
        if out is None:
            out = {}
        shapes = dict(u_main=((n_iters, N), torch.float32), dims_main=((n_iters, N, 2), torch.int32),
                      nz_main=((n_iters, c, N), torch.float64), u_orig=((n_iters, c, N), torch.float32),
                      dims_orig=((n_iters, c, N, 2), torch.int32), nz_orig=((n_iters, c, N), torch.float64))
        for k, (shp, dt) in shapes.items():
            if k not in out:
                out[k] = torch.empty(shp, dtype=dt)
            t = out[k]
            if tuple(t.shape) != shp or t.dtype != dt or not t.is_contiguous() or t.is_cuda:
                raise ValueError(f"ts_am_stream: buffer {k} must be a contiguous host tensor {shp} {dt}")
        states = np.empty((n_iters, STATE_BYTES), dtype=np.uint8) if want_states else None
        p = lambda k, ct: C.cast(C.c_void_p(out[k].data_ptr()), C.POINTER(ct))
        _lib.check(self.lib.nmb_host_rng_ts_am_stream(
            _u8p(self.state), n_iters, L, C.c_int64(N), d, p("u_main", C.c_float), p("dims_main", C.c_int32),
            p("nz_main", C.c_double), p("u_orig", C.c_float), p("dims_orig", C.c_int32), p("nz_orig", C.c_double),
            _u8p(states) if want_states else None, int(threads) if threads else min(16, os.cpu_count() or 1)), "nmb_host_rng_ts_am_stream")
        if want_states:
            out["states"] = states
        return out


class HostNoiseFeed:
    """Cost noise of the free-parameter algorithms in the reference's generator order, one Adam iteration at a
    time: ``sizes`` are the element counts of the ``randn_like`` calls of one ``acqf.forward`` (acqfs.py:421-441,
    624).  ``fill()`` makes the next iteration's draws resident in the static device buffer ``dev`` (what the
    captured CUDA graph of the iteration reads) and starts drawing the following iteration on a host thread; torch's
    generator always holds the state after the last iteration actually consumed."""

    def __init__(self, sizes, device, dtype=torch.float64, shard=None):
        """``shard = (n_restarts, lo, hi)``: restarts sharded over ranks - ``sizes`` are the GLOBAL call
        sizes, every rank walks the whole stream and keeps the entries of its restarts (the restart index is the
        fastest index of every level, so the shard is a strided slice)."""
        import threading
        self._threading = threading
        self.sizes = [int(s) for s in sizes]
        self.shard = shard
        n_glob = sum(self.sizes)
        if shard is not None:
            n_r, lo, hi = shard
            self.keep = [(s // n_r) * (hi - lo) for s in self.sizes]
            self.full = [torch.zeros(n_glob, dtype=torch.float64) for _ in range(2)]
        else:
            self.keep = self.sizes
        n = sum(self.keep)
        self.dev = torch.zeros(n, dtype=dtype, device=device)
        self.pinned = [torch.zeros(n, dtype=torch.float64).pin_memory() for _ in range(2)]
        self.cast = [torch.zeros(n, dtype=dtype).pin_memory() for _ in range(2)] if dtype != torch.float64 else None
        self.cpu = CpuStream()
        self.k = 0
        self._thread = None
        self._state_after = [None, None]
        self._copied = [None, None]
        self.bytes_per_iter = n * self.dev.element_size()

    def _generate(self, k: int) -> None:
        b = k & 1
        if self.shard is None:
            self.cpu.randn_f64(self.sizes, out=self.pinned[b])
        else:
            n_r, lo, hi = self.shard
            self.cpu.randn_f64(self.sizes, out=self.full[b])
            o = ok = 0
            for sz, kp in zip(self.sizes, self.keep):
                self.pinned[b][ok:ok + kp].copy_(self.full[b][o:o + sz].reshape(sz // n_r, n_r)[:, lo:hi].reshape(-1))
                o, ok = o + sz, ok + kp
        if self.cast is not None:
            self.cast[b].copy_(self.pinned[b])
        self._state_after[b] = self.cpu.state.copy()

    def fill(self) -> None:
        b = self.k & 1
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        else:
            self.cpu = CpuStream()          # first iteration: start from torch's current state
            self._generate(self.k)
        self.cpu.commit(self._state_after[b])
        src = self.pinned[b] if self.cast is None else self.cast[b]
        self.dev.copy_(src, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._copied[b] = ev
        self.k += 1
        nb = self.k & 1
        if self._copied[nb] is not None:
            self._copied[nb].synchronize()      # the other staging buffer has been copied out
        self._thread = self._threading.Thread(target=self._generate, args=(self.k,), daemon=True)
        self._thread.start()

    def close(self) -> None:
        if self._thread is not None:
            self._thread.join()         # speculative draw of an iteration that never ran: discarded
            self._thread = None
