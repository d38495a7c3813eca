"""Peer-mapped exchange buffers for the restart-sharded loop (SURVEY 8e, C2 + C3).

``torch.distributed._symmetric_memory`` is used for the plumbing only: it allocates memory every rank of the
group can map over NVLink and hands back the peers' addresses.  The exchange itself - publish, wait + decide,
gradient sum fused into AdamW - is the ``nmb_peer_*`` kernels (csrc/optim.cu), captured in the per-iteration
CUDA graph like every other launch of the loop.
"""
from __future__ import annotations

import ctypes as C
import warnings
from typing import Dict, Optional, Tuple

import torch

from . import _lib

MAX_PEERS = 16
TIMEOUT_STALL = -777


class PeerStruct(C.Structure):
    """``nmb_peer_exchange_t``"""
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("n", C.c_int64), ("base", C.c_void_p * MAX_PEERS)]


class PeerExchange:
    """One exchange buffer per (group, element count, dtype); reused across loops (the epoch counter lives in the
    buffer header and advances in lock step on every rank)."""

    _cache: Dict[Tuple, "PeerExchange"] = {}

    def __init__(self, group, n: int, dtype, device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if world > MAX_PEERS:
            raise _lib.NmbError(f"peer exchange supports at most {MAX_PEERS} ranks, got {world}")
        lib = _lib.lib()
        lib.nmb_peer_exchange_bytes.restype = C.c_int64
        nbytes = int(lib.nmb_peer_exchange_bytes(_lib.dtype_code(dtype), C.c_int64(n)))
        try:
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                symm.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass                                    # newer torch enables groups on rendezvous
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=device)
        self.handle = symm.rendezvous(self.buf, group)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group)                         # every header is zero before anyone publishes
        ptrs = list(self.handle.buffer_ptrs)
        if len(ptrs) != world or any(int(p) == 0 for p in ptrs):
            raise _lib.NmbError(f"symmetric-memory rendezvous returned {len(ptrs)} peer pointers for {world} ranks")
        self.struct = PeerStruct(rank=rank, world=world, n=n)
        for r, p in enumerate(ptrs):
            self.struct.base[r] = int(p)
        self.rank, self.world, self.n = rank, world, n

    @classmethod
    def get(cls, group, n: int, dtype, device) -> "PeerExchange":
        key = (id(group), int(n), dtype, torch.device(device).index)
        px = cls._cache.get(key)
        if px is None:
            px = cls(group, n, dtype, device)
            cls._cache[key] = px
        return px

    @classmethod
    def try_get(cls, group, n: int, dtype, device) -> Tuple[Optional["PeerExchange"], Optional[str]]:
        """(exchange, None) or (None, reason).  The outcome is agreed over the group so that either every rank
        uses peer memory or every rank uses NCCL."""
        import torch.distributed as dist
        px, why = None, None
        try:
            px = cls.get(group, n, dtype, device)
        except Exception as e:                      # no P2P mapping on this box / torch without symmetric memory
            why = f"{type(e).__name__}: {e}"
        ok = torch.tensor([1 if px is not None else 0], device=device, dtype=torch.int32)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok) == 0:
            return None, why or "a peer rank could not map the exchange buffer"
        return px, None

    def evict(self) -> None:
        """Forget this buffer (after a timeout the epoch counters of the ranks disagree)."""
        for k, v in list(self._cache.items()):
            if v is self:
                del self._cache[k]

    def byref(self):
        return C.byref(self.struct)
