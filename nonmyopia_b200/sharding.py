"""Restart sharding over ranks: one process per GPU, rank ``g`` owns a block of the
``n_restarts`` axis of every level tensor.  The data path has no collective; global semantics of the
reference's loop (actor.py:558-656) are kept by

* C3: one all-reduce(MIN) of two doubles per Adam iteration - ``min_r total`` and, for the snapshot,
  ``min_r (loss + cost + KL)`` - so every rank takes the same snapshot / early-stop decision
  (actor.py:622-639, 655);
* the global mean of actor.py:641 (``loss.mean()`` over ALL restarts): the local backward is seeded with
  ``1 / n_restarts_global``;
* C1: a final all-gather of ``best_loss + best_cost + best_KL`` -> arg-min -> the owner broadcasts the
  selected restart's ``next_X`` / ``actions`` / ``hidden_state`` (actor.py:662, 677-733).

Everything here is backend-agnostic host logic over ``torch.distributed`` (NCCL on the GPUs; the
world_size-2 ``gloo`` tests in tests/test_sharding_cpu.py drive it with the CPU oracle as the objective).
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import warnings

import torch
from torch import Tensor

# the graph-replayed loop runs its backward on the capture stream while the leaves were created on the caller's
# stream; the streams are ordered explicitly (wait_stream), so autograd's generic hint does not apply
warnings.filterwarnings("ignore", message="The AccumulateGrad node's stream does not match")


def _dist():
    import torch.distributed as dist
    return dist


def world_of(group) -> Tuple[int, int]:
    """(rank, world_size) of ``group``; (0, 1) when ``group`` is None (single GPU)."""
    if group is None:
        return 0, 1
    dist = _dist()
    return dist.get_rank(group), dist.get_world_size(group)


def restart_range(n_restarts: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of restarts owned by ``rank``; ragged counts give the first
    ``n_restarts % world`` ranks one extra restart."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    if n_restarts < world:
        raise ValueError(f"{n_restarts} restarts cannot be sharded over {world} ranks")
    base, extra = divmod(n_restarts, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_level(t: Tensor, n_restarts: int, lo: int, hi: int) -> Tensor:
    """Rows of restarts ``[lo, hi)`` of a flattened level tensor ``[cnt_s, ...]``.

    The reference flattens ``[*nf_rev[:s], n_r, ...]`` (actor.py:240-247, acqfs.py:333-345): the restart
    index is the FASTEST batch index, so a shard is a strided slice, not a contiguous block."""
    if t.shape[0] % n_restarts:
        raise ValueError(f"leading dim {t.shape[0]} is not a multiple of n_restarts={n_restarts}")
    v = t.reshape(t.shape[0] // n_restarts, n_restarts, *t.shape[1:])
    return v[:, lo:hi].reshape(-1, *t.shape[1:]).contiguous()


def unshard_level(parts: Sequence[Tensor], counts: Sequence[int]) -> Tensor:
    """Inverse of ``shard_level`` given every rank's shard (in rank order) and restart count."""
    views = [p.reshape(p.shape[0] // c, c, *p.shape[1:]) for p, c in zip(parts, counts)]
    full = torch.cat(views, dim=1)
    return full.reshape(-1, *full.shape[2:])


def allreduce_min(values: Tensor, group) -> Tensor:
    """In-place MIN over ranks (no-op for ``group is None``); ``values`` holds a few doubles."""
    if group is not None:
        dist = _dist()
        dist.all_reduce(values, op=dist.ReduceOp.MIN, group=group)
    return values


def allreduce_sum(values: Tensor, group) -> Tensor:
    if group is not None:
        dist = _dist()
        dist.all_reduce(values, op=dist.ReduceOp.SUM, group=group)
    return values


def assert_replicated(t: Tensor, group, what: str) -> None:
    """Raise unless ``t`` is bit-identical on every rank of ``group``.  The sharded loops keep replicated state
    (policy weights, hidden noise, Thompson path, base samples) consistent by construction - same host seed on
    every rank - and the per-iteration exchange applies identical updates, so a rank seeded differently would
    otherwise drift silently."""
    if group is None:
        return
    dist = _dist()
    world = dist.get_world_size(group)
    mine = t.detach().reshape(-1).contiguous()
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    bad = [r for r, q in enumerate(parts) if not torch.equal(q.view(torch.uint8), parts[0].view(torch.uint8))]
    if bad:
        raise RuntimeError(f"replicated state differs across ranks ({what}): ranks {bad} disagree with rank 0 - seed "
                           "every rank identically (torch.manual_seed) before Actor.reset_parameters / query")


def gather_restarts(local: Tensor, n_restarts: int, group) -> Tensor:
    """All-gather a per-restart vector ``[n_local, ...]`` into ``[n_restarts, ...]`` (rank order = restart
    order).  Ragged shards are padded to the largest shard for the collective."""
    rank, world = world_of(group)
    if world == 1:
        return local
    dist = _dist()
    counts = [restart_range(n_restarts, r, world) for r in range(world)]
    width = max(hi - lo for lo, hi in counts)
    pad = local.new_zeros((width, *local.shape[1:]))
    pad[:local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:hi - lo] for o, (lo, hi) in zip(out, counts)], dim=0)


def gather_level(local: Tensor, n_restarts: int, group) -> Tensor:
    """All ranks' shards of a flattened level tensor ``[cnt_s_local, ...]`` reassembled into the unsharded
    ``[cnt_s, ...]`` layout (restart index fastest)."""
    rank, world = world_of(group)
    if world == 1:
        return local
    lo, hi = restart_range(n_restarts, rank, world)
    n_loc = hi - lo
    v = local.reshape(local.shape[0] // n_loc, n_loc, *local.shape[1:]).transpose(0, 1).contiguous()
    full = gather_restarts(v, n_restarts, group)                  # [n_restarts, branches, ...]
    return full.transpose(0, 1).reshape(-1, *local.shape[1:]).contiguous()


def select_best(total_local: Tensor, n_restarts: int, group) -> Tuple[Tensor, int, int, int]:
    """C1: all-gather the per-restart totals, arg-min (first minimum, like ``torch.argmin`` on the
    unsharded vector).  Returns (all totals [n_restarts], global index, owner rank, index inside the owner)."""
    rank, world = world_of(group)
    totals = gather_restarts(total_local, n_restarts, group)
    idx = int(torch.argmin(totals))
    for r in range(world):
        lo, hi = restart_range(n_restarts, r, world)
        if lo <= idx < hi:
            return totals, idx, r, idx - lo
    raise AssertionError("unreachable")


def broadcast_from_owner(t: Optional[Tensor], shape, dtype, device, owner: int, group) -> Tensor:
    """The owner's tensor on every rank (``t`` is ignored on the other ranks)."""
    rank, world = world_of(group)
    if world == 1:
        return t
    dist = _dist()
    buf = t.contiguous().clone() if rank == owner else torch.empty(shape, dtype=dtype, device=device)
    dist.broadcast(buf, src=dist.get_global_rank(group, owner), group=group)
    return buf


def free_restart_loop(objective: Callable[[List[Tensor]], Dict], params: List[Tensor],
                      opt_step: Callable[[Sequence[Optional[Tensor]]], None], *, budget: float, n_iter: int,
                      past_initial: bool, n_global: Optional[int] = None, group=None, save_trajectory: bool = False,
                      log: Optional[Callable[[str], None]] = print, use_graph: bool = False,
                      pre_iteration: Optional[Callable[[], None]] = None, timing: Optional[dict] = None) -> Dict:
    """The reference's AdamW loop for free-parameter algorithms (actor.py:558-656; HES-TS, HES-MC, qEI ... qKG)
    over this rank's restarts.

    ``objective(local_maps)`` evaluates ``acqf.forward`` on ``[sigmoid(p) for p in params]`` and returns the
    reference's dict (``acqf_loss`` / ``acqf_cost`` per local restart, ``X``, ``actions``); ``opt_step(grads)``
    applies AdamW (a ``None`` gradient leaves its tensor untouched, actor.py:641-648).  With a process group
    the decision values are MIN-reduced and the backward is seeded with ``1 / n_global``.

    ``use_graph`` (CUDA only): after two eager iterations the device work of one iteration - sigmoid, objective,
    backward, AdamW, the two decision minima - is captured once and replayed as a CUDA graph; the host still reads
    the minima every iteration for the reference's snapshot / early-stop rule (one synchronisation per iteration
    instead of ~100 eager launches).  Requires an objective without per-call host randomness: host-stream draws
    are made resident by ``pre_iteration()`` (called before every iteration, eager or replayed) in a static device
    buffer the objective reads.  ``timing`` (a dict, optional) receives ``ms_per_iter``: CUDA-event time of the
    replayed iterations (everything after the capture), host reads included."""
    best = dict(loss=None, cost=None, KL=0)
    best_min = float("inf")
    snap, stall = {}, 0
    saved_trajectory, saved_loss, saved_cost = [], [], []
    n_done = 0
    keep = (lambda t: t.detach().clone()) if use_graph else (lambda t: t.detach())

    def iteration():
        local_maps = [torch.sigmoid(x) for x in params]
        rd = objective(local_maps)
        acqf_loss, acqf_cost = rd["acqf_loss"], rd["acqf_cost"]
        loss = acqf_loss + abs(budget - acqf_cost)
        cand = torch.stack([loss.detach().min(), (acqf_loss + acqf_cost).detach().min()])
        n_loc = acqf_loss.shape[0]
        n_glob = n_loc if n_global is None else int(n_global)
        seed = loss.mean() if n_glob == n_loc else loss.sum() / n_glob
        opt_step(torch.autograd.grad(seed, params, allow_unused=True))
        return local_maps, rd, cand

    graph, static, side, warm = None, None, None, 2
    if use_graph:
        side = torch.cuda.Stream(device=params[0].device)
    for ep in range(n_iter):
        if pre_iteration is not None:
            pre_iteration()
        if use_graph and ep >= warm and graph is not False:
            if graph is None:
                try:
                    torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        static = iteration()
                    graph = g
                except Exception as e:          # an op of this objective cannot be captured: stay eager
                    if log is not None:
                        log(f"free-parameter loop runs eagerly (graph capture failed: {type(e).__name__})")
                    torch.cuda.synchronize()
                    graph, keep = False, (lambda t: t.detach())
            if graph:
                if timing is not None and "t0" not in timing and ep > warm:
                    timing["t0"] = torch.cuda.Event(enable_timing=True)
                    timing["t0"].record()
                    timing["ep0"] = ep
                graph.replay()
                local_maps, rd, cand = static
            else:
                local_maps, rd, cand = iteration()
        elif use_graph:
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                local_maps, rd, cand = iteration()
            torch.cuda.current_stream().wait_stream(side)
        else:
            local_maps, rd, cand = iteration()
        acqf_loss, acqf_cost = rd["acqf_loss"], rd["acqf_cost"]
        if save_trajectory:
            # history stays on the device (one small copy per level); it is converted once after the loop - the
            # reference's per-iteration .cpu().tolist() (actor.py:561-564) would put a host sync and, at the
            # Hartmann tree's size, 1.28 M doubles of list building into every iteration
            saved_trajectory.append([x.detach().clone() for x in local_maps])
            saved_loss.append(acqf_loss.detach().clone())
            saved_cost.append(acqf_cost.detach().clone())
        cand = allreduce_min(cand.clone() if group is not None else cand, group).tolist()
        if ep == 0 or cand[0] < best_min:
            best = dict(loss=keep(acqf_loss), cost=keep(acqf_cost), KL=0)
            best_min = cand[1]
            snap = dict(X=[keep(x) for x in rd["X"]], actions=keep(rd["actions"]))
            stall = 0
        elif past_initial or ep >= 200:
            stall += 1
        n_done = ep + 1
        if log is not None and ep % 50 == 0:
            chosen = torch.argmin(acqf_loss + abs(budget - acqf_cost))
            log(f"Epoch {ep:05d}\tLoss {acqf_loss[chosen].item():.5f}\tCost: {acqf_cost[chosen].item():.5f}")
        if stall > 50:
            break
    if use_graph:
        torch.cuda.current_stream().wait_stream(side)
    if timing is not None and "t0" in timing:
        t1 = torch.cuda.Event(enable_timing=True)
        t1.record()
        torch.cuda.synchronize()
        timing["ms_per_iter"] = timing["t0"].elapsed_time(t1) / max(n_done - timing["ep0"], 1)
    if save_trajectory:
        saved_trajectory = [[x.cpu().tolist() for x in lv] for lv in saved_trajectory]
        saved_loss = [x.tolist() for x in saved_loss]
        saved_cost = [x.tolist() for x in saved_cost]
    return dict(best_loss=best["loss"], best_cost=best["cost"], best_KL=0, best_next_X=snap["X"],
                best_actions=snap["actions"], best_hidden_state=None, saved_trajectory=saved_trajectory,
                saved_loss=saved_loss, saved_cost=saved_cost, n_iter=n_done)
