"""GPU parity of the device-resident AdamW loop against the golden vectors recorded from the
reference's Actor.query (oracle/make_golden.py: case_query) - same draws, same iterations."""
import pytest
import torch

from tests import helpers as H
from tests.test_path_gpu import _product_path
from tests.test_ts_rollout_gpu import pack_draws

pytestmark = pytest.mark.gpu


def _make_loop(fx, **kw):
    from nonmyopia_b200 import ts_engine as te
    from nonmyopia_b200.ts_loop import TSAMLoop
    d, Hh, L, N = fx["d"], fx["H"], fx["L"], fx["n_r"]
    eng = te.TSRollout(_product_path(fx), d, Hh, L, N, fx["cost"])
    theta = te.flatten_params(fx["policy"])
    return TSAMLoop(eng, theta, fx["prev_X"], fx["prev_y"], fx["prev_h"], budget=fx["budget"], lr=fx["lr"],
                    past_initial=False, **kw)


def test_query_loop_matches_reference_actor():
    from nonmyopia_b200 import ts_engine as te
    fx = H.load("query_ts_am_ackley_L1")
    L, N = fx["L"], fx["n_r"]
    loop = _make_loop(fx, max_iter=fx["n_iter"], record_history=True)
    draws = []
    for dr in fx["draws"]:
        mq, md, mn = pack_draws(dr["main_rot"], dr["main_noise"], L, N)
        oq, od, on = pack_draws(dr["orig_rot"], dr["orig_noise"], L, N)
        draws.append((mq, md, mn, oq, od, on))
    st = loop.run(fx["n_iter"], draws_per_iter=draws)
    assert st.step == fx["n_iter"]
    res = loop.result()
    out = fx["out"]
    # "the selected next query must be identical after the final argmax"
    assert int(res["idx"]) == out["chosen_idx"]
    assert H.rel_err(res["next_X"], out["next_X"], atol=1e-3) < 1e-8
    assert H.rel_err(res["best_loss"], out["all_losses"], atol=1e-2) < 1e-8
    assert H.rel_err(res["best_cost"], out["all_costs"], atol=1e-3) < 1e-8
    assert H.rel_err(res["hidden_state"], out["hidden_state"], atol=1e-3) < 1e-8
    # AdamW trajectory of the 13,924 network parameters after n_iter steps
    final = te.unflatten_params(res["theta"].cpu(), fx["d"], fx["H"])
    for k, ref in out["final_policy"].items():
        assert H.rel_err(final[k], ref, atol=1e-3) < 1e-8, k
    # margin between best and runner-up restart (so a 1e-8 wobble cannot flip the argmin)
    tot = (res["best_loss"] + res["best_cost"] + res["best_KL"]).cpu()
    srt = torch.sort(tot).values
    assert float(srt[1] - srt[0]) > 1e-6


def test_graph_replay_equals_eager_and_early_stop_freezes_state():
    fx = H.load("query_ts_am_ackley_L1")
    a = _make_loop(fx, max_iter=40)
    b = _make_loop(fx, max_iter=40)
    sa = a.run(24, use_graph=True, poll_every=8)
    sb = b.run(24, use_graph=False, poll_every=8)
    assert sa.step == sb.step == 24
    # in-kernel Philox draws depend only on (seed, step, call, chain): graph and eager agree bitwise
    assert torch.equal(a.theta, b.theta)
    assert torch.equal(a.best_loss, b.best_loss)
    # force the stop flag: further replays must not change anything
    a.state[8:12] = torch.tensor([1, 0, 0, 0], dtype=torch.uint8, device="cuda")
    th = a.theta.clone()
    a.graph.replay(); a.graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(th, a.theta) and a.read_state().step == 24


def test_step_host_launch_ahead_equals_plain_replays():
    """``step_host`` with ``next_inputs`` (prefetch on a side stream + the following iteration launched behind this
    step's result copies) walks exactly the trajectory of plain graph replays, returns every step's own loss / cost,
    and refuses inputs other than the promised ones."""
    from nonmyopia_b200 import _lib
    fx = H.load("query_ts_am_ackley_L1")
    n = 12
    ref = _make_loop(fx, max_iter=40, record_history=True)
    ref.run(n, use_graph=True, poll_every=4)
    hx, hy, hh = (t.to(torch.float64).contiguous().pin_memory() for t in (fx["prev_X"], fx["prev_y"].reshape(-1), fx["prev_h"]))
    N = hx.shape[0]
    for promise in (False, True):
        loop = _make_loop(fx, max_iter=40)
        loop.capture()
        out = torch.empty(2, N, dtype=torch.float64).pin_memory()
        losses = []
        for k in range(n):
            nxt = (hx, hy, hh) if (promise and k + 1 < n) else None
            loop.step_host(hx, hy, hh, out, next_inputs=nxt)
            losses.append(out[0].clone())
        torch.cuda.synchronize()
        assert loop.read_state().step == n
        assert torch.equal(loop.theta, ref.theta), promise
        assert torch.equal(loop.best_loss, ref.best_loss), promise
        got = torch.stack(losses)
        assert torch.equal(got, ref.hist_loss[:n].cpu()), promise       # each call returned ITS step's losses
    # a promised step must be called with the promised tensors
    loop = _make_loop(fx, max_iter=40)
    loop.capture()
    out = torch.empty(2, N, dtype=torch.float64).pin_memory()
    loop.step_host(hx, hy, hh, out, next_inputs=(hx, hy, hh))
    other = hx.clone().pin_memory()
    with pytest.raises(_lib.NmbError):
        loop.step_host(other, hy, hh, out)


def test_fused_exchange_kernel_with_one_rank_equals_the_unsharded_iteration():
    """``nmb_peer_exchange_adamw_step`` (split-K reduce, LL push / receive, rank-ordered sum + AdamW, decision + advance
    in one kernel) run with a world of ONE rank over a plain device buffer must reproduce the unsharded iteration
    (objective-side decision, ``nmb_policy_adamw_step``) bit for bit: weights, Adam moments, snapshots, loop state.
    (The 2-rank tests of tests/test_multigpu_gpu.py need two GPUs; this one runs wherever the suite runs.)"""
    import ctypes as C
    from nonmyopia_b200 import _lib
    from nonmyopia_b200.peer import PeerStruct
    fx = H.load("query_ts_am_ackley_L1")

    class _SoloExchange:
        def __init__(self, n, dtype):
            lib = _lib.lib()
            lib.nmb_peer_exchange_bytes.restype = C.c_int64
            nbytes = int(lib.nmb_peer_exchange_bytes(_lib.dtype_code(dtype), C.c_int64(n)))
            self.buf = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
            self.struct = PeerStruct(rank=0, world=1, n=n)
            self.struct.base[0] = self.buf.data_ptr()

        def byref(self):
            return C.byref(self.struct)

    ref = _make_loop(fx, max_iter=40)
    fused = _make_loop(fx, max_iter=40)
    fused.group = "one rank"                     # objective leaves the decision to the exchange
    fused.px = _SoloExchange(fused.theta.numel(), fused.theta.dtype)
    fused.fused_exchange = True
    n = 14
    for _ in range(n):
        ref._iteration()
        fused._iteration()
    torch.cuda.synchronize()
    sa, sb = ref.read_state(), fused.read_state()
    assert sa.step == sb.step == n and sa.n_snapshots == sb.n_snapshots and sa.stall == sb.stall
    assert torch.equal(ref.theta, fused.theta) and torch.equal(ref.theta_t, fused.theta_t)
    assert torch.equal(ref.m, fused.m) and torch.equal(ref.v, fused.v)
    assert torch.equal(ref.best_loss, fused.best_loss) and torch.equal(ref.best_X, fused.best_X)
    # the exchange epoch advanced with the loop
    epoch = fused.px.buf[16 * 8:16 * 8 + 8].view(torch.int64)
    assert int(epoch) == n
