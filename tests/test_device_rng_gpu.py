"""GPU: statistical pins of the in-kernel ("device") random mode of the HES-TS-AM rollout.

The reference draws, per noisy policy call, ``theta = rand(N) * 2 pi - pi`` in float32 and a coordinate pair
``randperm(d)[:2]`` per sample (amortized_network.py:136,156), and one N(0,1) draw per cost call (acqfs.py:624).  The
device mode cannot be bit-compared with the host generator, so it is pinned by distribution: uniform angle (KS),
uniform ORDERED coordinate pairs with di != dj (chi-square), normal moments + KS of the cost noise, and
independence / distinctness of the streams across (chain, step, call, rollout) and shard-invariance (the stream is
keyed by the GLOBAL chain index: a rank's chain i draws what the unsharded run draws for chain offset + i)."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _draws(seed, step, stream, n_calls, N, d, offset=0):
    from nonmyopia_b200 import _lib
    lib = _lib.lib()
    cs = torch.empty(n_calls, N, 2, dtype=torch.float64, device="cuda")
    dims = torch.empty(n_calls, N, 2, dtype=torch.int32, device="cuda")
    nz = torch.empty(n_calls, N, dtype=torch.float64, device="cuda")
    _lib.check(lib.nmb_ts_device_draws(C.c_uint64(seed), C.c_int64(step), C.c_uint(stream), n_calls, C.c_int64(N),
                                       C.c_int64(offset), d,
                                       C.cast(_lib.ptr(cs), C.POINTER(C.c_double)),
                                       C.cast(_lib.ptr(dims), C.POINTER(C.c_int32)),
                                       C.cast(_lib.ptr(nz), C.POINTER(C.c_double)), _lib.stream_ptr()),
               "nmb_ts_device_draws")
    torch.cuda.synchronize()
    return cs.cpu().numpy(), dims.cpu().numpy(), nz.cpu().numpy()


def _ks_uniform(u):
    """Kolmogorov-Smirnov statistic of samples against U(0, 1), scaled by sqrt(n)."""
    u = np.sort(u)
    n = len(u)
    i = np.arange(1, n + 1)
    return math.sqrt(n) * max(np.max(i / n - u), np.max(u - (i - 1) / n))


def test_rotation_angle_is_uniform_and_unit_norm():
    cs, _, _ = _draws(1234, 7, 0, 4, 65536, 2)
    c, s = cs[..., 0].ravel(), cs[..., 1].ravel()
    # float32-rounded cos / sin as the reference builds them (amortized_network.py:144-153)
    assert np.array_equal(c, c.astype(np.float32).astype(np.float64))
    assert np.max(np.abs(c * c + s * s - 1)) < 1e-6
    th = np.arctan2(s, c)
    assert _ks_uniform((th + np.pi) / (2 * np.pi)) < 1.63          # 1 % critical value of the KS statistic
    assert abs(th.mean()) < 4 * (np.pi / math.sqrt(3)) / math.sqrt(len(th))


@pytest.mark.parametrize("d", [2, 3, 6, 8])
def test_coordinate_pairs_are_uniform_ordered_pairs(d):
    _, dims, _ = _draws(99, 3, 1, 3, 40000, d)
    di, dj = dims[..., 0].ravel(), dims[..., 1].ravel()
    assert di.min() >= 0 and di.max() < d and dj.min() >= 0 and dj.max() < d
    assert np.all(di != dj)                                       # randperm(d)[:2]: two distinct coordinates
    counts = np.zeros((d, d))
    np.add.at(counts, (di, dj), 1)
    k = d * (d - 1)
    exp = len(di) / k
    chi2 = ((counts[~np.eye(d, dtype=bool)] - exp) ** 2 / exp).sum()
    # chi-square with k - 1 degrees of freedom: mean k - 1, sd sqrt(2 (k - 1)); 5 sd bound
    assert chi2 < (k - 1) + 5 * math.sqrt(2 * (k - 1)) + 5, chi2


def test_cost_noise_is_standard_normal():
    _, _, nz = _draws(5, 11, 0, 4, 65536, 2)
    z = nz.ravel()
    n = len(z)
    assert abs(z.mean()) < 4 / math.sqrt(n)
    assert abs(z.var() - 1) < 4 * math.sqrt(2 / n)
    assert abs((z ** 3).mean()) < 4 * math.sqrt(15 / n)
    assert abs((z ** 4).mean() - 3) < 4 * math.sqrt(96 / n)
    cdf = 0.5 * (1 + np.vectorize(math.erf)(z / math.sqrt(2)))
    assert _ks_uniform(cdf) < 1.63
    assert np.isfinite(z).all() and np.abs(z).max() < 7


def test_streams_are_distinct_and_uncorrelated():
    N = 32768
    base = _draws(42, 5, 0, 2, N, 2)
    variants = {"step": _draws(42, 6, 0, 2, N, 2), "seed": _draws(43, 5, 0, 2, N, 2),
                "rollout": _draws(42, 5, 1, 2, N, 2)}
    th0 = np.arctan2(base[0][..., 1], base[0][..., 0])
    for what, v in variants.items():
        th = np.arctan2(v[0][..., 1], v[0][..., 0])
        assert np.mean(th == th0) < 1e-3, what
        assert abs(np.corrcoef(th.ravel(), th0.ravel())[0, 1]) < 5 / math.sqrt(th.size), what
        assert abs(np.corrcoef(v[2].ravel(), base[2].ravel())[0, 1]) < 5 / math.sqrt(th.size), what
    # calls within one iteration, and neighbouring chains
    assert abs(np.corrcoef(th0[0], th0[1])[0, 1]) < 5 / math.sqrt(N)
    assert abs(np.corrcoef(th0[0][:-1], th0[0][1:])[0, 1]) < 5 / math.sqrt(N)
    assert abs(np.corrcoef(base[2][0], base[2][1])[0, 1]) < 5 / math.sqrt(N)
    # the noise and the angle of the same (call, chain) come from different counters
    assert abs(np.corrcoef(base[2].ravel(), th0.ravel())[0, 1]) < 5 / math.sqrt(th0.size)
    assert len(np.unique(th0[0])) > 0.99 * N * (1 - N / 2 ** 25)      # 24-bit angles: birthday collisions only


def test_stream_is_keyed_by_the_global_chain_index():
    """Shard invariance: rank g's chain i draws what the unsharded run draws for chain lo + i."""
    full = _draws(7, 2, 0, 3, 4096, 6)
    lo = 1536
    part = _draws(7, 2, 0, 3, 1024, 6, offset=lo)
    for a, b in zip(full, part):
        assert np.array_equal(a[:, lo:lo + 1024], b)
