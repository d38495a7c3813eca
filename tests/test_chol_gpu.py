"""GPU parity of nmb_cholesky_root (csrc/chol.cu) - the Cholesky root and inverse root behind the GP caches -
against LAPACK on the host (torch.linalg on CPU tensors) and the oracle's psd_safe_cholesky."""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
F64 = torch.float64


def _spd(batch, n, seed, dtype=F64):
    g = torch.Generator().manual_seed(seed)
    B = torch.randn(batch, n, n + 3, generator=g, dtype=F64)
    return (B @ B.transpose(-1, -2) / (n + 3) + 0.05 * torch.eye(n, dtype=F64)).to(dtype)


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 64, 65, 100, 257, 500, 800])
def test_cholesky_root_fp64_matches_lapack(n):
    from nonmyopia_b200.gp import cholesky_root
    batch = 3 if n <= 257 else 1
    A = _spd(batch, n, n)
    L, Linv, info = cholesky_root(A.cuda())
    assert int(info.abs().sum()) == 0
    ref = torch.linalg.cholesky(A)
    # bit-level agreement is not defined across summation orders: 1e-10 relative to the largest entry
    assert float((L.cpu() - ref).abs().max()) / float(ref.abs().max()) < 1e-10
    assert float(L.cpu().triu(1).abs().max()) == 0.0 if n > 1 else True
    eye = torch.eye(n, dtype=F64)
    assert float((Linv.cpu() @ ref - eye).abs().max()) < 1e-9
    assert float(Linv.cpu().triu(1).abs().max()) == 0.0 if n > 1 else True


def test_cholesky_root_reports_first_bad_minor():
    from nonmyopia_b200.gp import cholesky_root
    A = _spd(2, 70, 0)
    A[1, 40, 40] = -1.0                       # leading minor of order 41 is the first non-PD one
    _, _, info = cholesky_root(A.cuda())
    _, ref_info = torch.linalg.cholesky_ex(A)
    assert info.cpu().tolist() == ref_info.tolist() == [0, 41]
    A[0, 3, 3] = float("nan")
    _, _, info = cholesky_root(A.cuda())
    assert int(info[0]) == 4


def test_cholesky_root_fp32_variant():
    from nonmyopia_b200.gp import cholesky_root
    A = _spd(2, 200, 5)
    L, Linv, info = cholesky_root(A.float().cuda())
    assert L.dtype == torch.float32 and int(info.abs().sum()) == 0
    ref = torch.linalg.cholesky(A)
    assert float((L.cpu().double() - ref).abs().max()) / float(ref.abs().max()) < 1e-3


def test_gp_caches_match_oracle_and_jitter_rule():
    """SurrogateGP (kernel matrix + nmb_cholesky_root + caches) vs the oracle's DenseGP on a golden fixture; a
    singular matrix goes through GPyTorch's jitter ladder like the oracle's psd_safe_cholesky."""
    from nonmyopia_b200.gp import _psd_safe_cholesky
    from oracle import gp_math
    from tests.test_dropin_gpu import product_gp
    fx = H.load("ts_am_cosine8_L3")
    ref, gp = H.dense_gp(fx["gp"]), product_gp(fx["gp"])
    assert H.rel_err(gp.L, ref.L, atol=1e-3) < 1e-8
    assert H.rel_err(gp.alpha, ref.alpha, atol=1e-3) < 1e-8
    assert H.rel_err(gp.Linv.T, ref.R, atol=1e-3) < 1e-8
    v = torch.randn(40, 3, dtype=F64, generator=torch.Generator().manual_seed(1))
    S = v @ v.T                                # rank 3: needs jitter
    L, _ = _psd_safe_cholesky(S.cuda())
    Lr = gp_math.psd_safe_cholesky(S)
    assert float((L.cpu() @ L.cpu().T - Lr @ Lr.T).abs().max()) < 1e-6


@pytest.mark.parametrize("n", [128, 161, 333, 500, 800])
def test_multi_cta_path_is_bit_identical_to_the_single_cta_kernel(n, monkeypatch):
    """One large matrix runs as diag / panel / update launches over many CTAs (chol.cu): every element goes through the
    same FMA sequence as the single-CTA kernel, so L and L^-1 agree bit for bit; a non-PD minor is reported alike."""
    import time
    from nonmyopia_b200.gp import cholesky_root
    A = _spd(1, n, 100 + n).cuda()
    monkeypatch.setenv("NMB_CHOL_SINGLE_CTA", "1")
    L1, X1, i1 = cholesky_root(A)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); cholesky_root(A); torch.cuda.synchronize(); t_single = time.perf_counter() - t0
    monkeypatch.delenv("NMB_CHOL_SINGLE_CTA")
    Lm, Xm, im = cholesky_root(A)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); cholesky_root(A); torch.cuda.synchronize(); t_multi = time.perf_counter() - t0
    print(f"n={n}: single CTA {1e3 * t_single:.2f} ms, multi CTA {1e3 * t_multi:.2f} ms")
    assert int(i1) == int(im) == 0
    assert torch.equal(L1, Lm) and torch.equal(X1, Xm)
    B = A.clone()
    B[0, n - 7, n - 7] = -3.0
    _, _, ib = cholesky_root(B)
    assert int(ib) == n - 6
