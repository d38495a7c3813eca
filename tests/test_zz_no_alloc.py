"""GPU: the allocation-free entry point ``nmb_gp_mll_value_grad_static`` (workspace in a file-scope device array of
the library; no counterpart in the reference, added for callers that cannot call cudaMalloc) against the existing
``nmb_gp_mll_value_grad`` (caller workspace), through the C-ABI, on two calls in a row.

Collected last on purpose (``test_zz_``).  Passes alone in a fresh process and
after the other tests: the only shared state is the library's array, which every call fully overwrites before
reading."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
F64 = torch.float64
STATIC_MAX_N = 256


def _inputs(n, d, kernel, seed):
    """Training data, K + s2 I and its root / inverse root - everything allocated HERE, before the calls compared."""
    from nonmyopia_b200 import _lib
    from nonmyopia_b200.gp import _psd_safe_cholesky, kernel_matrix
    g = torch.Generator().manual_seed(seed)
    X = torch.rand(n, d, dtype=F64, generator=g).cuda()
    y = torch.randn(n, dtype=F64, generator=g).cuda()
    ls = (0.3 + torch.rand(d, dtype=F64, generator=g)).cuda()
    Khat = kernel_matrix(X, X, kernel, ls, 1.3, 1e-2)
    L, Linv = _psd_safe_cholesky(Khat)
    return dict(kid=_lib.KERNEL_IDS[kernel], n=n, d=d, X=X.contiguous(), y=y.contiguous(),
                inv_ls=(1.0 / ls).contiguous(), L=L.contiguous(), Linv=Linv.contiguous(), c=0.17)


def _dp(t):
    from nonmyopia_b200 import _lib
    return C.cast(_lib.ptr(t), C.POINTER(C.c_double))


def _existing(lib, a, out):
    from nonmyopia_b200 import _lib
    lib.nmb_gp_mll_workspace_elems.restype = C.c_int64
    work = torch.full((int(lib.nmb_gp_mll_workspace_elems(a["n"], a["d"])),), float("nan"), dtype=F64, device="cuda")
    _lib.check(lib.nmb_gp_mll_value_grad(a["kid"], a["d"], a["n"], _dp(a["X"]), _dp(a["inv_ls"]), _dp(a["y"]),
                                         C.c_double(a["c"]), _dp(a["L"]), C.c_int64(a["n"]), _dp(a["Linv"]),
                                         C.c_int64(a["n"]), _dp(work), _dp(out), _lib.stream_ptr()),
               "nmb_gp_mll_value_grad")


def _static(lib, a, out):
    return lib.nmb_gp_mll_value_grad_static(a["kid"], a["d"], a["n"], _dp(a["X"]), _dp(a["inv_ls"]), _dp(a["y"]),
                                            C.c_double(a["c"]), _dp(a["L"]), C.c_int64(a["n"]), _dp(a["Linv"]),
                                            C.c_int64(a["n"]), _dp(out), _stream())


def _stream():
    from nonmyopia_b200 import _lib
    return _lib.stream_ptr()


def test_static_workspace_entry_point_matches_existing_on_two_calls_in_a_row():
    from nonmyopia_b200 import _lib
    lib = _lib.lib()
    # a larger problem first, then a smaller one of another shape / kernel: the second call finds the first call's
    # contents in the shared array and must not be affected by them
    cases = [_inputs(STATIC_MAX_N, 6, "matern52", 1), _inputs(90, 2, "rbf", 2)]
    want = [torch.empty(5 + a["d"], dtype=F64, device="cuda") for a in cases]
    got = [torch.full((5 + a["d"],), float("nan"), dtype=F64, device="cuda") for a in cases]
    for a, w in zip(cases, want):
        _existing(lib, a, w)
    torch.cuda.synchronize()
    # the two calls under test, back to back on one stream, nothing allocated in between
    before = torch.cuda.memory_allocated()
    for a, o in zip(cases, got):
        _lib.check(_static(lib, a, o), "nmb_gp_mll_value_grad_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for a, w, o in zip(cases, want, got):
        # same kernels, same launch order, deterministic reductions: the tests of nmb_gp_mll_value_grad use 1e-9 on
        # the value and 1e-8 on the gradient against the oracle; the two entry points agree to the bit
        assert torch.isfinite(o).all()
        assert torch.equal(o, w), (a["n"], a["d"], (o - w).abs().max().item())
    # and once more in the opposite order (small -> large)
    for a, w in zip(reversed(cases), reversed(want)):
        o = torch.full_like(w, float("nan"))
        _lib.check(_static(lib, a, o), "nmb_gp_mll_value_grad_static")
        torch.cuda.synchronize()
        assert torch.equal(o, w)


def test_static_workspace_entry_point_rejects_sizes_beyond_its_limit():
    from nonmyopia_b200 import _lib
    lib = _lib.lib()
    a = _inputs(STATIC_MAX_N + 1, 2, "matern52", 3)
    out = torch.zeros(5 + a["d"], dtype=F64, device="cuda")
    rc = _static(lib, a, out)
    assert rc == -2                                                   # NMB_ELIMIT
    msg = lib.nmb_last_error().decode(errors="replace")
    assert "256" in msg and "nmb_gp_mll_value_grad_static" in msg
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0                              # nothing launched
    # the existing entry point takes the same problem with caller memory
    _existing(lib, a, out)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
