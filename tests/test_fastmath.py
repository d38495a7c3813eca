"""Host-side accuracy check of the branch-free FP64 transcendentals (csrc/fastmath.cuh) against long-double libm."""
import os
import re
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not on PATH")
def test_fastmath_matches_libm(tmp_path):
    exe = str(tmp_path / "fastmath_check")
    subprocess.check_call(["nvcc", "-O2", "-Wno-deprecated-gpu-targets", "-o", exe,
                           os.path.join(HERE, "host", "fastmath_check.cu")])
    out = subprocess.check_output([exe], text=True)
    taylor, table = out.splitlines()[0], out.splitlines()[1]
    vals = dict(re.findall(r"(exp rel|expm1 rel|sin abs|cos abs|sigmoid abs|tanh rel) ([0-9.e+-]+)", taylor))
    assert len(vals) == 6, out
    for k, v in vals.items():
        assert float(v) < 1e-15, (k, v)
    # table-driven variants: expm1 / tanh / elu carry one rounding of a table entry (absolute 1.1e-16), so their error
    # is measured relative to max(|result|, 2^-7) (and relative to the result itself where the entry is exactly 1)
    assert table.startswith("tab:")
    tvals = dict(re.findall(r"(exp rel|expm1 rel|sin abs|cos abs|sigmoid abs|tanh rel|elu rel) ([0-9.e+-]+)", table))
    assert len(tvals) == 7, out
    for k, v in tvals.items():
        assert float(v) < (1e-14 if k in ("expm1 rel", "tanh rel", "elu rel") else 1e-15), (k, v)
    for tag in ("edge: ", "tabedge: "):
        line = [l for l in out.splitlines() if l.startswith(tag)][0]
        assert "expm1(-100)=-1 " in line and "sigmoid(100)=1 " in line and "tanh(-50)=-1 " in line


@pytest.mark.gpu
def test_fastmath_device_tables_match_libm():
    """The table-driven kernels as the rollout kernels run them (shared-memory tables, rcp.approx + Newton) against
    float64 libm / mpmath-free references computed in extended precision on the host."""
    import ctypes as C

    import numpy as np
    import torch

    from nonmyopia_b200 import _lib

    lib = _lib.lib()
    g = torch.Generator().manual_seed(5)

    def probe(which, x):
        xd = x.cuda()
        y0, y1 = torch.empty_like(xd), torch.empty_like(xd)
        _lib.check(lib.nmb_fastmath_probe(which, _lib.ptr(xd), _lib.ptr(y0), _lib.ptr(y1), C.c_int64(xd.numel()),
                                          _lib.stream_ptr()), "nmb_fastmath_probe")
        return y0.cpu().numpy().astype(np.longdouble), y1.cpu().numpy().astype(np.longdouble)

    n = 1 << 20
    u = torch.rand(n, generator=g, dtype=torch.float64)
    x = torch.cat([(u - 0.5) * 120, (u - 0.5) * 2e-3, (u - 0.5) * 4])
    xl = x.numpy().astype(np.longdouble)
    e_tab, e_tay = probe(0, x)
    ref = np.exp(xl)
    assert np.max(np.abs((e_tab - ref) / ref)) < 1e-15 and np.max(np.abs((e_tay - ref) / ref)) < 1e-15
    m_tab, m_tay = probe(1, x)
    ref = np.expm1(xl)
    assert np.max(np.abs(m_tab - ref) / np.maximum(np.abs(ref), 2.0**-7)) < 1e-14
    # table entry 1 (k = 0): the degree-5 polynomial alone; its first dropped term r^6 / 720 is r^5 / 720 relative to
    # expm1(r) ~ r, i.e. 6.4e-15 at the edge |r| = ln2 / 128 (absolute error < 4e-17)
    small = np.abs(xl) < 0.005
    assert np.max(np.abs((m_tab[small] - ref[small]) / np.where(ref[small] == 0, 1, ref[small]))) < 1e-14
    xs = torch.cat([(u - 0.5) * 400, (u - 0.5) * 1e-3])
    s, c = probe(2, xs)
    xsl = xs.numpy().astype(np.longdouble) * np.longdouble("3.14159265358979323846264338327950288")
    assert np.max(np.abs(s - np.sin(xsl))) < 2e-15 and np.max(np.abs(c - np.cos(xsl))) < 2e-15     # |x| pi rounding
    xt = torch.cat([(u - 0.5) * 50, (u - 0.5) * 1e-2])
    xtl = xt.numpy().astype(np.longdouble)
    g_tab, g_tay = probe(3, xt)
    ref = 1 / (1 + np.exp(-xtl))
    assert np.max(np.abs(g_tab - ref)) < 1e-15 and np.max(np.abs(g_tay - ref)) < 1e-15
    t_tab, t_tay = probe(4, xt)
    ref = np.tanh(xtl)
    assert np.max(np.abs(t_tab - ref) / np.maximum(np.abs(ref), 2.0**-7)) < 1e-14
    xr = torch.exp((u - 0.5) * 40)
    r2, r3 = probe(5, xr)
    ref = 1 / xr.numpy().astype(np.longdouble)
    print("rcp 2 Newton steps rel err", float(np.max(np.abs((r2 - ref) / ref))), "3 steps",
          float(np.max(np.abs((r3 - ref) / ref))))
    assert np.max(np.abs((r3 - ref) / ref)) < 3e-16
    l_tab, l_tay = probe(6, x)
    ref = np.where(xl > 0, xl, np.expm1(xl))
    assert np.max(np.abs(l_tab - ref) / np.maximum(np.abs(ref), 2.0**-7)) < 1e-14
