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
    vals = dict(re.findall(r"(exp rel|expm1 rel|sin abs|cos abs|sigmoid abs|tanh rel) ([0-9.e+-]+)", out))
    assert len(vals) == 6, out
    for k, v in vals.items():
        assert float(v) < 1e-15, (k, v)
    assert "expm1(-100)=-1 " in out and "sigmoid(100)=1 " in out and "tanh(-50)=-1 " in out
