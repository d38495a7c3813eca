"""CPU: the C-ABI library builds, loads and exports every symbol include/*.h declares."""
import ctypes
import os
import re

from tests.conftest import ROOT


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "nonmyopia_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nmb_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_header_symbols():
    from nonmyopia_b200 import build
    path = build.build()
    lib = ctypes.CDLL(path)
    syms = _declared_symbols()
    assert len(syms) >= 5
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    lib.nmb_version.restype = ctypes.c_int
    assert lib.nmb_version() >= 100


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "nonmyopia_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f


def test_reference_module_names_are_mirrored():
    """The names the reference's main.py / actor.py / compute_metrics.py import from acqfs, actor,
    amortized_network and utils (main.py:19-33, actor.py:13-22, compute_metrics.py:23) exist here."""
    import importlib
    want = {
        "acqfs": ["qBOAcqf", "qMultiStepHEntropySearch", "qLossFunctionTopK", "qCostFunction",
                  "qCostFunctionEditDistance"],
        "actor": ["Actor"],
        "amortized_network": ["AmortizedNetwork", "Project2Range"],
        "utils": ["eval_func", "set_seed", "generate_random_points_batch", "generate_random_rotation_matrix",
                  "rotate_points"],
    }
    for mod, names in want.items():
        m = importlib.import_module("nonmyopia_b200." + mod)
        for n in names:
            assert hasattr(m, n), f"{mod}.{n}"


def test_eval_func_fails_loudly_without_cuda():
    """No CPU fallback: on a box without a GPU the product call raises instead of computing on the host."""
    import types

    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a CPU-only box")
    from nonmyopia_b200 import acqfs
    from nonmyopia_b200.utils import eval_func
    parms = types.SimpleNamespace(
        algo="HES-TS-AM-20", algo_lookahead_steps=3, x_dim=2, n_actions=1, n_restarts=4, n_initial_points=5,
        acq_opt_lr=1e-2, torch_dtype=torch.float64, loss_function_class=acqfs.qLossFunctionTopK,
        loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5), cost_function_class=acqfs.qCostFunction,
        cost_func_hypers=dict(radius=0.1))
    X, Y = torch.rand(5, 2, dtype=torch.float64), torch.randn(5, 1, dtype=torch.float64)
    model = types.SimpleNamespace(train_inputs=(X,), train_targets=Y.reshape(-1))
    with pytest.raises(Exception):
        eval_func(lambda a: a.sum(), model, parms, dict(x=X, y=Y, cost=torch.zeros(5)), 4)


def test_tile_stride_constant_is_consistent():
    """NMB_TILE_STRIDE (header) == kSW (kernels) == the host packer's default, and it is 4 mod 16 doubles - the
    bank-conflict-free stride of the DMMA B-fragment loads (a half-warp reads k = 0..3 x n = 0..3)."""
    import inspect
    hdr = open(os.path.join(ROOT, "include", "nonmyopia_b200.h")).read()
    cuh = open(os.path.join(ROOT, "nonmyopia_b200", "csrc", "policy_mma.cuh")).read()
    h = int(re.search(r"#define\s+NMB_TILE_STRIDE\s+(\d+)", hdr).group(1))
    k = int(re.search(r"constexpr int kSW = (\d+);", cuh).group(1))
    from nonmyopia_b200 import mc_tree
    py = inspect.signature(mc_tree._col_tiles).parameters["stride"].default
    assert h == k == py
    assert h % 16 in (4, 12) and h >= 64
