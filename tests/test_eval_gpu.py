"""GPU parity of ``nonmyopia_b200.utils.eval_func`` (reference utils.py:143-299) against the
golden vectors recorded from the reference's own ``utils.eval_func`` (oracle/make_golden.py ``case_eval``)."""
import types

import pytest
import torch

from tests import helpers as H
from tests.test_dropin_gpu import product_gp

pytestmark = pytest.mark.gpu
F64 = torch.float64


def _parms(fx, **over):
    from nonmyopia_b200 import acqfs
    p = dict(algo=fx["algo"], algo_lookahead_steps=fx["lookahead"], x_dim=fx["d"], y_dim=1, n_actions=1,
             n_restarts=fx["n_r"], n_initial_points=fx["n_initial"], acq_opt_lr=fx["lr"], torch_dtype=F64,
             num_categories=None, loss_function_class=acqfs.qLossFunctionTopK,
             loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5), cost_function_class=acqfs.qCostFunction,
             cost_func_hypers=dict(fx["cost"]))
    p.update(over)
    return types.SimpleNamespace(**p)


def _env(fx):
    from nonmyopia_b200.synthetic import objective
    return lambda a: objective(fx["env"], a.reshape(-1, fx["d"]).cpu())


@pytest.mark.parametrize("name", ["eval_hes_ackley", "eval_hes_hartmann_init", "eval_qei_ackley", "eval_qkg_ackley"])
def test_eval_func_matches_reference_golden(name):
    """All 1000 steps on the reference's recorded draws: the chosen action agrees to 1e-7 (FP64, 1000 AdamW
    steps of accumulated rounding), utilities to 1e-6."""
    from nonmyopia_b200.utils import eval_func
    fx = H.load(name)
    (u_obs, u_post, regret), A = eval_func(_env(fx), product_gp(fx["gp"]), _parms(fx), fx["buffer"], fx["iteration"],
                                           _replay_draws=fx["draws"])
    g = fx["out"]
    assert A.shape == g["A_chosen"].shape and not A.is_cuda
    assert H.scaled_err(A, g["A_chosen"]) < 1e-7
    assert abs(u_post - g["u_posterior"]) < 1e-6 and abs(regret - g["regret"]) < 1e-6
    assert u_obs == g["u_observed"]


def test_eval_func_graph_replay_equals_eager():
    """Noise-free cost makes the qEI branch deterministic: the captured-graph loop (device draws) and the eager
    loop (host draws) walk the same trajectory bit for bit."""
    from nonmyopia_b200.utils import eval_func
    fx = H.load("eval_qei_ackley")
    cost = dict(fx["cost"], max_noise=0.0)
    out = {}
    for mode in ("device", "host"):
        p = _parms(fx, rng_mode=mode, cost_func_hypers=cost)
        out[mode] = eval_func(_env(fx), product_gp(fx["gp"]), p, fx["buffer"], fx["iteration"], n_steps=60,
                              _z0=fx["draws"]["z0"])             # same initial jitter for both
    assert torch.equal(out["device"][1], out["host"][1])
    assert out["device"][0] == out["host"][0]


def test_eval_func_device_mode_runs_hes():
    """Default (device generator, graph replay) HES branch: finite result inside the unit cube's neighbourhood,
    statistically the same optimum as the host-draw run (posterior-mean ascent from the last query)."""
    from nonmyopia_b200.utils import eval_func
    fx = H.load("eval_hes_ackley")
    (u_obs, u_post, regret), A = eval_func(_env(fx), product_gp(fx["gp"]), _parms(fx), fx["buffer"], fx["iteration"])
    assert torch.isfinite(A).all() and A.shape == (1, fx["d"])
    assert abs(regret - (3 - u_post)) < 1e-12
    assert float((A - fx["out"]["A_chosen"]).abs().max()) < 0.05


def test_eval_func_refuses_cpu_and_multi_action():
    from nonmyopia_b200.utils import eval_func
    fx = H.load("eval_qei_ackley")
    with pytest.raises(NotImplementedError):
        eval_func(_env(fx), product_gp(fx["gp"]), _parms(fx, n_actions=2), fx["buffer"], fx["iteration"])
    with pytest.raises(NotImplementedError):
        eval_func(_env(fx), product_gp(fx["gp"]), _parms(fx), fx["buffer"], fx["iteration"], embedder=object())
