"""Per-iteration wall time of the free-parameter loop (HES-TS free form, HES-MC) eager vs CUDA-graph replay."""
import os, sys, time
sys.path.insert(0, os.getcwd())
import torch
from tests import helpers as H
from tests.test_dropin_gpu import make_acqf
from tests.test_path_gpu import _product_path
from nonmyopia_b200 import sharding as sh
from nonmyopia_b200.actor import _FlatAdamW
F64 = torch.float64
for name, reps in (("ts_free_ackley_L2", 8), ("mc_hartmann_L3", 5)):
    fx = H.load(name)
    ts = fx["kind"] == "ts_free"
    L, d = fx["L"], fx["d"]
    n0 = fx["n_r"]
    N = n0 * reps
    acqf = make_acqf(fx, [1] * L if ts else list(fx["nf"]), ts, n_samples=fx.get("S_a", 64))
    acqf.dump_model(f=_product_path(fx) if ts else None)
    prev_X, prev_y = fx["prev_X"].repeat(reps, 1).cuda(), fx["prev_y"].repeat(reps, 1).cuda()
    prev_h = torch.zeros(N, 4, dtype=F64, device="cuda")
    logits = [sh.unshard_level([l] * reps, [n0] * reps) for l in fx["logits"]]
    for use_graph in (False, True, False, True):
        params = [l.clone().cuda().requires_grad_(True) for l in logits]
        opt = _FlatAdamW(params, lr=1e-2)
        objective = lambda maps: acqf.forward(prev_X=prev_X, prev_y=prev_y, prev_hid_state=prev_h, maps=maps, prev_cost=0.0,
                                              enable_noise=False)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        res = sh.free_restart_loop(objective, params, opt.step, budget=fx["budget"], n_iter=200, past_initial=False, log=None,
                                   use_graph=use_graph)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print(f"{name} n_r={N} graph={use_graph}: {dt / res['n_iter'] * 1e3:.3f} ms/iteration ({res['n_iter']} iterations)")
