"""Time ``nonmyopia_b200.utils.eval_func`` (1000 AdamW steps, graph replay) at the reference's shape
(Ackley 2D, 64 restarts) and a larger one; prints one JSON line per case."""
import json
import os
import sys
import time
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from nonmyopia_b200 import acqfs
from nonmyopia_b200.gp import SurrogateGP
from nonmyopia_b200.synthetic import make_problem, objective
from nonmyopia_b200.utils import eval_func


def run(env, n, n_r, algo, look):
    prob = make_problem(env, seed=2)
    X, Y = prob.train_X[-n:], prob.train_Y[-n:]
    d = prob.d
    ls = torch.full((d,), float(prob.lengthscale), dtype=torch.float64)
    gp = SurrogateGP(X, Y, kernel=prob.kernel, lengthscale=ls, outputscale=prob.outputscale, noise=prob.noise,
                     mean_const=prob.mean_const)
    parms = types.SimpleNamespace(
        algo=algo, algo_lookahead_steps=look, x_dim=d, n_actions=1, n_restarts=n_r, n_initial_points=n,
        acq_opt_lr=1e-2, torch_dtype=torch.float64, loss_function_class=acqfs.qLossFunctionTopK,
        loss_func_hypers=dict(dist_weight=1, dist_threshold=0.5), cost_function_class=acqfs.qCostFunction,
        cost_func_hypers=dict(radius=prob.radius, k=1, max_noise=1e-5, p_norm=2.0))
    buffer = dict(x=X, y=Y, cost=torch.zeros(n, dtype=torch.float64))
    envf = lambda a: objective(env, a.reshape(-1, d).cpu())
    ts = []
    for _ in range(3):
        torch.cuda.synchronize()
        t = time.perf_counter()
        out = eval_func(envf, gp, parms, buffer, n - 1)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t)
    print(json.dumps(dict(env=env, n_train=n, n_restarts=n_r, algo=algo, seconds=min(ts),
                          ms_per_step=min(ts), u_posterior=out[0][1])), flush=True)


if __name__ == "__main__":
    run("Ackley", 50, 64, "HES-TS-AM-20", 3)
    run("Ackley", 50, 64, "qEI", 0)
    run("Hartmann", 500, 256, "HES-MC-3", 3)
