"""Seconds per GP hyper-parameter fit on the device vs the CPU oracle, one JSON line per problem."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nonmyopia_b200.gp_fit import fit_exact_mll
from nonmyopia_b200.synthetic import make_problem
from oracle import gp_fit as ofit           # CPU baseline leg only

for env in ("Ackley", "Cosine8", "Hartmann"):
    p = make_problem(env, seed=2)
    X, Y = p.train_X, p.train_Y
    fit_exact_mll(X[:20], Y[:20])           # warm-up (library load, allocator)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    gp, raw, res = fit_exact_mll(X, Y)
    torch.cuda.synchronize()
    t_gpu = time.perf_counter() - t0
    torch.set_num_threads(os.cpu_count() or 1)
    t0 = time.perf_counter()
    h, raw_ref, res_ref = ofit.fit(X, Y)
    t_cpu = time.perf_counter() - t0
    print(json.dumps(dict(case="gp_fit", env=env, n=X.shape[0], d=X.shape[1], seconds_b200=t_gpu, evals_b200=res.n_device_evals,
                          ms_per_eval_b200=1e3 * t_gpu / res.n_device_evals, seconds_cpu_oracle=t_cpu, evals_cpu=int(res_ref.nfev),
                          cores=os.cpu_count(), objective=float(res.fun), objective_cpu=float(res_ref.fun),
                          max_raw_diff=float(abs(raw - raw_ref.numpy()).max()))), flush=True)
