"""Per-kernel CUDA-event times of one HES-TS-AM Adam iteration (eager launches), FP64 vs the FP32 / TF32 variant.
usage: python tools/time_step_kernels.py [env] [restarts] [L]"""
import math, os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nonmyopia_b200 import ts_engine as te
from nonmyopia_b200.amortized_network import AmortizedNetwork
from nonmyopia_b200.gp import SurrogateGP, draw_thompson_path
from nonmyopia_b200.path import ThompsonPath
from nonmyopia_b200.synthetic import make_problem
from nonmyopia_b200.ts_loop import TSAMLoop
env = sys.argv[1] if len(sys.argv) > 1 else "Cosine8"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
L = int(sys.argv[3]) if len(sys.argv) > 3 else 3
F64 = torch.float64
prob = make_problem(env, seed=2)
d, H = prob.d, 64
gp = SurrogateGP(prob.train_X, prob.train_Y, kernel=prob.kernel, lengthscale=torch.full((d,), prob.lengthscale, dtype=F64),
                 outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
torch.manual_seed(2)
path64 = draw_thompson_path(gp, 1)
net = AmortizedNetwork(d + 1, d, H, 1, torch.tensor([[0.0, 1.0]] * d, dtype=F64)).double()
theta = net.flat_parameters().detach()
prev_h = torch.randn(N, H, generator=torch.Generator().manual_seed(7)).to(F64)
prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
for dt in (torch.float64, torch.float32):
    path = path64 if dt == F64 else ThompsonPath(path64.W.to(dt), path64.w.to(dt), path64.v.to(dt), path64.train_X.to(dt),
                                                 path64.lengthscale.to(dt), kernel=path64.kernel,
                                                 outputscale=path64.outputscale, mean_const=path64.mean_const)
    eng = te.TSRollout(path, d, H, L, N, dict(radius=prob.radius, k=1.0, max_noise=1e-5, p_norm=2.0), dtype=dt, seed=99)
    loop = TSAMLoop(eng, theta, prev_x, prev_y, prev_h, budget=2 * math.sqrt(d), lr=1e-3, past_initial=False,
                    max_iter=10**9, patience=-1)
    names, evs = [], []
    def mark(name):
        e = torch.cuda.Event(enable_timing=True); e.record(); names.append(name); evs.append(e)
    for _ in range(5):
        loop._iteration()
    torch.cuda.synchronize()
    for _ in range(20):
        mark("start"); loop._iteration(mark=mark)
    torch.cuda.synchronize()
    acc = {}
    for i in range(1, len(evs)):
        if names[i] != "start":
            acc.setdefault(names[i], []).append(evs[i - 1].elapsed_time(evs[i]))
    print(json.dumps(dict(env=env, restarts=N, L=L, dtype=str(dt), ms={k: round(sum(v) / len(v), 4) for k, v in acc.items()})))
