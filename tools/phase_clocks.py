"""Phase accounting of the forward rollout (build with NMB_EXTRA_NVCC_FLAGS=-DNMB_PHASE_CLOCKS, experiment only):
runs the cfg-2b iteration a few times and prints where one consumer warp per CTA spends its cycles."""
import ctypes as C, os, sys, math
sys.path.insert(0, os.getcwd())
import torch
import bench
from nonmyopia_b200 import _lib, ts_engine as te
from nonmyopia_b200.gp import SurrogateGP, draw_thompson_path
from nonmyopia_b200.ts_loop import TSAMLoop
n_r, S, Lk = 64, (int(sys.argv[1]) if len(sys.argv) > 1 else 64), (int(sys.argv[2]) if len(sys.argv) > 2 else 3)
N = n_r * S
prob, prev_x, prev_y, prev_h = bench.host_workload(n_r, S, 0)
gp = SurrogateGP(prob.train_X, prob.train_Y, kernel=prob.kernel, lengthscale=torch.full((2,), prob.lengthscale),
                 outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
torch.manual_seed(2)
path = draw_thompson_path(gp, S)
net = bench.init_policy(2, 64)
eng = te.TSRollout(path, 2, 64, Lk, N, bench.cost_hypers(), chains_per_path=n_r, seed=1)
loop = TSAMLoop(eng, net.flat_parameters().detach(), prev_x, prev_y, prev_h, budget=2 * math.sqrt(2), lr=1e-3,
                past_initial=False, n_global=N, max_iter=10**9, patience=-1)
lib = _lib.lib()
for _ in range(5):
    loop._iteration()
buf = (C.c_ulonglong * 16)()
lib.nmb_debug_phase_clocks(buf, 1)
n = 20
for _ in range(n):
    loop._iteration()
lib.nmb_debug_phase_clocks(buf, 0)
names = ["embedding", "gemm loops", "activation epilogues", "named barriers", "output layer", "rotation + path", "cost/tail",
         "input load"]
tot = sum(buf[i] for i in range(8))
ctas = buf[8]
print("CTA-warps sampled", ctas, "mean cycles per CTA", tot / max(ctas, 1))
for i, nm in enumerate(names):
    print(f"{nm:22s} {100.0 * buf[i] / tot:5.1f} %   {buf[i] / max(ctas, 1):10.0f} cycles")
