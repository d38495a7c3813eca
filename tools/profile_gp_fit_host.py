import cProfile, pstats, sys, time, torch
sys.path.insert(0, "/root/repo")
from nonmyopia_b200.synthetic import make_problem
from nonmyopia_b200.gp_fit import fit_exact_mll
p = make_problem("Ackley", seed=2)
X, Y = p.train_X, p.train_Y
for _ in range(2): fit_exact_mll(X, Y)
torch.cuda.synchronize()
t0=time.perf_counter(); gp, raw, res = fit_exact_mll(X, Y); torch.cuda.synchronize(); print("fit s", time.perf_counter()-t0, res.nfev)
pr = cProfile.Profile(); pr.enable(); fit_exact_mll(X, Y); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(30)
