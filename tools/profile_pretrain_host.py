"""cProfile of Actor.reset_parameters in host-stream mode (where does the host time go?)."""
import cProfile, io, os, pstats, sys, tempfile, contextlib
sys.path.insert(0, os.getcwd())
import torch
from tests import helpers as H
from tests.test_dropin_gpu import _parms, product_gp
from nonmyopia_b200 import actor as A
fx = dict(H.load("query_ts_am_ackley_L1"), L=3)
n = fx["gp"]["train_X"].shape[0]
g = fx["gp"]
buffer = dict(x=g["train_X"].cuda(), y=g["train_Y"].cuda(), h=torch.zeros(n, fx["H"], dtype=torch.float64, device="cuda"),
              cost=torch.zeros(n, dtype=torch.float64, device="cuda"))
for rep in range(2):
    parms = _parms(fx, tempfile.mkdtemp(), 500, "host")
    torch.manual_seed(0)
    act = A.Actor(parms)
    act.construct_acqf(product_gp(g), buffer)
    pr = cProfile.Profile()
    with contextlib.redirect_stdout(io.StringIO()):
        pr.enable(); act.reset_parameters(buffer); torch.cuda.synchronize(); pr.disable()
    if rep == 1:
        s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(25); print(s.getvalue())
