import sys, os, time, tempfile, types
sys.path.insert(0, os.getcwd())
import torch
from tests import helpers as H
from tests.test_dropin_gpu import _parms, product_gp
from nonmyopia_b200 import actor as A
fx = dict(H.load("query_ts_am_ackley_L1"), L=3)
n = fx["gp"]["train_X"].shape[0]
for mode in ("host", "device", "host", "device"):
    parms = _parms(fx, tempfile.mkdtemp(), 500, mode)
    torch.manual_seed(0)
    act = A.Actor(parms)
    g = fx["gp"]
    buffer = dict(x=g["train_X"].cuda(), y=g["train_Y"].cuda(), h=torch.zeros(n, fx["H"], dtype=torch.float64, device="cuda"),
                  cost=torch.zeros(n, dtype=torch.float64, device="cuda"))
    act.construct_acqf(product_gp(g), buffer)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    import io, contextlib
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        act.reset_parameters(buffer)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    lines = [l for l in buf.getvalue().splitlines() if l.startswith("Loss")]
    print(mode, "reset_parameters %.3f s" % dt, "epochs printed", len(lines), lines[:1], lines[-1:], [l for l in buf.getvalue().splitlines() if "eagerly" in l])
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(buf):
        out = act.query(buffer, iteration=n, save_trajectory=False)
    torch.cuda.synchronize()
    print(mode, "query %.3f s" % (time.perf_counter() - t0), out["next_X"].tolist())
