"""Shared test helpers: rebuild oracle objects from golden fixtures (no /root/reference needed)."""
import os

import torch

from oracle import gp_math, hes_oracle as ho

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def dense_gp(g):
    d = g["train_X"].shape[-1]
    ls = torch.full((1, d if g["ard"] else 1), g["lengthscale"], dtype=torch.float64)
    return gp_math.DenseGP.fit_caches(g["train_X"], g["train_Y"], kind=g["kernel"], lengthscale=ls,
                                      outputscale=g["outputscale"], noise=g["noise"],
                                      mean_const=g["mean_const"])


def oracle_path(fx):
    g = fx["gp"]
    d = g["train_X"].shape[-1]
    ls = torch.full((1, d if g["ard"] else 1), g["lengthscale"], dtype=torch.float64)
    p = fx["path"]
    return gp_math.MatheronPath(p["W"], p["w"], p["v"], g["train_X"], kind=g["kernel"], lengthscale=ls,
                                outputscale=g["outputscale"], mean_const=g["mean_const"])


def cost_spec(c):
    return ho.CostSpec(radius=c["radius"], k=c["k"], max_noise=c["max_noise"], p_norm=c["p_norm"],
                       discount=c["discount"], discount_threshold=c["discount_threshold"])


def rel_err(a, b, atol=0.0):
    """max |a-b| / max(|b|, atol) elementwise."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    den = torch.clamp(b.abs(), min=atol) if atol > 0 else b.abs().clamp_min(1e-300)
    return float(((a - b).abs() / den).max()) if a.numel() else 0.0


def scaled_err(a, b):
    """max |a-b| / max(1, max|b|)."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    if a.numel() == 0:
        return 0.0
    return float((a - b).abs().max() / max(1.0, float(b.abs().max())))
