"""N1 evidence: would TF32 tensor-core GEMMs in the policy network meet north_star's 1e-3 tolerance?
CPU experiment on BASELINE configs[3] (Cosine8 8-D, n = 200, HES-TS-AM, L = 3 and 20): the oracle rollout in FP64 vs
the same rollout with every policy GEMM's operands rounded to TF32 (10-bit mantissa, round-to-nearest-even; FP32
accumulation emulated by rounding the products' sum to float32), and vs plain FP32 everywhere.  Prints one JSON line
per case.  Test infrastructure only (imports oracle/)."""
import json, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from nonmyopia_b200.synthetic import make_problem
from oracle import gp_math, hes_oracle as ho

F64 = torch.float64


def tf32(x):
    """round an FP32-representable tensor to TF32 (keep 10 mantissa bits, RNE)"""
    b = x.to(torch.float32).contiguous().view(torch.int32)
    r = (b + 0xFFF + ((b >> 13) & 1)) & ~0x1FFF
    return r.view(torch.float32)


def run(L, mode):
    prob = make_problem("Cosine8", seed=2)
    d, N, H = prob.d, 64, 64
    ls = torch.full((1, d), prob.lengthscale, dtype=F64)
    gp = gp_math.DenseGP.fit_caches(prob.train_X, prob.train_Y, kind=prob.kernel, lengthscale=ls,
                                    outputscale=prob.outputscale, noise=prob.noise, mean_const=prob.mean_const)
    torch.manual_seed(2)
    path = gp_math.draw_matheron_path(gp, prob.train_Y, 1)
    g = torch.Generator().manual_seed(3)
    P = ho.init_policy_params(d, H, generator=g)
    prev_x = prob.train_X[-1:].expand(N, -1).contiguous()
    prev_y = prob.train_Y[-1:].expand(N, -1).contiguous()
    prev_h = torch.randn(N, H, dtype=F64, generator=g)
    spec = ho.HESSpec(L, [1] * L, 1, ho.CostSpec(radius=prob.radius, k=1, max_noise=0.0, p_norm=2.0), enable_ts=True)
    orig = F.linear

    def lin(x, w, b=None):
        if mode == "f64" or w.shape[-1] < 16:          # the K = d + 1 embedding stays on the FMA pipe in every variant
            return orig(x, w, b)
        if mode == "f32":
            return orig(x.float(), w.float(), None if b is None else b.float()).double()
        y = orig(tf32(x), tf32(w)).float()              # TF32 operands, FP32 accumulate
        return (y + (0 if b is None else b.float())).double()

    F.linear = lin
    try:
        torch.manual_seed(11)
        o = ho.hes_forward(spec, prev_x, prev_y, prev_h, P, path=path, enable_noise=False)
    finally:
        F.linear = orig
    return o


for L in (3, 20):
    ref = run(L, "f64")
    for mode in ("f32", "tf32"):
        o = run(L, mode)
        rel = lambda a, b: float(((a - b).abs() / b.abs().clamp_min(1.0)).max())
        tot_r = ref["acqf_loss"] + (2 * math.sqrt(8) - ref["acqf_cost"]).abs()
        tot_o = o["acqf_loss"] + (2 * math.sqrt(8) - o["acqf_cost"]).abs()
        print(json.dumps(dict(case="cfg4_tf32_tolerance", lookahead=L, gemm_precision=mode,
                              rel_dev=dict(acqf_loss=rel(o["acqf_loss"], ref["acqf_loss"]),
                                           acqf_cost=rel(o["acqf_cost"], ref["acqf_cost"]),
                                           actions=rel(o["actions"], ref["actions"])),
                              tolerance=1e-3, same_argmin=bool(tot_r.argmin() == tot_o.argmin()),
                              argmin_margin=float(torch.sort(tot_r).values[1] - torch.sort(tot_r).values[0]))))
