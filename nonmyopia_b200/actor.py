"""Actor - host-side mirror of the reference's ``synthetic_experiment/actor.py``:
``Actor(parms)``, ``construct_acqf``, ``reset_parameters``, ``query`` with the same arguments
and the same return dictionary.

``query`` (reference actor.py:439-734) runs the AdamW loop over the acquisition objective:

* HES-TS-AM: the device-resident loop of ``ts_loop.TSAMLoop`` - one CUDA graph per Adam
  iteration (two rollouts in one launch, objective + best-restart bookkeeping, snapshot,
  reverse rollout, split-K weight gradients, fused AdamW), no host synchronisation except a
  poll of the early-stop flag every ``poll_every`` iterations;
* every other algorithm (free-parameter HES-TS, HES-MC, qEI/qPI/qSR/qUCB/qKG): the same loop
  in eager form over ``acqf.forward`` (custom autograd nodes around the kernels) with the
  fused ``nmb_adamw_flat`` optimiser step.
"""
from __future__ import annotations

import ctypes as C
import gc
import math
import os
import pickle
import random
from typing import List, Optional

import torch
from torch import Tensor

from . import _lib
from . import sharding as sh
from . import ts_engine as te
from .acqfs import SobolQMCNormalSampler, qBOAcqf, qMultiStepHEntropySearch
from .amortized_network import AmortizedNetwork
from .ts_loop import TSAMLoop


# ---- helpers restated from the reference's utils.py:313-365 (host-side random initialisation)
def generate_random_points_batch(inputs: Tensor, radius: float, n: int) -> Tensor:
    """``n`` uniform points in the radius-ball around every input point; new leading dim ``n``.
    float32 host draws cast to the input dtype, like the reference."""
    batch, dim = inputs.shape[:-1], inputs.shape[-1]
    direction = torch.randn(n, *batch, dim).to(inputs)
    direction = direction / torch.norm(direction, dim=-1, keepdim=True)
    r = torch.rand(n, *batch).pow(1 / dim).unsqueeze(-1).to(inputs) * radius
    return inputs.unsqueeze(0).expand(n, *batch, dim) + r * direction


def generate_random_rotation_matrix(D: int) -> Tensor:
    Q, _ = torch.linalg.qr(torch.randn(D, D))
    if torch.det(Q) < 0:
        Q[:, random.randint(0, D - 1)] *= -1
    return Q


def rotate_points(inputs: Tensor, R: Tensor, p: Tensor) -> Tensor:
    return torch.matmul(inputs - p, R.transpose(-1, -2).to(inputs)) + p


def fantasy_widths(parms, lookahead: int) -> List[int]:
    """reference actor.py:89-104 (duplicated at :399-414)."""
    if parms.algo_ts:
        return [1] * lookahead
    if "MSL" in parms.algo:
        return [32, 2, 1, 1][:lookahead]
    table = {0: [], 1: [64], 2: [64, 8], 3: [64, 4, 2]}
    return table[lookahead] if lookahead in table else [64, 4, 2, 1] + [1] * (lookahead - 4)


def _logit(x: Tensor) -> Tensor:
    return -torch.log(1 / x - 1)


class Actor:
    def __init__(self, parms):
        self.parms = parms
        if parms.algo == "HES":
            self.acqf_class = qMultiStepHEntropySearch
            if parms.amortized:
                self.maps = AmortizedNetwork(
                    input_dim=parms.x_dim + parms.y_dim, output_dim=parms.x_dim, hidden_dim=parms.hidden_dim,
                    n_actions=parms.n_actions, output_bounds=parms.bounds, discrete=parms.env_discretized,
                    num_categories=parms.num_categories).to(dtype=parms.torch_dtype, device=parms.device)
                print("Number of AmortizedNet params:",
                      sum(p.numel() for p in self.maps.parameters() if p.requires_grad))
                self._parameters = list(self.maps.parameters())
            else:
                self.maps = []
        else:
            self.acqf_class = qBOAcqf
            self.maps = []
        self.algo_lookahead_steps = parms.algo_lookahead_steps
        self.acqf = None
        # "host" (default): rotation angles / cost noise follow torch's global CPU generator in the reference's order
        # (same seed => same next query as the reference run on the CPU), fed to the CUDA-graph loop through a
        # device ring; "device": drawn in-kernel (Philox), no host work
        self.rng_mode = getattr(parms, "rng_mode", "host")
        # restarts sharded over the ranks of this torch.distributed group (None: single GPU); SURVEY 8e
        self.group = getattr(parms, "process_group", None)

    # ------------------------------------------------------------------ construct_acqf
    def construct_acqf(self, surr_model, buffer, **kwargs):
        """reference actor.py:382-437."""
        self.acqf = None
        gc.collect()
        p = self.parms
        nf = fantasy_widths(p, self.algo_lookahead_steps)
        sampler = SobolQMCNormalSampler(sample_shape=p.n_samples, seed=0, resample=False) if p.algo != "HES" else None
        self.acqf = self.acqf_class(
            name=p.algo, model=surr_model, lookahead_steps=self.algo_lookahead_steps, n_actions=p.n_actions,
            n_fantasy_at_design_pts=nf, n_fantasy_at_action_pts=p.n_samples,
            loss_function_class=p.loss_function_class, loss_func_hypers=p.loss_func_hypers,
            cost_function_class=p.cost_function_class, cost_func_hypers=p.cost_func_hypers,
            enable_ts=p.algo_ts, sampler=sampler, best_f=buffer["y"].max(), rng_mode=self.rng_mode)

    # ------------------------------------------------------------------ reset_parameters
    def _random_tree(self, prev_point: Tensor, nf: List[int], n_first: int, clamp_designs: bool = False) -> List[Tensor]:
        """Random designs level by level, each in the radius-ball of its parent
        (reference actor.py:116-160 / 226-293).  Returns L+1 host tensors [cnt_s, d].  ``clamp_designs``: the
        amortized branch clamps every design level to [0.01, 0.99] BEFORE the next level is drawn around it."""
        p = self.parms
        L, r = self.algo_lookahead_steps, p.cost_func_hypers["radius"]
        out: List[Tensor] = []
        for s in range(L + 1):
            parent = prev_point if s == 0 else out[-1]
            if s == 0:
                n = n_first if L > 0 else n_first * p.n_actions
            elif s < L:
                n = nf[s - 1]
            else:
                n = nf[-1] * p.n_actions
            x = generate_random_points_batch(parent, r, n)
            if clamp_designs and s < L:
                x = torch.clamp(x, max=0.99, min=0.01)
            out.append(x)
        return out

    def reset_parameters(self, buffer, bo_iter=0, embedder=None, prev_chosen_idx=0):
        """reference actor.py:66-380.  Amortized: imitation pre-training of the network on
        random ball designs (<= 300 AdamW epochs); otherwise random (or rotated previous-best)
        designs stored as logits."""
        if embedder is not None:
            raise NotImplementedError("discrete environments are outside the B200 hot-path scope")
        print("Resetting actor parameters...")
        p = self.parms
        L, d = self.algo_lookahead_steps, p.x_dim
        prev_X = buffer["x"][-1:].expand(p.n_restarts, -1)
        prev_y = buffer["y"][-1:].expand(p.n_restarts, -1)
        nf = fantasy_widths(p, L)
        dev, dt = p.device, p.torch_dtype
        if p.amortized:
            self._pretrain(buffer, prev_X, prev_y, nf)
            return
        rank, world = sh.world_of(self.group)
        lo, hi = sh.restart_range(p.n_restarts, rank, world)
        if bo_iter == 0 or not p.algo_ts:
            tree = self._random_tree(prev_X[0], nf, p.n_restarts)
            full = [torch.clamp(x.reshape(-1, d).to(device=dev, dtype=dt), min=0.01, max=0.99) for x in tree]
        else:
            # warm start: rotate the previous best trajectory about its first point (actor.py:294-379)
            maps = [torch.sigmoid(sh.gather_level(x.detach(), p.n_restarts, self.group)) for x in self.maps]
            idx = prev_chosen_idx.long() if torch.is_tensor(prev_chosen_idx) else int(prev_chosen_idx)
            R = [torch.eye(d).to(prev_X)] + [generate_random_rotation_matrix(d).to(prev_X)
                                             for _ in range(p.n_restarts - 1)]
            R = torch.stack(R, dim=0)
            maps[0][:] = maps[1][idx]
            for lah in range(2, L + 1):
                pts = maps[lah].reshape(*nf[:lah], p.n_restarts, d)[..., idx, :]
                rot = rotate_points(pts.reshape(-1, d), R, maps[0][idx]).transpose(0, 1)
                maps[lah - 1] = torch.clamp(rot, min=0.01, max=0.99).reshape(-1, d)
            a = generate_random_points_batch(maps[L - 1], p.cost_func_hypers["radius"], nf[-1] * p.n_actions)
            maps[L] = torch.clamp(a.to(device=dev, dtype=dt), min=0.01, max=0.99).reshape(-1, d)
            full = maps
        # every rank draws the full-size initialisation from the same host seed and keeps its restarts (8e)
        self.maps = [_logit(sh.shard_level(x, p.n_restarts, lo, hi) if world > 1 else x).requires_grad_(True)
                     for x in full]
        self._parameters = self.maps

    def _pretrain(self, buffer, prev_X, prev_y, nf):
        """reference actor.py:106-220: fit the network to random ball designs, with draws of
        the surrogate as the observations fed to the next step."""
        p = self.parms
        L, d, n = self.algo_lookahead_steps, p.x_dim, p.n_restarts
        dev, dt = p.device, p.torch_dtype
        prev_h = buffer["h"][-1:].clone().expand(n, -1)
        opt = torch.optim.AdamW(self._parameters, lr=0.01, capturable=True)
        tree = self._random_tree(prev_X[0:1], nf, n, clamp_designs=True)
        # design levels were clamped on the host in their float32 form; the action level after the cast
        # (reference actor.py:126-137 vs 158-162)
        target = torch.stack([x.to(device=dev, dtype=dt).reshape(-1, d) if s < L else
                              torch.clamp(x.to(device=dev, dtype=dt).reshape(-1, d), min=0.01, max=0.99)
                              for s, x in enumerate(tree)], dim=0).detach()
        self.hidden_noise = torch.randn(n, p.hidden_dim).to(target[-1])
        prev_h = prev_h + self.hidden_noise
        gp = self.acqf.gp()
        x0, y0 = prev_X[0].expand(n, d), prev_y[0].expand(n, p.y_dim)
        from . import pretrain as pt
        if pt.supported(d, p.hidden_dim, dt, dev) and p.y_dim == 1 and p.n_actions == 1:
            # own kernels: the epoch (fused policy steps, surrogate draws, L1 seeds, reverse rollout, weight gradients,
            # AdamW, early-stop rule) is one CUDA graph; nothing is read back before the loop ends
            seed = int(torch.randint(0, 2**31 - 1, (1,)).item()) if self.rng_mode == "device" else 0
            loop = pt.PretrainLoop(gp, self.maps.flat_parameters().detach(), x0, y0, prev_h, target, H=p.hidden_dim,
                                   lr=0.01, max_epochs=300, patience=50, rng=self.rng_mode, seed=seed)
            st = loop.run()
            o = 0
            with torch.no_grad():
                for prm in self.maps.parameters():
                    prm.copy_(loop.theta[o:o + prm.numel()].view_as(prm))
                    o += prm.numel()
            self.pretrain_losses = loop.losses(int(st.step))
            for ep in range(0, int(st.step), 10):
                print("Loss:", float(self.pretrain_losses[ep]))
            return

        # shapes outside the tensor-pipe kernel (float32 variant, odd d): the same epoch as device tensor ops
        def epoch():
            """One imitation epoch, device work only: rollout of the network on draws of the surrogate, L1 loss to
            the random ball designs, backward, AdamW."""
            outs, h, x, y = [], prev_h, x0, y0
            for j in range(L):
                x, h = self.maps(x, y, h, return_actions=False, enable_noise=False)
                outs.append(x)
                y = _prior_free_sample(gp, x.detach()).reshape(-1, 1)
            a, h = self.maps(x, y, h, return_actions=True, enable_noise=False)
            outs.append(a)
            loss = torch.mean(abs(target - torch.stack(outs, dim=0)))
            loss.backward()
            opt.step()
            return loss.detach()

        # The epoch is ~300 small launches: after three eager epochs (they count) it is captured once and replayed
        # as a CUDA graph; the host reads the loss every epoch for the reference's early-stop rule (actor.py:200-216).
        graph, static_loss, side = None, None, torch.cuda.Stream(device=target.device)
        stall, best = 0, float("inf")
        for ep in range(300):
            if graph is None and ep == 3 and self.rng_mode != "host":
                try:
                    opt.zero_grad(set_to_none=True)
                    torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        static_loss = epoch()
                    graph = g
                    graph.replay()              # capture only records: this replay is epoch 3
                    loss = static_loss
                except Exception as e:          # capture unsupported for some op: keep the eager epochs
                    print("pre-training runs eagerly (graph capture failed: %s)" % type(e).__name__)
                    torch.cuda.synchronize()
                    graph, static_loss = False, None
                    opt.zero_grad(set_to_none=True)
                    loss = epoch()
            elif graph:
                graph.replay()
                loss = static_loss
            elif ep < 3 and self.rng_mode != "host":
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    opt.zero_grad(set_to_none=True)
                    loss = epoch()
                torch.cuda.current_stream().wait_stream(side)
            else:
                opt.zero_grad(set_to_none=True)
                loss = epoch()
            loss_v = float(loss)
            if ep % 10 == 0:
                print("Loss:", loss_v)
            if loss_v < best:
                best, stall = loss_v, 0
            else:
                stall += 1
            if stall >= 50:
                break
        torch.cuda.current_stream().wait_stream(side)

    # ------------------------------------------------------------------ query
    def query(self, buffer, iteration: int, embedder=None, f_posterior=None, save_trajectory=True):
        """reference actor.py:439-734: optimise the acquisition objective over all restarts and
        return the best restart's next design."""
        assert self.acqf is not None, "Acquisition function is not initialized."
        if embedder is not None:
            raise NotImplementedError("discrete environments are outside the B200 hot-path scope")
        p = self.parms
        rank, world = sh.world_of(self.group)
        lo, hi = sh.restart_range(p.n_restarts, rank, world)
        n_r = hi - lo                                   # this rank's restarts (all of them on one GPU)
        prev_X = buffer["x"][iteration - 1:iteration].expand(n_r, -1)
        prev_y = buffer["y"][iteration - 1:iteration].expand(n_r, -1)
        prev_h = buffer["h"][iteration - 1:iteration].expand(n_r, -1)
        if p.amortized:
            prev_h = prev_h + self.hidden_noise[lo:hi]
        prev_cost = (float(buffer["cost"][p.n_initial_points:iteration].sum())
                     if iteration > p.n_initial_points else 0.0)
        is_hes = p.algo.startswith("HES")
        if is_hes:
            self.acqf.dump_model(f=f_posterior)
            if p.algo_ts and f_posterior is None and rank == 0:
                with open(os.path.join(p.save_dir, f"posterior_sample_{iteration}.pkl"), "wb") as fh:
                    pickle.dump(self.acqf.f, fh)
        past_initial = iteration > p.n_initial_points
        if p.amortized:
            res = self._query_amortized(prev_X, prev_y, prev_h, prev_cost, past_initial, save_trajectory)
        else:
            res = self._query_free(prev_X, prev_y, prev_h, prev_cost, past_initial, save_trajectory)
        if is_hes:
            self.acqf.clean_dump_model()
        best_loss, best_cost, best_KL = res["best_loss"], res["best_cost"], res["best_KL"]
        if world > 1:
            return self._select_sharded(res, buffer, iteration, prev_cost, save_trajectory)
        idx = torch.argmin(best_loss + best_cost + best_KL)
        if save_trajectory:
            with open(p.save_dir + f"/trajectory_{iteration}.pkl", "wb") as fh:
                pickle.dump((res["saved_trajectory"], idx), fh)
            with open(p.save_dir + f"/lossncost_{iteration}.pkl", "wb") as fh:
                pickle.dump((res["saved_loss"], res["saved_cost"]), fh)
        next_X = res["best_next_X"][0][idx].reshape(1, -1)
        actions = res["best_actions"][..., idx, :, :]
        hidden_state = res["best_hidden_state"][0][idx:idx + 1] if p.amortized else None
        acqf_cost = self.acqf.cost_function(prev_X=buffer["x"][iteration - 1:iteration], current_X=next_X,
                                            previous_cost=prev_cost).detach()
        return {"all_losses": best_loss, "all_costs": best_cost, "loss": best_loss[idx], "cost": acqf_cost,
                "next_X": next_X, "actions": actions, "hidden_state": hidden_state, "chosen_idx": idx}

    def _select_sharded(self, res, buffer, iteration, prev_cost, save_trajectory):
        """Final selection over ALL ranks' restarts: all-gather of the per-restart totals,
        arg-min, the owner broadcasts the chosen restart's next_X / actions / hidden state.  Every rank
        returns the same dictionary as the unsharded ``query``."""
        p = self.parms
        n_all, d = p.n_restarts, p.x_dim
        total = res["best_loss"] + res["best_cost"] + res["best_KL"]
        _, idx, owner, li = sh.select_best(total, n_all, self.group)
        all_losses = sh.gather_restarts(res["best_loss"], n_all, self.group)
        all_costs = sh.gather_restarts(res["best_cost"], n_all, self.group)
        dev, dt = total.device, total.dtype
        acts = res["best_actions"]
        a_shape = (*acts.shape[:-3], *acts.shape[-2:])
        bc = lambda t, shape: sh.broadcast_from_owner(t, shape, dt, dev, owner, self.group)
        mine = sh.world_of(self.group)[0] == owner
        next_X = bc(res["best_next_X"][0][li].reshape(1, -1) if mine else None, (1, d))
        actions = bc(acts[..., li, :, :] if mine else None, a_shape)
        hidden_state = None
        if p.amortized:
            hidden_state = bc(res["best_hidden_state"][0][li:li + 1] if mine else None, (1, p.hidden_dim))
        idx_t = torch.tensor(idx, device=dev)
        if save_trajectory:
            tag = f"{iteration}.rank{sh.world_of(self.group)[0]}"
            with open(p.save_dir + f"/trajectory_{tag}.pkl", "wb") as fh:
                pickle.dump((res["saved_trajectory"], idx_t), fh)
            with open(p.save_dir + f"/lossncost_{tag}.pkl", "wb") as fh:
                pickle.dump((res["saved_loss"], res["saved_cost"]), fh)
        acqf_cost = self.acqf.cost_function(prev_X=buffer["x"][iteration - 1:iteration], current_X=next_X,
                                            previous_cost=prev_cost).detach()
        return {"all_losses": all_losses, "all_costs": all_costs, "loss": all_losses[idx], "cost": acqf_cost,
                "next_X": next_X, "actions": actions, "hidden_state": hidden_state, "chosen_idx": idx_t}

    def _query_amortized(self, prev_X, prev_y, prev_h, prev_cost, past_initial, save_trajectory):
        """HES-TS-AM: device-resident AdamW loop (one CUDA graph per iteration)."""
        p, acqf = self.parms, self.acqf
        L, d, H, N = self.algo_lookahead_steps, p.x_dim, p.hidden_dim, prev_X.shape[0]
        eng = acqf._ts_engine(acqf.f, d, H, N, p.torch_dtype)
        eng.seed = acqf.rng_seed
        theta0 = self.maps.flat_parameters().detach()
        record = bool(save_trajectory)
        rank, world = sh.world_of(self.group)
        own = sh.restart_range(p.n_restarts, rank, world)
        eng.chain_offset = own[0]
        if world > 1:
            # replicated by construction (same host seed on every rank): checked, not assumed
            f = acqf.f
            chk = torch.stack([theta0.double().sum(), theta0.double().abs().sum(),
                               self.hidden_noise.double().sum(), self.hidden_noise.double().abs().sum(),
                               f.W.double().sum(), f.w.double().sum(), f.v.double().sum()])
            sh.assert_replicated(chk, self.group, "policy weights / hidden noise / Thompson path")
        recorded = getattr(self, "_recorded_draws", None)      # tests: explicit per-iteration draws (golden fixtures)
        stream = self.rng_mode == "host" and p.torch_dtype == torch.float64 and recorded is None
        if self.rng_mode == "host" and recorded is None:
            # the reference builds its frozen copy of the network twice before the loop (actor.py:524-555); the
            # random initialisation of those modules consumes the host generator ahead of the loop's draws
            for _ in range(2):
                AmortizedNetwork(input_dim=p.x_dim + p.y_dim, output_dim=p.x_dim, hidden_dim=p.hidden_dim,
                                 n_actions=p.n_actions, output_bounds=p.bounds)
        loop = TSAMLoop(eng, theta0, prev_X, prev_y, prev_h, budget=float(p.budget), lr=p.acq_opt_lr,
                        past_initial=past_initial, prev_cost=prev_cost, max_iter=p.acq_opt_iter,
                        record_history=record, record_trajectory=record, n_global=p.n_restarts, group=self.group,
                        rng="host" if stream else "device", own=own, exchange=getattr(p, "exchange", "auto"))
        if recorded is not None:
            st = loop.run(p.acq_opt_iter, draws_per_iter=recorded)
        elif self.rng_mode == "host" and not stream:
            # float32 variant: torch's float32 normal stream (AVX2 normal_fill) is drawn by torch itself, eagerly
            draws = _HostDraws(acqf, L, p.n_restarts, d, prev_X, own)
            st = loop.run(p.acq_opt_iter, draws_per_iter=draws)
        else:
            st = loop.run(p.acq_opt_iter)
        n_done = int(st.step)
        # last-iterate weights go back into the module: the reference's best_map holds references to
        # the live parameters, so its load_state_dict(best_map) is a no-op
        flat = loop.theta
        o = 0
        with torch.no_grad():
            for prm in self.maps.parameters():
                prm.copy_(flat[o:o + prm.numel()].view_as(prm))
                o += prm.numel()
        r = loop.result()
        out = dict(best_loss=r["best_loss"], best_cost=r["best_cost"], best_KL=r["best_KL"],
                   best_next_X=[loop.best_X[s].reshape(*([1] * s), N, 1, d) for s in range(L)],
                   best_actions=loop.best_A.reshape(*([1] * L), N, p.n_actions, d),
                   best_hidden_state=[loop.best_HS[s + 1] for s in range(L)],
                   saved_trajectory=[], saved_loss=[], saved_cost=[])
        if record:
            out["saved_loss"] = loop.hist_loss[:n_done].tolist()
            out["saved_cost"] = loop.hist_cost[:n_done].tolist()
            tr = loop.hist_X[:n_done].cpu()
            out["saved_trajectory"] = [[tr[e, s].reshape(*([1] * s), N, 1, d).tolist() for s in range(L)]
                                       for e in range(n_done)]
        return out

    def _query_free(self, prev_X, prev_y, prev_h, prev_cost, past_initial, save_trajectory):
        """Free-parameter algorithms: eager loop over ``acqf.forward`` + fused AdamW kernel
        (``sharding.free_restart_loop`` restates actor.py:558-656; the parameters hold this rank's restarts)."""
        p, acqf = self.parms, self.acqf
        params = list(self._parameters)
        opt = _FlatAdamW(params, lr=p.acq_opt_lr)

        def objective(local_maps):
            return acqf.forward(prev_X=prev_X, prev_y=prev_y, prev_hid_state=prev_h, maps=local_maps, embedder=None,
                                prev_cost=prev_cost, enable_noise=False)

        # The iteration is replayed as one CUDA graph.  "device": the objective draws its cost noise with the
        # graph-safe device generator.  "host": the noise of every iteration is drawn on torch's CPU generator in the
        # reference's order (nmb_host_rng_randn_f64, ahead of time on a host thread) and copied into the static
        # buffer the graph reads.  Sharded: every rank walks the full stream and keeps its restarts' entries.
        use_graph = params[0].is_cuda
        feed = None
        if self.rng_mode == "host":
            from .host_rng import HostNoiseFeed
            rank, world = sh.world_of(self.group)
            lo, hi = sh.restart_range(p.n_restarts, rank, world)
            feed = HostNoiseFeed(acqf.cost_noise_sizes(p.n_restarts), prev_X.device, prev_X.dtype,
                                 shard=(p.n_restarts, lo, hi) if world > 1 else None)
            acqf._fed_noise = feed.dev
        try:
            return sh.free_restart_loop(objective, params, opt.step, budget=p.budget, n_iter=p.acq_opt_iter,
                                        past_initial=past_initial, n_global=p.n_restarts, group=self.group,
                                        save_trajectory=save_trajectory, use_graph=use_graph,
                                        pre_iteration=feed.fill if feed is not None else None)
        finally:
            if feed is not None:
                feed.close()
                acqf._fed_noise = None


def _prior_free_sample(gp, x: Tensor) -> Tensor:
    """One draw of the surrogate's marginal posterior at each point of ``x`` [n, d]: the
    reference calls ``model(x.unsqueeze(-2)).sample()`` (actor.py:186-191), i.e. independent
    q = 1 posteriors."""
    k = gp.k(x, gp.train_X)                                    # [n, n_train]
    u = k @ gp.Linv.T
    mean = gp.mean_const + u @ gp.beta
    var = (gp.amplitude - (u * u).sum(-1)).clamp_min(1e-12)
    return mean + var.sqrt() * torch.randn(x.shape[0], dtype=x.dtype, device=x.device)


class _HostDraws:
    """Per-iteration draws in the reference's order: main rollout (action rotation, cost noise),
    then the frozen "original" rollout (every call rotated)."""

    def __init__(self, acqf, L, N, d, like, own=None):
        self.args = (acqf, L, N, d, like)
        self.own = own              # (lo, hi): this rank's restarts of the full-size draws

    def __len__(self):
        return 1 << 30

    def __getitem__(self, ep):
        acqf, L, N, d, like = self.args
        out = (*acqf._host_draws(L, N, d, like, False), *acqf._host_draws(L, N, d, like, True))
        if self.own is not None and self.own != (0, N):
            lo, hi = self.own
            out = tuple(t[:, lo:hi].contiguous() for t in out)
        return out


class _FlatAdamW:
    """torch.optim.AdamW semantics (actor.py:489,646-648) on a list of free tensors through the
    ``nmb_adamw_flat`` kernel; a ``None`` gradient leaves its tensor untouched."""

    def __init__(self, params, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        self.params = params
        self.hp = te.AdamWStruct(lr=lr, beta1=betas[0], beta2=betas[1], eps=eps, weight_decay=weight_decay)
        self.m = [torch.zeros_like(q) for q in params]
        self.v = [torch.zeros_like(q) for q in params]
        self.t = [0] * len(params)
        self.state = [torch.zeros(te.LOOP_STATE_BYTES, dtype=torch.uint8, device=q.device) for q in params]

    def step(self, grads):
        lib = _lib.lib()
        for i, (q, g) in enumerate(zip(self.params, grads)):
            if g is None:
                continue
            _lib.check(lib.nmb_adamw_flat(_lib.dtype_code(q.dtype), C.c_int64(q.numel()), C.byref(self.hp),
                                          _lib.ptr(q.data), _lib.ptr(self.m[i]), _lib.ptr(self.v[i]),
                                          _lib.ptr(g.contiguous()), C.c_void_p(self.state[i].data_ptr()), 1,
                                          _lib.stream_ptr()), "nmb_adamw_flat")
