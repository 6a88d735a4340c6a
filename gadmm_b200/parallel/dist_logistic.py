"""``backend="nccl"`` / ``"gloo"`` logistic GADMM (exact and inexact local solves): the plain
torch.distributed baseline of the logistic family.

Contiguous chain segments per rank; a rank solves the local problems of its workers of the active
colour with the oracle's own solvers (``_logistic_local_newton`` for group_ADMM_logistic.m:48-55,
``logReg_GD`` for group_ADMM_logistic_GD.m:24-52) and the ranks re-assemble the model matrix with
one ``all_gather`` per half-iteration; duals and the objective are ``all_reduce``-free because the
assembled models are replicated.  No custom kernels, no peer memory: this is the arm the
``backend="b200"`` engines (:mod:`gadmm_b200.engine.logistic`, ``logistic_scaled``) are compared
with.
"""
from __future__ import annotations

import time

import numpy as np
import torch
import torch.distributed as dist

from ..models.problems import LogisticProblem
from ..oracle.gadmm import _Timer, _logistic_local_newton, logReg_GD
from .nccl_baseline import _Ctx
from .runtime import partition


class _Chain:
    def __init__(self, XX, YY, N, d, s, lam, naive, backend):
        self.ctx = _Ctx(backend)
        self.prob = LogisticProblem(XX, YY, N, d, s, lam, naive)
        self.N, self.d = self.prob.N, self.prob.d
        world = max(self.ctx.world, 1)
        self.owner = partition(self.N, world)
        rank = self.ctx.rank if self.ctx.world > 1 else 0
        mine = np.nonzero(self.owner == rank)[0]
        self.lo, self.nloc = (int(mine[0]) if mine.size else 0), int(mine.size)
        self.counts = [int((self.owner == r).sum()) for r in range(world)]
        self.maxloc = max(self.counts)

    def local(self, colour):
        return [n for n in range(self.lo, self.lo + self.nloc) if n % 2 == colour]

    def assemble(self, theta):
        """Every rank contributes the columns it owns (the neighbour exchange of the baseline)."""
        if self.ctx.world == 1:
            return theta
        pad = torch.zeros((self.maxloc, self.d), dtype=torch.float64, device=self.ctx.device)
        pad[: self.nloc] = torch.from_numpy(np.ascontiguousarray(theta[:, self.lo:self.lo + self.nloc].T))
        parts = [torch.empty_like(pad) for _ in range(self.ctx.world)]
        dist.all_gather(parts, pad)
        full = torch.cat([p[:c] for p, c in zip(parts, self.counts)], 0).cpu().numpy()
        return np.ascontiguousarray(full.T)

    def total(self, v: int) -> int:
        if self.ctx.world == 1:
            return int(v)
        t = torch.tensor([int(v)], dtype=torch.int64, device=self.ctx.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return int(t.item())


def group_admm_logistic(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                        lambda_logistic, acc, backend="gloo", faithful=False, naive=True,
                        return_state=False, **_ignored):
    """``group_ADMM_logistic`` (exact local solves) over the process group."""
    ch = _Chain(XX, YY, no_workers, num_feature, per_samples, lambda_logistic, naive, backend)
    prob, N, d = ch.prob, ch.N, ch.d
    Xs, ys = prob.Xs, prob.ys
    lam, theta = np.zeros((d, N)), np.zeros((d, N))
    obj, loss = [], []
    timer = _Timer()
    Iter, hit = (100 if faithful else num_iter), False
    for i in range(1, num_iter + 1):
        for colour in (0, 1):
            t0 = time.perf_counter()
            for n in ch.local(colour):
                lin, centre, c = np.zeros(d), np.zeros(d), 0
                if n > 0:
                    lin -= lam[:, n - 1]
                    centre += rho * theta[:, n - 1]
                    c += 1
                if n < N - 1:
                    lin += lam[:, n]
                    centre += rho * theta[:, n + 1]
                    c += 1
                theta[:, n] = _logistic_local_newton(Xs[n], ys[n], theta[:, n], lin, c * rho, centre,
                                                     lambda_logistic, naive)
            if colour == 0:
                timer.add(i - 1, (time.perf_counter() - t0) / max(len(ch.local(0)), 1))
            theta = ch.assemble(theta)
        lam[:, : N - 1] += rho * (theta[:, : N - 1] - theta[:, 1:])
        f = prob.objective(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter, hit = i, True
            break
    n_out = Iter if hit else len(obj)
    res = (np.asarray(obj[:n_out]), np.asarray(loss[:n_out]), Iter, timer.vector(n_out))
    if return_state:
        return res + ({"theta": theta, "dual": lam, "iteration": Iter},)
    return res


def group_admm_logistic_gd(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                           lambda_logistic, acc, gdStep, backend="gloo", naive=True,
                           return_state=False, **_ignored):
    """``group_ADMM_logistic_GD`` (inexact local solves) over the process group."""
    ch = _Chain(XX, YY, no_workers, num_feature, per_samples, lambda_logistic, naive, backend)
    prob, N, d = ch.prob, ch.N, ch.d
    Xs, ys = prob.Xs, prob.ys
    lam, theta, z = np.zeros((d, N)), np.zeros((d, N)), np.zeros(d)
    obj, loss = [], []
    timer = _Timer()
    Iter, inner = num_iter, 0
    for i in range(1, num_iter + 1):
        for colour in (0, 1):
            t0 = time.perf_counter()
            for n in ch.local(colour):
                C1 = lam[:, n - 1] if n > 0 else z
                t1 = rho * (theta[:, n] - theta[:, n - 1]) if n > 0 else z
                C2 = lam[:, n] if n < N - 1 else z
                t2 = rho * (theta[:, n] - theta[:, n + 1]) if n < N - 1 else z
                theta[:, n], st = logReg_GD(Xs[n], ys[n], theta[:, n], C1, C2, t1, t2, lambda_logistic,
                                            gdStep, naive, return_steps=True)
                inner += st
            if colour == 0:
                timer.add(i - 1, (time.perf_counter() - t0) / max(len(ch.local(0)), 1))
            theta = ch.assemble(theta)
        lam[:, : N - 1] += rho * (theta[:, : N - 1] - theta[:, 1:])
        f = prob.objective(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter = i
            break
    res = (np.asarray(obj[:Iter]), np.asarray(loss[:Iter]), Iter, timer.vector(Iter))
    if return_state:
        return res + ({"theta": theta, "dual": lam, "iteration": Iter, "inner_steps": ch.total(inner)},)
    return res
