"""Baseline star ADMM with an ``all_reduce`` for the server's reduction (standared_ADMM.m:57-88)
and a broadcast-free replicated server solve.  Library ops only; runs on NCCL or gloo."""
from __future__ import annotations

import time

import numpy as np
import torch
import torch.distributed as dist

from .nccl_baseline import _Ctx
from .runtime import partition


def star_admm(XX, YY, rho, N, d, s, num_iter, obj0, acc, backend="nccl"):
    ctx = _Ctx(backend)
    dev = ctx.device
    owner = partition(N, ctx.world)
    gids = np.nonzero(owner == ctx.rank)[0]
    XX = np.asarray(XX, dtype=np.float64)
    YY = np.asarray(YY, dtype=np.float64).reshape(-1)

    def shard(g):
        X = torch.from_numpy(XX[g * s:(g + 1) * s, :d]).to(dev)
        y = torch.from_numpy(YY[g * s:(g + 1) * s]).to(dev)
        return X.T @ X, X.T @ y, 0.5 * float(y @ y)

    eye = torch.eye(d, dtype=torch.float64, device=dev)
    wk = [g for g in gids if g < N - 1]
    nW = len(wk)
    if nW:
        Gs, bs, hs = zip(*[shard(g) for g in wk])
        G, b = torch.stack(Gs), torch.stack(bs)
        hyy = torch.tensor(hs, dtype=torch.float64, device=dev)
        M = torch.cholesky_inverse(torch.linalg.cholesky(G + rho * eye))
    else:
        G = M = torch.zeros((0, d, d), dtype=torch.float64, device=dev)
        b = torch.zeros((0, d), dtype=torch.float64, device=dev)
        hyy = torch.zeros((0,), dtype=torch.float64, device=dev)
    Gsv, bsv, hsv = shard(N - 1)            # server shard replicated on every rank
    Ms = torch.cholesky_inverse(torch.linalg.cholesky(Gsv + max(N - 1, 1) * rho * eye))
    lam = torch.zeros((nW, d), dtype=torch.float64, device=dev)
    z = torch.zeros((d,), dtype=torch.float64, device=dev)
    obj, loss, tv = [], [], []
    Iter = num_iter
    for i in range(1, num_iter + 1):
        th = torch.bmm(M, (b - lam + rho * z).unsqueeze(2)).squeeze(2)
        u = (lam + rho * th).sum(0)
        if ctx.world > 1:
            dist.all_reduce(u, op=dist.ReduceOp.SUM)
        z = Ms @ (bsv + u)
        lam = lam + rho * (th - z)
        f = (0.5 * torch.einsum("nd,nde,ne->", th, G, th) - (b * th).sum() + hyy.sum()).reshape(1)
        if ctx.world > 1:
            dist.all_reduce(f, op=dist.ReduceOp.SUM)
        fs = float(f.item()) + float(0.5 * z @ (Gsv @ z) - bsv @ z) + hsv
        obj.append(fs)
        loss.append(abs(fs - obj0))
        tv.append(time.perf_counter())
        if loss[-1] < acc:
            Iter = i
            break
    t = np.asarray(tv[:Iter])
    return np.asarray(obj[:Iter]), np.asarray(loss[:Iter]), Iter, t - t[0]
