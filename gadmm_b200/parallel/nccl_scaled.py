"""Timed baselines at the scaled shape: every algorithm of BASELINE.json config 2 written with
library ops only -- dense fp64 matrices, ``torch.bmm`` (cuBLAS) for the local products,
``torch.distributed`` NCCL ``all_reduce`` / ``batch_isend_irecv`` for every exchange, the stop test
on the host.  None of the hand-written kernels run here.  This is the implementation the product
is measured against (``benchmarks/config2_arms.py --impl nccl``); the data set is the one of
``gadmm_b200.engine.scaled.worker_grams`` (same Philox streams, any partition).

Semantics follow the oracle (GD_DGD_LAG.m, dual_averaging.m, standared_ADMM.m) including the
ones-initialised gradients and the iteration-1 guards; dual averaging is run in Jacobi order (the
reference's Gauss-Seidel sweep is sequential along the chain and has no sensible library-op form).
"""
from __future__ import annotations

import time

import numpy as np
import torch
import torch.distributed as dist

from .runtime import partition


def _sync_time(dev):
    torch.cuda.synchronize(dev)
    return time.perf_counter()


class DenseWorkers:
    """This rank's workers as dense Gram matrices (N_loc, d, d), b (N_loc, d), 0.5 y'y."""

    def __init__(self, gids, d, s, D, seed=0):
        from ..engine import scaled as SC
        from ..ops import native as nv
        dev = D.device
        dp = nv.padded(d)
        self.gids = np.asarray(gids)
        n = len(gids)
        self.G = torch.empty((n, d, d), dtype=torch.float64, device=dev)
        self.b = torch.empty((n, d), dtype=torch.float64, device=dev)
        self.hyy = torch.empty((n,), dtype=torch.float64, device=dev)
        for k0, G, b, hyy in SC.worker_grams(gids, d, s, dp, seed, dev=dev):
            nb = G.shape[0]
            self.G[k0:k0 + nb] = G[:, :d, :d]
            self.b[k0:k0 + nb] = b[:, :d]
            self.hyy[k0:k0 + nb] = hyy

    def grad_obj(self, theta):
        """g_n = G_n theta_n - b_n and f_n = 0.5 th'G th - b'th + 0.5 y'y; theta (n, d) or (d,)."""
        th = theta if theta.dim() == 2 else theta.unsqueeze(0).expand(self.G.shape[0], -1)
        Gt = torch.bmm(self.G, th.unsqueeze(2)).squeeze(2)
        f = 0.5 * (th * Gt).sum(1) - (self.b * th).sum(1) + self.hyy
        return Gt - self.b, f


def _allsum(t, world):
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def run_shared(W: DenseWorkers, kind, N, d, alpha, Hmax, T, obj0, D):
    """GD / LAG-PS / LAG-WK / cyclic IAG with a replicated server (all_reduce of the column sums)."""
    dev, world = D.device, D.world
    nl = W.G.shape[0]
    g0 = int(W.gids[0])
    theta = torch.zeros((d,), dtype=torch.float64, device=dev)
    ghat = torch.ones((nl, d), dtype=torch.float64, device=dev)
    that = torch.zeros((nl, d), dtype=torch.float64, device=dev)
    hm = torch.as_tensor(Hmax[g0:g0 + nl], device=dev)
    dn = []
    thr_ps = 10.0 / (alpha ** 2 * N ** 2) / 10
    thr_wk = thr_ps / 10
    step = alpha / N if kind == "iag" else alpha
    loss, tv, comm, up_tot = [], [], [], 1
    gids = torch.as_tensor(W.gids, device=dev)
    for it in range(1, T + 1):
        g, f = W.grad_obj(theta)
        if kind == "gd":
            S = g.sum(0) if it > 1 else torch.ones_like(theta) * (1.0 if g0 == 0 else 0.0)
            ups = nl
        else:
            if kind in ("lagps", "lagwk"):
                if kind == "lagps" and it > 1 and g0 == 0:
                    ghat[0] = g[0]
                ups = 0
                if it > 10:
                    trig = float(sum(dn[-10:]))
                    if kind == "lagps":
                        v = ((that - theta) ** 2).sum(1) * hm * hm
                        up = v > thr_ps * trig
                    else:
                        up = ((g - ghat) ** 2).sum(1) > thr_wk * trig
                    ghat = torch.where(up.unsqueeze(1), g, ghat)
                    if kind == "lagps":
                        that = torch.where(up.unsqueeze(1), theta.unsqueeze(0).expand_as(that), that)
                    ups = int(up.sum().item())
            else:
                sel = it % N
                if it > 1:
                    ghat = torch.where((gids == sel).unsqueeze(1), g, ghat)
                ups = 1
            S = ghat.sum(0)
        pack = torch.cat([S, f.sum().reshape(1), torch.tensor([float(ups)], dtype=torch.float64, device=dev)])
        _allsum(pack, world)
        stepv = step * pack[:d]
        theta = theta - stepv
        dn.append(float((stepv * stepv).sum().item()))
        fs = float(pack[d].item())                       # host-side stop test every iteration
        up_tot += int(pack[d + 1].item())
        loss.append(abs(fs - obj0))
        comm.append(up_tot)
        tv.append(_sync_time(dev))
    return np.asarray(loss), np.asarray(tv) - tv[0], np.asarray(comm, dtype=np.float64)


def _halo(rows_send, peers, dev, d):
    """Exchange boundary rows with the chain neighbours: rows_send/peers are (left, right)."""
    recv = [torch.zeros((d,), dtype=torch.float64, device=dev) for _ in range(2)]
    ops = []
    for i in range(2):
        if peers[i] is not None:
            ops.append(dist.P2POp(dist.isend, rows_send[i].contiguous(), peers[i]))
            ops.append(dist.P2POp(dist.irecv, recv[i], peers[i]))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return recv


def run_chain_first_order(W: DenseWorkers, kind, N, d, step, T, obj0, D):
    """DGD (gradient averaging over {n, n+-1}) and dual averaging (Jacobi order) on the chain;
    boundary rows travel by NCCL send/recv."""
    dev, world, rank = D.device, D.world, D.rank
    nl = W.G.shape[0]
    g0 = int(W.gids[0])
    Theta = torch.zeros((nl, d), dtype=torch.float64, device=dev)
    Z = torch.zeros((nl, d), dtype=torch.float64, device=dev)
    peers = (rank - 1 if rank > 0 else None, rank + 1 if rank < world - 1 else None)
    gid = torch.arange(g0, g0 + nl, device=dev)
    hasL = (gid > 0).to(torch.float64).unsqueeze(1)
    hasR = (gid < N - 1).to(torch.float64).unsqueeze(1)
    loss, tv = [], []
    for it in range(1, T + 1):
        g, f = W.grad_obj(Theta)
        fs = _allsum(f.sum().reshape(1), world)
        src = g if kind == "dgd" else Z
        if kind == "dgd" and it == 1:
            src = torch.ones_like(g)
        left_in, right_in = _halo((src[0], src[-1]), peers, dev, d) if world > 1 else (None, None)
        L = torch.cat([(left_in if left_in is not None else torch.zeros_like(src[0])).unsqueeze(0), src[:-1]])
        R = torch.cat([src[1:], (right_in if right_in is not None else torch.zeros_like(src[0])).unsqueeze(0)])
        cnt = hasL + hasR
        if kind == "dgd":
            avg = (src + hasL * L + hasR * R) / (1.0 + cnt)
            Theta = Theta - step * avg
        else:
            Z = (hasL * L + hasR * R) / torch.clamp(cnt, min=1.0) + g
            Theta = -step * Z
        loss.append(abs(float(fs.item()) - obj0))
        tv.append(_sync_time(dev))
    return np.asarray(loss), np.asarray(tv) - tv[0]


def run_star(N, d, s, rho, T, obj0, D):
    """standared_ADMM.m with an all_reduce for the server's reduction, replicated server solve."""
    dev, world = D.device, D.world
    own = np.nonzero(partition(N - 1, world) == D.rank)[0]
    W = DenseWorkers(own, d, s, D)
    S = DenseWorkers([N - 1], d, s, D)
    eye = torch.eye(d, dtype=torch.float64, device=dev)
    M = torch.empty_like(W.G)
    for k0 in range(0, len(own), 16):
        M[k0:k0 + 16] = torch.cholesky_inverse(torch.linalg.cholesky(W.G[k0:k0 + 16] + rho * eye))
    Ms = torch.cholesky_inverse(torch.linalg.cholesky(S.G[0] + (N - 1) * rho * eye))
    lam = torch.zeros_like(W.b)
    z = torch.zeros((d,), dtype=torch.float64, device=dev)
    loss, tv = [], []
    for it in range(1, T + 1):
        r = W.b - lam + rho * z
        th = torch.bmm(M, r.unsqueeze(2)).squeeze(2)
        f = 0.5 * (th * (r - rho * th)).sum() - (W.b * th).sum() + W.hyy.sum()
        pack = torch.cat([(lam + rho * th).sum(0), f.reshape(1)])
        _allsum(pack, world)
        rs = S.b[0] + pack[:d]
        z = Ms @ rs
        lam = lam + rho * (th - z)
        fs = float(pack[d].item()) + float((0.5 * z @ (rs - (N - 1) * rho * z) - S.b[0] @ z).item()) \
            + float(S.hyy[0].item())
        loss.append(abs(fs - obj0))
        tv.append(_sync_time(dev))
        if loss[-1] < 1e-4:
            break
    return np.asarray(loss), np.asarray(tv) - tv[0], len(own)


def run_arms(a, D, obj0, Gsum, algos, record, hbm):
    """All requested arms, one after the other (each frees its dense matrices before the next)."""
    import gc
    from ..parallel.nccl_baseline import NcclChainGadmm
    from ..engine import scaled as SC
    from ..ops import native as nv
    N, d, s = a.workers, a.d, a.s
    dev, world = D.device, D.world
    dp = nv.padded(d)
    dense = d * d * 8
    own = np.nonzero(partition(N, world) == D.rank)[0]

    def tmax(tv):
        t = torch.as_tensor(np.asarray(tv), device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().numpy()

    for name in ("gadmm", "dgadmm"):
        if name not in algos:
            continue
        eng = NcclChainGadmm(N, d, a.rho, obj0, 1e-4, a.cap + 8, dist=D, dynamic=(name == "dgadmm"),
                             stop=False)
        for k0, G, b, hyy in SC.worker_grams(eng.gids, d, s, dp, dev=dev):
            eng.ingest_gram(G, b, hyy, k0)
        eng.set_path(np.arange(N))
        if name == "dgadmm":
            from ..topology import findPath2
            g = np.random.default_rng(11)

            def nxt(i):
                out = findPath2(N, g) if D.rank == 0 else None
                return np.asarray(D.broadcast_object(out)[0]) - 1
            it = eng.run(a.cap, rechain=nxt, coherence=a.tau)
        else:
            it = eng.run(a.cap)
        obj, loss, it, tv = eng.results(it)
        record(D, a.out, "GADMM (static chain)" if name == "gadmm" else f"D-GADMM tau={a.tau}", "nccl",
               loss, tv, len(own) * dense, hbm, {"rho": a.rho})
        del eng
        gc.collect()
        torch.cuda.empty_cache()
    if "star" in algos:
        loss, tv, nw = run_star(N, d, s, a.star_rho, a.cap, obj0, D)
        record(D, a.out, "standard (star) ADMM", "nccl", loss, tmax(tv), (nw + 1) * dense, hbm,
               {"rho": a.star_rho})
        gc.collect()
        torch.cuda.empty_cache()
    fo = [x for x in ("gd", "lagps", "lagwk", "iag", "dgd", "dualavg_jacobi") if x in algos]
    if fo:
        W = DenseWorkers(own, d, s, D)
        lmax = float(torch.linalg.eigvalsh(Gsum[:d, :d])[-1].item())
        alpha = 1.0 / lmax
        hm = torch.zeros((N,), dtype=torch.float64, device=dev)
        v = torch.ones((len(own), d, 1), dtype=torch.float64, device=dev)
        for _ in range(40):                                  # batched power iteration (set-up)
            w = torch.bmm(W.G, v)
            v = w / w.norm(dim=1, keepdim=True)
        hm[int(own[0]):int(own[0]) + len(own)] = (v * torch.bmm(W.G, v)).sum((1, 2))
        _allsum(hm, world)
        Hmax = hm.cpu().numpy()
        names = {"gd": "GD", "lagps": "LAG-PS", "lagwk": "LAG-WK", "iag": "cyclic IAG", "dgd": "DGD",
                 "dualavg_jacobi": "dual averaging (Jacobi order)"}
        for key in fo:
            extra = {"step": alpha}
            if key in ("dgd", "dualavg_jacobi"):
                loss, tv = run_chain_first_order(W, key, N, d, alpha / 100 if key == "dgd" else alpha,
                                                 a.fo_iters, obj0, D)
            else:
                loss, tv, comm = run_shared(W, key, N, d, alpha, Hmax, a.fo_iters, obj0, D)
                if key.startswith("lag"):
                    extra["uploads"] = float(comm[-1])
            record(D, a.out, names[key], "nccl", loss, tmax(tv), len(own) * dense, hbm, extra)
