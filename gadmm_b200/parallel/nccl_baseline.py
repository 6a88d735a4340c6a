"""Baseline: chain GADMM / D-GADMM / star ADMM with plain ``torch.distributed`` collectives.

This is the implementation the product has to beat: dense fp64 inverses,
``torch.bmm`` (cuBLAS) for the local solves, ``batch_isend_irecv`` (NCCL send/recv) for the
neighbour exchange, ``all_reduce`` for the objective and the star reduction, a host-side stop
test every iteration.  The same code runs on CPU tensors over gloo -- that is the "plumbing"
configuration of BASELINE.json (world_size=2, no GPU).

None of the hand-written kernels are used here.
"""
from __future__ import annotations

import time

import numpy as np
import torch
import torch.distributed as dist

from ..topology import chain_tables
from .runtime import env_rank_world, init_process_group, partition


class _Ctx:
    def __init__(self, backend: str):
        self.rank, self.world, self.local = env_rank_world()
        if self.world > 1 and not dist.is_initialized():
            init_process_group(backend)
        self.cuda = backend == "nccl"
        self.device = torch.device("cuda", self.local if self.world > 1 else
                                   torch.cuda.current_device()) if self.cuda \
            else torch.device("cpu")

    def sync(self):
        if self.cuda:
            torch.cuda.synchronize(self.device)


class NcclChainGadmm:
    """Node-dual chain GADMM with NCCL/gloo send/recv halos (one instance per rank)."""

    def __init__(self, N, d, rho, obj0, acc, max_iter, dist=None, backend: str = "nccl",
                 dynamic: bool = False, stop: bool = True):
        self.ctx = _Ctx(backend)
        self.dist = dist
        self.dev = self.ctx.device
        self.N, self.d, self.rho = int(N), int(d), float(rho)
        self.obj0, self.acc, self.max_iter = float(obj0), float(acc), int(max_iter)
        self.stop, self.dynamic = stop, dynamic
        self.rank, self.world = self.ctx.rank, self.ctx.world
        self.owner = partition(self.N, self.world)
        self.gids = np.nonzero(self.owner == self.rank)[0]
        self.Nloc = len(self.gids)
        self.g2l = {int(g): k for k, g in enumerate(self.gids)}
        f64 = torch.float64
        # static chains: one dense inverse per worker (chain ends get c=1 in place);
        # dynamic chains keep c=1 and c=2 for every worker
        self.M = {2: torch.empty((self.Nloc, d, d), dtype=f64, device=self.dev)}
        if dynamic:
            self.M[1] = torch.empty((self.Nloc, d, d), dtype=f64, device=self.dev)
        self._M1_end = {}
        self.b = torch.zeros((self.Nloc, d), dtype=f64, device=self.dev)
        self.hyy = torch.zeros((self.Nloc,), dtype=f64, device=self.dev)
        self.G = None
        self.nu = torch.zeros((self.Nloc, d), dtype=f64, device=self.dev)
        self.theta = torch.zeros((self.N, d), dtype=f64, device=self.dev)   # replica (own + halo rows)
        self.obj_hist, self.loss_hist, self.t_hist = [], [], []
        self.it = 0
        self.hit = 0
        self.launches = 0

    # ---------------------------------------------------------------------------- setup
    def ingest_gram(self, G, b, hyy, k0):
        nb, d = G.shape[0], self.d
        G = G[:, :d, :d].to(self.dev)
        self.b[k0:k0 + nb] = b[:, :d].to(self.dev)
        self.hyy[k0:k0 + nb] = hyy.to(self.dev)
        eye = torch.eye(d, dtype=torch.float64, device=self.dev)
        for c in ((1, 2) if self.dynamic else (2,)):
            L = torch.linalg.cholesky(G + c * self.rho * eye)
            self.M[c][k0:k0 + nb] = torch.cholesky_inverse(L)
        if not self.dynamic:
            for g_end in (0, self.N - 1):
                k = self.g2l.get(g_end)
                if k is not None and k0 <= k < k0 + nb:
                    L = torch.linalg.cholesky(G[k - k0] + self.rho * eye)
                    self._M1_end[k] = torch.cholesky_inverse(L)

    def ingest_host(self, XX, YY, s, batch: int = 64):
        XX = np.asarray(XX, dtype=np.float64)
        YY = np.asarray(YY, dtype=np.float64).reshape(-1)
        d = self.d
        for k0 in range(0, self.Nloc, batch):
            k1 = min(self.Nloc, k0 + batch)
            g0, g1 = int(self.gids[k0]), int(self.gids[k1 - 1]) + 1
            X = torch.from_numpy(XX[g0 * s:g1 * s, :d].reshape(g1 - g0, s, d)).to(self.dev)
            y = torch.from_numpy(YY[g0 * s:g1 * s].reshape(g1 - g0, s)).to(self.dev)
            G = X.transpose(1, 2) @ X
            b = torch.einsum("nsd,ns->nd", X, y)
            self.ingest_gram(G, b, 0.5 * (y * y).sum(1), k0)

    def set_path(self, path0):
        path0 = np.asarray(path0, dtype=np.int64)
        prev = getattr(self, "path0", None)
        self.path0 = path0
        self.left, self.right, self.color, self.cmul = chain_tables(path0)
        dev = self.dev
        g = self.gids
        self.loc = {}
        for col in (0, 1):
            ks = [k for k, gid in enumerate(g) if self.color[gid] == col]
            if not ks:
                self.loc[col] = None
                continue
            kt = torch.tensor(ks, device=dev)
            gl = g[ks]
            le, ri = self.left[gl], self.right[gl]
            if self.dynamic:
                Msel = torch.stack([self.M[int(self.cmul[gid])][k] for k, gid in zip(ks, gl)])
            else:
                for k, gid in zip(ks, gl):
                    if self.cmul[gid] == 1 and k in self._M1_end:
                        self.M[2][k] = self._M1_end.pop(k)       # c is fixed per worker here
                step = ks[1] - ks[0] if len(ks) > 1 else 1
                regular = all(b2 - a2 == step for a2, b2 in zip(ks[:-1], ks[1:]))
                Msel = self.M[2][ks[0]:ks[-1] + 1:step] if regular else self.M[2][kt]
            self.loc[col] = dict(
                k=kt, gid=torch.tensor(gl, device=dev), M=Msel,
                l=torch.tensor(np.maximum(le, 0), device=dev),
                r=torch.tensor(np.maximum(ri, 0), device=dev),
                hl=torch.tensor((le >= 0).astype(np.float64), device=dev).unsqueeze(1),
                hr=torch.tensor((ri >= 0).astype(np.float64), device=dev).unsqueeze(1),
                c=torch.tensor(self.cmul[gl].astype(np.float64), device=dev))
        # halo plan: after a colour's update, send own rows whose neighbour lives elsewhere
        self.plan = {0: [], 1: []}
        for gid in range(self.N):
            for nb in (self.left[gid], self.right[gid]):
                if nb < 0 or self.owner[nb] == self.owner[gid]:
                    continue
                col = int(self.color[gid])
                if self.owner[gid] == self.rank:
                    self.plan[col].append(("send", int(gid), int(self.owner[nb])))
                elif self.owner[nb] == self.rank:
                    self.plan[col].append(("recv", int(gid), int(self.owner[gid])))
        for col in (0, 1):   # de-duplicate (a worker with both neighbours on one peer)
            self.plan[col] = sorted(set(self.plan[col]), key=lambda t: (t[1], t[0], t[2]))
        allk = torch.arange(self.Nloc, device=dev)
        gl = g
        le, ri = self.left[gl], self.right[gl]
        self.dual = dict(gid=torch.tensor(gl, device=dev),
                         l=torch.tensor(np.maximum(le, 0), device=dev),
                         r=torch.tensor(np.maximum(ri, 0), device=dev),
                         hl=torch.tensor((le >= 0).astype(np.float64), device=dev).unsqueeze(1),
                         hr=torch.tensor((ri >= 0).astype(np.float64), device=dev).unsqueeze(1))
        if prev is not None and self.world > 1:
            self._refresh_replica()

    def _refresh_replica(self):
        """After a re-chain every rank needs its new neighbours' rows: all_gather the table."""
        own = torch.zeros_like(self.theta)
        idx = torch.tensor(self.gids, device=self.dev)
        own[idx] = self.theta[idx]
        dist.all_reduce(own, op=dist.ReduceOp.SUM)
        self.theta.copy_(own)

    # ------------------------------------------------------------------------------ run
    def _exchange(self, col):
        ops = []
        for kind, gid, peer in self.plan[col]:
            row = self.theta[gid]
            ops.append(dist.P2POp(dist.isend if kind == "send" else dist.irecv, row, peer))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            self.launches += len(ops)

    def _half(self, col):
        L = self.loc[col]
        if L is not None:
            th = self.theta
            r = self.b[L["k"]] - self.nu[L["k"]] + self.rho * (L["hl"] * th[L["l"]] +
                                                                L["hr"] * th[L["r"]])
            new = torch.bmm(L["M"], r.unsqueeze(2)).squeeze(2)
            th[L["gid"]] = new
            # objective piece: 0.5 th'(r - c rho th) - b'th + 0.5 y'y
            gth = r - (self.rho * L["c"]).unsqueeze(1) * new
            self._objw[L["k"]] = 0.5 * (new * gth).sum(1) - (self.b[L["k"]] * new).sum(1) \
                + self.hyy[L["k"]]
            self.launches += 8
        if self.world > 1:
            self._exchange(col)

    def iterate(self, n_iter: int, use_graph: bool = False):
        if not hasattr(self, "_objw"):
            self._objw = torch.zeros((self.Nloc,), dtype=torch.float64, device=self.dev)
        for _ in range(int(n_iter)):
            if self.stop and self.hit:
                return
            self.it += 1
            self._half(0)
            self._half(1)
            D = self.dual
            th = self.theta
            me = th[D["gid"]]
            self.nu += self.rho * D["hr"] * (me - th[D["r"]]) - self.rho * D["hl"] * (th[D["l"]] - me)
            tot = self._objw.sum().reshape(1)
            if self.world > 1:
                dist.all_reduce(tot, op=dist.ReduceOp.SUM)
            self.launches += 6
            f = float(tot.item())            # host-side stop test every iteration
            self.obj_hist.append(f)
            self.loss_hist.append(abs(f - self.obj0))
            self.t_hist.append(time.perf_counter())
            if self.loss_hist[-1] < self.acc and not self.hit:
                self.hit = self.it

    def run(self, num_iter=None, window: int = 32, rechain=None, coherence=None, **_):
        num_iter = self.max_iter if num_iter is None else min(int(num_iter), self.max_iter)
        while self.it < num_iter:
            nxt = self.it + 1
            if rechain is not None and coherence is not None and nxt > 1 and nxt % coherence == 0:
                self.set_path(rechain(nxt))
            self.iterate(1)
            if self.stop and self.hit:
                return self.hit
        return num_iter

    def results(self, Iter):
        obj = np.asarray(self.obj_hist[:Iter])
        loss = np.asarray(self.loss_hist[:Iter])
        t = np.asarray(self.t_hist[:Iter])
        tv = t - t[0] if Iter > 0 else t
        if self.world > 1:
            tt = torch.from_numpy(tv.copy()).to(self.dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            tv = tt.cpu().numpy()
        return obj, loss, Iter, tv


def build_scaled_linreg(eng: NcclChainGadmm, s: int, seed: int = 0, noise: float = 0.1,
                        verbose: bool = False, batch: int = 16):
    """Same synthetic problem as :func:`gadmm_b200.engine.scaled.build_scaled_linreg` (same
    generator seeds per worker), built with library ops only."""
    dev, d = eng.dev, eng.d
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)
    theta_true = torch.randn(d, dtype=torch.float64, device=dev, generator=gen)
    Gsum = torch.zeros((d, d), dtype=torch.float64, device=dev)
    bsum = torch.zeros((d,), dtype=torch.float64, device=dev)
    hsum = torch.zeros((), dtype=torch.float64, device=dev)
    scale = 1.0 / np.sqrt(s)
    for k0 in range(0, eng.Nloc, batch):
        nb = min(batch, eng.Nloc - k0)
        gen.manual_seed(seed * 1000003 + int(eng.gids[k0]))
        X = torch.randn((nb, s, d), dtype=torch.float32, device=dev, generator=gen).double()
        X *= scale
        y = torch.einsum("nsd,d->ns", X, theta_true)
        y += noise * torch.randn((nb, s), dtype=torch.float64, device=dev, generator=gen)
        G = X.transpose(1, 2) @ X
        b = torch.einsum("nsd,ns->nd", X, y)
        hyy = 0.5 * (y * y).sum(1)
        Gsum += G.sum(0)
        bsum += b.sum(0)
        hsum += hyy.sum()
        eng.ingest_gram(G, b, hyy, k0)
    if eng.world > 1:
        for t in (Gsum, bsum, hsum):
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
    th = torch.linalg.solve(Gsum, bsum)
    obj0 = 0.5 * float(th @ (Gsum @ th)) - float(bsum @ th) + float(hsum)
    eng.obj0 = obj0
    return obj0
