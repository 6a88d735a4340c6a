"""``backend="nccl"`` / ``"gloo"`` first-order baselines: GD, DGD, LAG-PS, LAG-WK, IAG and dual
averaging with the workers split over the ranks and nothing but library ops + collectives.

This is the *baseline* arm of BASELINE.json ("a plain torch.distributed ... implementation of the
same algorithms"): every rank keeps the shards of its contiguous block of workers, computes their
gradients with torch (fp64, on its GPU for nccl / on the CPU for gloo) and the ranks assemble the
``N x d`` gradient matrix with ``all_gather`` (parameter-server uploads and chain halos alike);
objectives and the global Gram are ``all_reduce`` sums.  The algorithm loops themselves are the
oracle's (:mod:`gadmm_b200.oracle.baselines`, :mod:`gadmm_b200.oracle.dual_avg`), which are written
against a gradient/objective provider -- here :class:`DistFamily` -- so every curve, trigger
decision and communication counter is the reference's (GD_DGD_LAG.m, dual_averaging.m).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from ..models.problems import _as_f64
from ..oracle import baselines as OB
from ..oracle import dual_avg as OD
from .nccl_baseline import _Ctx
from .runtime import partition


class DistFamily:
    """Drop-in for ``oracle.baselines._Family`` over a process group."""

    def __init__(self, XX, YY, N, d, s, logistic: bool, lam: float = 0.0, naive: bool = True,
                 backend: str = "gloo"):
        self.ctx = _Ctx(backend)
        self.N, self.d, self.s, self.logistic = int(N), int(d), int(s), bool(logistic)
        self.lam, self.naive = float(lam), bool(naive)
        world = self.ctx.world
        owner = partition(self.N, max(world, 1))
        mine = np.nonzero(owner == (self.ctx.rank if world > 1 else 0))[0]
        self.owner = owner
        self.lo = int(mine[0]) if mine.size else 0
        self.nloc = int(mine.size)
        self.counts = [int((owner == r).sum()) for r in range(max(world, 1))]
        self.maxloc = max(self.counts)
        dev = self.ctx.device
        XXa = _as_f64(XX)[: self.N * self.s, : self.d].reshape(self.N, self.s, self.d)
        YYa = _as_f64(YY).reshape(-1)[: self.N * self.s].reshape(self.N, self.s)
        self.X = torch.from_numpy(np.ascontiguousarray(XXa[self.lo:self.lo + self.nloc])).to(dev)
        self.y = torch.from_numpy(np.ascontiguousarray(YYa[self.lo:self.lo + self.nloc])).to(dev)
        self.G = torch.einsum("nsi,nsj->nij", self.X, self.X)
        self.b = torch.einsum("nsi,ns->ni", self.X, self.y)
        self.Gsum = self._sum(self.G.sum(0)).cpu().numpy()          # global Gram (Hmaxall)
        self.bsum = self._sum(self.b.sum(0)).cpu().numpy()

    # ------------------------------------------------------------------ collectives
    def _sum(self, t: torch.Tensor) -> torch.Tensor:
        t = t.contiguous()
        if self.ctx.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def _gather_rows(self, local: torch.Tensor) -> np.ndarray:
        """(nloc, d) on every rank -> (N, d) everywhere (the 'upload' of every worker's vector)."""
        if self.ctx.world == 1:
            return local.cpu().numpy()
        pad = torch.zeros((self.maxloc, self.d), dtype=torch.float64, device=self.ctx.device)
        pad[: self.nloc] = local
        parts = [torch.empty_like(pad) for _ in range(self.ctx.world)]
        dist.all_gather(parts, pad)
        return torch.cat([p[:c] for p, c in zip(parts, self.counts)], 0).cpu().numpy()

    def _dev(self, a) -> torch.Tensor:
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.ctx.device)

    # ------------------------------------------------------------------ provider interface
    def _w(self, z: torch.Tensor) -> torch.Tensor:
        if self.naive:
            return self.y / (1.0 + torch.exp(z))
        return self.y * 0.5 * (1.0 - torch.tanh(0.5 * z))

    def _grads_local(self, Th_loc: torch.Tensor) -> torch.Tensor:
        """Gradient of each local worker at its own row of ``Th_loc`` (nloc, d)."""
        if not self.logistic:
            return torch.einsum("nij,nj->ni", self.G, Th_loc) - self.b
        z = self.y * torch.einsum("nsd,nd->ns", self.X, Th_loc)
        return -torch.einsum("nsd,ns->nd", self.X, self._w(z)) + self.lam * Th_loc

    def grads_shared(self, th):
        t = self._dev(th)
        return self._gather_rows(self._grads_local(t.unsqueeze(0).expand(self.nloc, self.d)))

    def grads_own(self, Th):
        T = self._dev(np.asarray(Th).T[self.lo:self.lo + self.nloc])
        return self._gather_rows(self._grads_local(T))

    def grad_one(self, n, th):
        return self.grads_shared(th)[int(n)]

    def full_grad(self, th):
        t = self._dev(th)
        g = self._grads_local(t.unsqueeze(0).expand(self.nloc, self.d)).sum(0)
        return self._sum(g).cpu().numpy()

    def _loss_rows(self, Th_loc: torch.Tensor) -> torch.Tensor:
        z = -self.y * torch.einsum("nsd,nd->ns", self.X, Th_loc)
        return torch.log(1.0 + torch.exp(z)) if self.naive else torch.nn.functional.softplus(z)

    def obj_shared(self, th):
        t = self._dev(th)
        Tl = t.unsqueeze(0).expand(self.nloc, self.d)
        if not self.logistic:
            r = torch.einsum("nsd,nd->ns", self.X, Tl) - self.y
            return 0.5 * float(self._sum((r * r).sum().reshape(1)).item())
        part = float(self._sum(self._loss_rows(Tl).sum().reshape(1)).item())
        return 0.5 * self.N * self.lam * float(np.dot(th, th)) + part

    def obj_own(self, Th):
        T = self._dev(np.asarray(Th).T[self.lo:self.lo + self.nloc])
        if not self.logistic:
            r = torch.einsum("nsd,nd->ns", self.X, T) - self.y
            return 0.5 * float(self._sum((r * r).sum().reshape(1)).item())
        part = self._loss_rows(T).sum() + 0.5 * self.lam * (T * T).sum()
        return float(self._sum(part.reshape(1)).item())


def gd_dgd_lag(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede, y_fede, num_sample,
               rng=None, backend="gloo"):
    """``GD_DGD_LAG`` (19 outputs, GD_DGD_LAG.m:1-4) over the process group."""
    fam = DistFamily(X_fede, y_fede, num_workers, num_feature, num_sample, False, backend=backend)
    Hmax = _as_f64(Hmax).reshape(-1)[:num_workers]
    alpha = 1.0 / float(np.linalg.eigvalsh(fam.Gsum)[-1])
    thrd = 10.0 / (alpha ** 2 * num_workers ** 2) / OB.TRIGGERSLOT
    thrd5 = 1.0 / (alpha ** 2 * num_workers ** 2) / OB.TRIGGERSLOT
    nonprob = Hmax / Hmax.sum()
    g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng
    obj_GD, time_GD, _ = OB._gd(fam, alpha, num_iter, logistic=False)
    loss_GD = np.abs(obj_GD - obj0)
    obj_DGD, loss_DGD, time_DGD = OB._dgd(fam, alpha / 100.0, num_iter, obj0)
    o_ps, l_ps, t_ps, c_ps = OB._lag_like(fam, "lag_ps", alpha, thrd, Hmax, num_iter, obj0, None)
    o_wk, l_wk, t_wk, c_wk = OB._lag_like(fam, "lag_wk", alpha, thrd5, Hmax, num_iter, obj0, None)
    o_c, l_c = OB._lag_like(fam, "iag_cyclic", alpha, 0.0, Hmax, num_iter, obj0, None)
    o_r, l_r = OB._lag_like(fam, "iag_random", alpha, 0.0, Hmax, num_iter, obj0, None, g, nonprob)
    return (obj_GD, loss_GD, time_GD, obj_DGD, loss_DGD, time_DGD,
            o_ps, l_ps, t_ps, c_ps, o_wk, l_wk, t_wk, c_wk, o_c, l_c, o_r, l_r, alpha)


def gd_dgd_lag_logistic(X, y, num_workers, num_feature, Hmax, num_iter, X_fede, y_fede, num_sample,
                        lam, accuracy, rng=None, naive=True, backend="gloo"):
    """``GD_DGD_LAG_logistic`` (20 outputs, GD_DGD_LAG_logistic.m:1-5) over the process group."""
    fam = DistFamily(X_fede, y_fede, num_workers, num_feature, num_sample, True, lam, naive, backend)
    Hmax = _as_f64(Hmax).reshape(-1)[:num_workers]
    alpha = 1.0 / (0.25 * float(np.linalg.eigvalsh(fam.Gsum)[-1]) + lam)
    thrd = 10.0 / (alpha ** 2 * num_workers ** 2) / OB.TRIGGERSLOT
    thrd5 = 1.0 / (alpha ** 2 * num_workers ** 2) / OB.TRIGGERSLOT
    nonprob = Hmax / Hmax.sum()
    g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng
    obj_GD, time_GD, _ = OB._gd(fam, alpha, num_iter, logistic=True)
    obj1 = float(obj_GD[-1])
    loss_GD = np.abs(obj_GD - obj1)
    obj_DGD, loss_DGD, time_DGD = OB._dgd(fam, alpha / 100.0, num_iter, obj1)
    o_ps, l_ps, t_ps, c_ps = OB._lag_like(fam, "lag_ps", alpha, thrd, Hmax, num_iter, obj1, accuracy)
    o_wk, l_wk, t_wk, c_wk = OB._lag_like(fam, "lag_wk", alpha, thrd5, Hmax, num_iter, obj1, accuracy)
    o_c, l_c = OB._lag_like(fam, "iag_cyclic", alpha, 0.0, Hmax, num_iter, obj1, accuracy)
    o_r, l_r = OB._lag_like(fam, "iag_random", alpha, 0.0, Hmax, num_iter, obj1, accuracy, g, nonprob)
    return (obj_GD, loss_GD, time_GD, obj_DGD, loss_DGD, time_DGD,
            o_ps, l_ps, t_ps, c_ps, o_wk, l_wk, t_wk, c_wk, o_c, l_c, o_r, l_r, alpha, obj1)


def _dual_avg(fam: DistFamily, num_iter, obj0, acc, alpha, mode, return_state):
    res, st = OD._dual_avg(fam.N, fam.d, None, fam.obj_own, num_iter, obj0, acc, alpha, mode,
                           grads_all=fam.grads_own)
    return res + (st,) if return_state else res


def dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc, alpha,
                   mode="faithful", return_state=False, backend="gloo"):
    """``dual_averaging`` (dual_averaging.m:1) over the process group."""
    fam = DistFamily(XX, YY, no_workers, num_feature, noSamples, False, backend=backend)
    return _dual_avg(fam, num_iter, obj0, acc, alpha, mode, return_state)


def dual_averaging_logistic(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc, alpha,
                            eta, mode="faithful", naive=True, return_state=False, backend="gloo"):
    """``dual_averaging_logisticReg`` (dual_averaging_logisticReg.m:1) over the process group."""
    fam = DistFamily(XX, YY, no_workers, num_feature, noSamples, True, eta, naive, backend)
    return _dual_avg(fam, num_iter, obj0, acc, alpha, mode, return_state)
