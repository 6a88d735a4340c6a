"""Logistic GADMM with exact local solves at scale, single- and multi-GPU (K9, BASELINE config 3).

``group_ADMM_logistic.m:48-55`` hands every worker's strongly convex sub-problem to CVX.  At the
reference's 50x50 shards an explicit Hessian is the right tool (:func:`gadmm_logistic_newton`);
at 512 workers x d = 8192 with s < d it is not -- forming ``X'DX`` costs ``2 s d^2`` flop per
Newton step and worker, a Hessian *product* costs ``4 s d`` and one read of the shard.  Here the
local solve is a batched Newton-CG: all workers of one colour advance in lock-step, and each
gradient evaluation or Hessian product is ONE pass over the fp32 shards by
``logistic_xtwx_kernel`` (csrc/logistic_scaled.cu), accumulated in fp64.

Multi-GPU: contiguous chain segments per rank (as in :class:`ChainEngine`).  The model table lives
in symmetric memory; after each colour's update a rank stores its boundary model straight into
the neighbouring rank's table over NVLink (a peer view of the symmetric buffer) and the ranks
meet at a barrier -- the local solve is hundreds of kernel launches long, so unlike the
closed-form chain there is nothing to gain from in-kernel flags here.  Edge duals are kept by the
left end-point's rank and recomputed redundantly for the one edge that crosses into a rank.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Optional, Tuple

import numpy as np
import torch

from ..ops import native as nv
from ..ops import topology as TP
from ..parallel.runtime import Dist, partition

_SM_COUNT = 148


def _pick_ctas_per_shard(n_shards: int, s: int, rb: int = 2) -> int:
    """CTAs per shard so that ``n_shards * G`` fills whole waves of one CTA per SM."""
    best, best_cost = 1, None
    for g in range(1, 17):
        if g * rb * 4 > s and g > 1:
            break
        waves = -(-n_shards * g // _SM_COUNT)
        # every CTA writes (and the reduction re-reads) a d-vector of partial sums: 16 g d bytes per
        # shard against 4 s d bytes of shard data
        cost = waves / g * (1.0 + 4.0 * g / max(s, 1))
        if best_cost is None or cost < best_cost * (1 - 1e-9):
            best, best_cost = g, cost
    return best


def newton_cg(evaluate: Callable, hessvec: Callable, x0: torch.Tensor, tol: float = 1e-9,
              max_newton: int = 40, max_cg: int = 250, check_every: int = 4, stats: dict = None):
    """Batched damped Newton-CG on rows of ``x0`` (B, d).

    ``evaluate(x) -> (f (B,), g (B, d))`` must also leave the curvature state used by
    ``hessvec(p) -> H(x) p`` at the evaluated point.  Rows stop when
    ``|g| <= tol * max(1, |x|)``; the inner CG uses the forcing term ``min(0.1, sqrt|g|)`` and an
    Armijo back-tracking line search keeps every row monotone.  Rows are masked, never
    compacted, so every product is one fixed-shape launch."""
    x = x0.clone()
    f, g = evaluate(x)
    n_eval, n_hv = 1, 0
    for _ in range(max_newton):
        gn = g.norm(dim=1)
        live = gn > tol * torch.clamp(x.norm(dim=1), min=1.0)
        if not bool(live.any().item()):
            break
        livef = live.to(x.dtype)
        # ---- CG on H p = g, zero start, per-row scalars
        eta = torch.clamp(torch.sqrt(gn), max=0.1) * gn
        p = torch.zeros_like(x)
        r = g * livef.unsqueeze(1)
        q = r.clone()
        rs = (r * r).sum(1)
        for it in range(max_cg):
            hq = hessvec(q)
            n_hv += 1
            qhq = (q * hq).sum(1)
            alpha = torch.where(qhq > 0, rs / torch.clamp(qhq, min=1e-300), torch.zeros_like(rs))
            p += alpha.unsqueeze(1) * q
            r -= alpha.unsqueeze(1) * hq
            rs_new = (r * r).sum(1)
            beta = torch.where(rs > 0, rs_new / torch.clamp(rs, min=1e-300), torch.zeros_like(rs))
            q = r + beta.unsqueeze(1) * q
            rs = rs_new
            if (it + 1) % check_every == 0 and not bool((torch.sqrt(rs) > eta).any().item()):
                break
        # ---- back-tracking from the full step; the last evaluation is at the accepted points
        slope = (g * p).sum(1)
        t = livef.clone()
        for _ls in range(30):
            f_try, g_try = evaluate(x - t.unsqueeze(1) * p)
            n_eval += 1
            # a predicted decrease below the rounding noise of f cannot be tested: accept it
            noise = 64.0 * torch.finfo(x.dtype).eps * torch.clamp(f.abs(), min=1.0)
            bad = (f_try > f - 1e-4 * t * slope + noise) & live & (t > 1e-9)
            if not bool(bad.any().item()):
                break
            t = torch.where(bad, 0.5 * t, t)
        x = x - t.unsqueeze(1) * p
        f, g = f_try, g_try
    if stats is not None:
        stats["evals"] = stats.get("evals", 0) + n_eval
        stats["hessvecs"] = stats.get("hessvecs", 0) + n_hv
    return x, f, g


class ScaledLogisticGadmm:
    """GADMM for consensus logistic regression with Newton-CG local solves.

    ``data=None`` draws a synthetic problem on the owning GPU (Philox stream = global worker id,
    so the data set does not depend on the number of ranks): ``X_n = Z_n / sqrt(d)``,
    ``y = sign(X theta* + 0.5 noise)``.  ``data=(X, y)`` takes host arrays ``(N, s, d)`` /
    ``(N, s)`` (every rank passes the full set and keeps its slice)."""

    def __init__(self, N: int, d: int, s: int, lam_log: float, rho: float,
                 dist: Optional[Dist] = None, seed: int = 0, data: Optional[Tuple] = None,
                 fast_hessvec: bool = True):
        self.dist = dist or Dist()
        # fp32-multiply Hessian products only where they are faster: for d > 5120 the kernel streams
        # two rows per block and keeps the converted block in registers, and the exact fp64 product
        # runs at the same 5.6 TB/s
        self.hv_mode = 2 if (fast_hessvec and int(d) * 40 <= 200 * 1024) else 1
        self.dev = self.dist.device
        self.N, self.d, self.s = int(N), int(d), int(s)
        self.lam_log, self.rho = float(lam_log), float(rho)
        if self.d % 4 or self.d > 8192:
            raise ValueError("ScaledLogisticGadmm needs d % 4 == 0 and d <= 8192")
        owner = partition(self.N, self.dist.world)
        mine = np.nonzero(owner == self.dist.rank)[0]
        if mine.size == 0:
            raise ValueError("more ranks than workers")
        self.owner = owner
        self.g0, self.Nloc = int(mine[0]), int(mine.size)
        f64, dev = torch.float64, self.dev
        # ---- shards
        self.X = torch.empty((self.Nloc, self.s, self.d), dtype=torch.float32, device=dev)
        self.y = torch.empty((self.Nloc, self.s), dtype=f64, device=dev)
        if data is None:
            th_true = TP.philox_normal((self.d,), seed * 7919 + 17, 0xFFFF, 0.0, 2.0, device=dev)
            for k in range(self.Nloc):
                gid = self.g0 + k
                TP.philox_normal(None, seed, gid, 0.0, 1.0 / np.sqrt(self.d), out=self.X[k])
                z = self.X[k].double() @ th_true
                z += TP.philox_normal((self.s,), seed + 1, gid, 0.0, 0.5, device=dev)
                self.y[k] = torch.where(z >= 0, 1.0, -1.0)
        else:
            Xh, yh = data
            Xh = np.asarray(Xh).reshape(self.N, self.s, self.d)[self.g0:self.g0 + self.Nloc]
            yh = np.asarray(yh).reshape(self.N, self.s)[self.g0:self.g0 + self.Nloc]
            self.X.copy_(torch.from_numpy(np.ascontiguousarray(Xh, dtype=np.float32)))
            self.y.copy_(torch.from_numpy(np.ascontiguousarray(yh, dtype=np.float64)))
        self.Dw = torch.zeros((self.Nloc, self.s), dtype=f64, device=dev)
        # ---- model table in symmetric memory: row g = theta of global worker g
        self._sym = self.dist.symm_bytes(self.N * self.d * 8)
        self.table = self._sym.local[: self.N * self.d * 8].view(f64).view(self.N, self.d)
        # lam[k] couples global workers (g0-1+k, g0+k), k = 0..Nloc  (edge -1 / N-1 do not exist)
        self.lam = torch.zeros((self.Nloc + 1, self.d), dtype=f64, device=dev)
        gids = torch.arange(self.g0, self.g0 + self.Nloc, device=dev)
        self.hasL = (gids > 0).to(f64).unsqueeze(1)
        self.hasR = (gids < self.N - 1).to(f64).unsqueeze(1)
        self.cvec = (self.hasL + self.hasR).squeeze(1)
        self.stats = {}
        self._last_loss = {}
        self._last_eval = {}       # colour -> (X'w, loss) of the last accepted evaluation
        self.launches = 0
        self._scratch = None
        self._scratch_key = None

    # ------------------------------------------------------------------ one pass over shards
    def _xtwx(self, x: torch.Tensor, first: int, count: int, stride: int, hv: int,
              shared_x: bool = False):
        """``v[k] = X' w(X x[k])`` for local shards ``first, first+stride, ...`` (``count``).
        ``hv``: 0 gradient weights (+ loss, curvature state), 1 Hessian product in fp64,
        2 Hessian product with fp32 multiplies (inner CG)."""
        G = _pick_ctas_per_shard(count, self.s)
        need = int(nv.lib().gadmm_logistic_xtwx_scratch_elems(count, self.d, G))
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty((need,), dtype=torch.float64, device=self.dev)
        v = torch.empty((count, self.d), dtype=torch.float64, device=self.dev)
        loss = torch.zeros((count,), dtype=torch.float64, device=self.dev)
        esz = self.s * self.d * 4
        nv.check(nv.lib().gadmm_logistic_xtwx(
            nv.ptr(self.X.data_ptr() + first * esz), nv.ptr(self.y.data_ptr() + first * self.s * 8),
            nv.ptr(x), C.c_longlong(0 if shared_x else self.d),
            nv.ptr(self.Dw.data_ptr() + first * self.s * 8), nv.ptr(v), self.d, nv.ptr(loss),
            count, self.s, self.d, G, int(hv), stride, nv.ptr(self._scratch),
            nv.stream_ptr()))
        self.launches += 2
        return v, loss

    # ------------------------------------------------------------------ consensus optimum
    def solve_optimum(self, tol: float = 1e-10):
        """``min_theta sum_n lam/2 |theta|^2 + loss_n(theta)`` by Newton-CG over all shards of all
        ranks (one shared vector, partial sums all-reduced).  Returns ``(obj0, theta)``."""
        mu = self.N * self.lam_log

        def evaluate(x):
            v, loss = self._xtwx(x.contiguous(), 0, self.Nloc, 1, 0, shared_x=True)
            gsum = torch.cat([v.sum(0), loss.sum(0, keepdim=True)])
            self.dist.all_reduce_sum_(gsum)
            f = gsum[-1:] + 0.5 * mu * (x * x).sum(1)
            return f, (gsum[:-1] + mu * x[0]).unsqueeze(0)

        def hessvec(p):
            v, _ = self._xtwx(p.contiguous(), 0, self.Nloc, 1, self.hv_mode, shared_x=True)
            hs = v.sum(0)
            self.dist.all_reduce_sum_(hs)
            return (hs + mu * p[0]).unsqueeze(0)

        x0 = torch.zeros((1, self.d), dtype=torch.float64, device=self.dev)
        x, f, _ = newton_cg(evaluate, hessvec, x0, tol=tol, max_cg=2000)
        self._last_eval.clear()            # the curvature state now belongs to the consensus point
        return float(f.item()), x[0]

    # ------------------------------------------------------------------ GADMM
    def objective(self, reuse: bool = False) -> float:
        """``sum_n lam/2 |theta_n|^2 + loss_n(theta_n)`` over all ranks.  ``reuse=True`` takes the
        data losses that the two half-iterations' last evaluations already produced (the last
        evaluation of :func:`newton_cg` is at the accepted point and a colour's models do not move
        during the other colour's half), which saves one pass over every shard per iteration."""
        th = self.table[self.g0:self.g0 + self.Nloc]
        if reuse and all(c in self._last_loss for c in (0, 1)):
            data = self._last_loss[0].sum() + self._last_loss[1].sum()
        else:
            _, loss = self._xtwx(th, 0, self.Nloc, 1, 0)      # same points: curvature state unchanged
            data = loss.sum()
        part = (data + 0.5 * self.lam_log * (th * th).sum()).reshape(1)
        self.dist.all_reduce_sum_(part)
        return float(part.item())

    def _push(self, colour: int):
        """Store this rank's boundary models of ``colour`` into the neighbouring ranks' tables."""
        if self.dist.world == 1:
            return
        hdl = self._sym.handle
        for gid, nb in ((self.g0, self.g0 - 1), (self.g0 + self.Nloc - 1, self.g0 + self.Nloc)):
            if gid % 2 != colour or nb < 0 or nb >= self.N:
                continue
            peer = int(self.owner[nb])
            view = hdl.get_buffer(peer, (self.N, self.d), torch.float64)
            view[gid].copy_(self.table[gid])

    def _half(self, colour: int, tol: float):
        first = (colour - self.g0) % 2                  # local index of the first worker of colour
        idx = torch.arange(first, self.Nloc, 2, device=self.dev)
        count = int(idx.numel())
        if count:
            g0, rho = self.g0, self.rho
            gl = idx + g0
            left = self.table[torch.clamp(gl - 1, min=0)] * self.hasL[idx]
            right = self.table[torch.clamp(gl + 1, max=self.N - 1)] * self.hasR[idx]
            lin = -self.lam[idx] * self.hasL[idx] + self.lam[idx + 1] * self.hasR[idx]
            lin -= rho * (left + right)
            mu = (self.lam_log + self.cvec[idx] * rho).unsqueeze(1)

            fresh = {"first": True}

            def evaluate(x):
                # The data terms at the starting point are those of this colour's last accepted
                # evaluation one iteration ago (its models have not moved since, and the curvature
                # state of its shards is still in place): the first evaluation costs no pass.
                if fresh["first"] and colour in self._last_eval:
                    v, loss = self._last_eval[colour]
                else:
                    v, loss = self._xtwx(x.contiguous(), first, count, 2, 0)
                    self._last_eval[colour] = (v, loss)
                fresh["first"] = False
                self._last_loss[colour] = loss
                f = loss + 0.5 * mu.squeeze(1) * (x * x).sum(1) + (lin * x).sum(1)
                return f, v + mu * x + lin

            def hessvec(p):
                v, _ = self._xtwx(p.contiguous(), first, count, 2, self.hv_mode)
                return v + mu * p

            x, _, _ = newton_cg(evaluate, hessvec, self.table[gl], tol=tol, stats=self.stats)
            self.table[gl] = x
        else:
            self._last_loss[colour] = torch.zeros((0,), dtype=torch.float64, device=self.dev)
        self._push(colour)
        self.dist.barrier()

    def run(self, num_iter: int, obj0: float, acc: float, newton_tol: float = 1e-9,
            faithful: bool = False, verbose: bool = False):
        """``(obj, loss, Iter, time)`` as ``group_ADMM_logistic`` returns them; time is the device
        clock at the end of every iteration (seconds since the first launch)."""
        obj, loss, tv = [], [], []
        Iter, hit = (100 if faithful else int(num_iter)), False
        self.dist.barrier()
        ev0 = torch.cuda.Event(enable_timing=True)
        ev0.record()
        lo = max(self.g0 - 1, 0)
        for i in range(1, int(num_iter) + 1):
            self._half(0, newton_tol)
            self._half(1, newton_tol)
            # edges (g, g+1) for g = g0-1 .. g0+Nloc-1 -> lam rows g - (g0-1)
            hi = min(self.g0 + self.Nloc, self.N - 1)          # exclusive upper edge id
            seg = self.table[lo:hi + 1]
            self.lam[lo - (self.g0 - 1): hi - (self.g0 - 1)] += self.rho * (seg[:-1] - seg[1:])
            fobj = self.objective(reuse=True)
            obj.append(fobj)
            loss.append(abs(fobj - obj0))
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            ev.synchronize()
            tv.append(ev0.elapsed_time(ev) * 1e-3)
            if verbose and self.dist.rank == 0:
                print(f"  iter {i}: obj {fobj:.9e} gap {loss[-1]:.3e} t {tv[-1]:.3f}s", flush=True)
            if loss[-1] < acc:
                Iter, hit = i, True
                break
        n_out = Iter if hit else len(obj)
        tvec = self.dist_max_times(np.asarray(tv[:n_out]))
        return np.asarray(obj[:n_out]), np.asarray(loss[:n_out]), Iter, tvec

    def dist_max_times(self, tvec: np.ndarray) -> np.ndarray:
        if self.dist.world == 1 or tvec.size == 0:
            return tvec
        t = torch.from_numpy(tvec).to(self.dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return t.cpu().numpy()

    def state(self):
        return {"theta": self.table[self.g0:self.g0 + self.Nloc].cpu().numpy().T,
                "dual": self.lam[1:].cpu().numpy().T, "g0": self.g0,
                "launches": self.launches, **self.stats}


def gadmm_logistic_scaled(N: int, d: int, s: int, rho: float, lam_log: float, num_iter: int,
                          acc: float, dist: Optional[Dist] = None, seed: int = 0, data=None,
                          obj0: Optional[float] = None, newton_tol: float = 1e-9,
                          verbose: bool = False, return_state: bool = False,
                          fast_hessvec: bool = True):
    """Build the problem, solve the consensus optimum (unless ``obj0`` is given) and run GADMM."""
    eng = ScaledLogisticGadmm(N, d, s, lam_log, rho, dist=dist, seed=seed, data=data,
                              fast_hessvec=fast_hessvec)
    if obj0 is None:
        obj0, _ = eng.solve_optimum()
    res = eng.run(num_iter, obj0, acc, newton_tol=newton_tol, verbose=verbose)
    if return_state:
        st = eng.state()
        st["obj0"] = obj0
        return res + (st,)
    return res


def gadmm_logistic_cg(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                      lambda_logistic, acc, dist: Optional[Dist] = None, newton_tol: float = 1e-9,
                      faithful: bool = False, fast_hessvec: bool = True, return_state: bool = False,
                      **_ignored):
    """``group_ADMM_logistic`` (group_ADMM_logistic.m:1) on host data with the Newton-CG engine:
    same signature and outputs as the dense-Hessian :func:`gadmm_logistic_newton`, but it scales to
    wide shards and runs multi-GPU when ``dist`` spans several ranks.  Features are zero-padded to
    a multiple of 4 (16-byte rows for the bulk copies; a zero column keeps its model entry at 0)."""
    N, d, s = int(no_workers), int(num_feature), int(per_samples)
    dpad = (d + 3) // 4 * 4
    X = np.zeros((N, s, dpad), dtype=np.float32)
    X[:, :, :d] = np.asarray(XX, dtype=np.float64)[: N * s, :d].reshape(N, s, d)
    y = np.asarray(YY, dtype=np.float64).reshape(-1)[: N * s].reshape(N, s)
    eng = ScaledLogisticGadmm(N, dpad, s, lambda_logistic, rho, dist=dist, data=(X, y),
                              fast_hessvec=fast_hessvec)
    res = eng.run(num_iter, obj0, acc, newton_tol=newton_tol, faithful=faithful)
    if return_state:
        st = eng.state()
        st["theta"], st["dual"] = st["theta"][:d], st["dual"][:d]
        st["iteration"] = res[2]
        return res + (st,)
    return res
