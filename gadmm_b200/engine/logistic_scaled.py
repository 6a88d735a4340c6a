"""Logistic GADMM with exact local solves at scale, single- and multi-GPU (K9, BASELINE config 3).

``group_ADMM_logistic.m:48-55`` hands every worker's strongly convex sub-problem to CVX.  At the
reference's 50x50 shards an explicit Hessian is the right tool (:func:`gadmm_logistic_newton`);
at 512 workers x d = 8192 with s < d it is not -- forming ``X'DX`` costs ``2 s d^2`` flop per
Newton step and worker, a Hessian *product* costs ``4 s d`` and one read of the shard.  Here the
local solve is a batched Newton-CG: all workers of one colour advance in lock-step, and each
gradient evaluation or Hessian product is ONE pass over the fp32 shards by
``logistic_xtwx_kernel`` (csrc/logistic_scaled.cu), accumulated in fp64.

The GADMM run itself is device-resident (csrc/logistic_scaled.cu, ``GadmmLogi``): the CG
recurrences, the forcing-term test, the Armijo line search, the publication of the models and the
dual / objective / stop step are kernels, and the loops around them (iterations, Newton steps, CG
steps, line-search trials) are WHILE nodes of ONE CUDA graph whose conditions are set on the
device.  Between the graph launch and the stop there is no host round trip, no torch op and no
NCCL call.

Multi-GPU: contiguous chain segments per rank (as in :class:`ChainEngine`).  The model table lives
in symmetric memory; the publish kernel of a colour stores the boundary models straight into the
neighbouring GPUs' tables over NVLink and raises a per-worker version flag (``st.release.sys``);
the consumers' setup kernel spins on it (``ld.acquire.sys``).  The objective is summed over the
ranks through the same kind of peer stores + flags, which also keeps the ranks within one
iteration of each other.  Edge duals are kept by the left end-point's rank and recomputed
redundantly for the one edge that crosses into a rank.

Shards are fp32 when the data is exactly representable in fp32 (the synthetic configs; half the
HBM bytes per pass) and fp64 otherwise, so user data is never rounded.
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




_loops_ok = {}


def graph_loops_available(dev=None) -> bool:
    """Conditional (WHILE) graph nodes usable on this device?  Probed once per device with a
    three-iteration loop (``gadmm_graph_loops_selftest``)."""
    dev = dev if dev is not None else torch.cuda.current_device()
    key = int(dev.index) if isinstance(dev, torch.device) and dev.index is not None else int(
        torch.cuda.current_device() if isinstance(dev, torch.device) else dev)
    if key not in _loops_ok:
        with torch.cuda.device(key):
            _loops_ok[key] = nv.lib().gadmm_graph_loops_selftest(3, nv.stream_ptr()) == 0
    return _loops_ok[key]


class ScaledLogisticGadmm:
    """GADMM for consensus logistic regression with Newton-CG local solves.

    ``data=None`` draws a synthetic problem on the owning GPU (Philox stream = global worker id,
    so the data set does not depend on the number of ranks): ``X_n = Z_n / sqrt(d)``,
    ``y = sign(X theta* + 0.5 noise)``.  ``data=(X, y)`` takes host arrays ``(N, s, d)`` /
    ``(N, s)`` (every rank passes the full set and keeps its slice).  ``store``: ``"f32"`` /
    ``"f64"`` shards, ``"auto"`` = fp32 only if that is exact for the given data."""

    def __init__(self, N: int, d: int, s: int, lam_log: float, rho: float,
                 dist: Optional[Dist] = None, seed: int = 0, data: Optional[Tuple] = None,
                 fast_hessvec: bool = True, store: str = "auto", max_iter: int = 1000):
        self.dist = dist or Dist()
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
        Xh = yh = None
        if data is not None:
            Xh, yh = data
            Xh = np.asarray(Xh).reshape(self.N, self.s, self.d)[self.g0:self.g0 + self.Nloc]
            yh = np.asarray(yh).reshape(self.N, self.s)[self.g0:self.g0 + self.Nloc]
        if store == "auto":
            exact = data is None or Xh.dtype == np.float32 or \
                bool(np.array_equal(Xh.astype(np.float32).astype(np.float64), Xh))
            if self.dist.world > 1 and data is not None:      # one storage type for all ranks
                flag = torch.tensor([0.0 if exact else 1.0], dtype=f64, device=dev)
                self.dist.all_reduce_sum_(flag)
                exact = float(flag.item()) == 0.0
            store = "f32" if exact else "f64"
        if store not in ("f32", "f64"):
            raise ValueError("store must be 'auto', 'f32' or 'f64'")
        self.store = nv.STORE_F32 if store == "f32" else nv.STORE_F64
        xdt = torch.float32 if store == "f32" else f64
        # fp32-multiply Hessian products only where they are faster (fp32 shards, d <= 5120): for
        # wider rows the kernel streams two rows per block and keeps the converted block in
        # registers, and the exact fp64 product runs at the same rate
        self.hv_mode = 2 if (fast_hessvec and store == "f32" and self.d * 40 <= 200 * 1024) else 1
        self.X = torch.empty((self.Nloc, self.s, self.d), dtype=xdt, device=dev)
        self.y = torch.empty((self.Nloc, self.s), dtype=f64, device=dev)
        if data is None:
            th_true = TP.philox_normal((self.d,), seed * 7919 + 17, 0xFFFF, 0.0, 2.0, device=dev)
            tmp = self.X[0] if store == "f32" else torch.empty((self.s, self.d),
                                                               dtype=torch.float32, device=dev)
            for k in range(self.Nloc):
                gid = self.g0 + k
                buf = self.X[k] if store == "f32" else tmp
                TP.philox_normal(None, seed, gid, 0.0, 1.0 / np.sqrt(self.d), out=buf)
                if store != "f32":
                    self.X[k].copy_(buf)
                z = self.X[k].double() @ th_true
                z += TP.philox_normal((self.s,), seed + 1, gid, 0.0, 0.5, device=dev)
                self.y[k] = torch.where(z >= 0, 1.0, -1.0)
        else:
            npdt = np.float32 if store == "f32" else np.float64
            self.X.copy_(torch.from_numpy(np.ascontiguousarray(Xh, dtype=npdt)))
            self.y.copy_(torch.from_numpy(np.ascontiguousarray(yh, dtype=np.float64)))
        self.Dw = torch.zeros((self.Nloc, self.s), dtype=f64, device=dev)
        self.max_iter = int(max_iter)
        self._alloc_state()
        self.stats = {}
        self.launches = 0
        self._scratch = None
        self._graph = None
        self._graph_key = None
        self._primed = False

    def _alloc_state(self):
        """Model table, flags and objective ring in symmetric memory; work vectors of the batched
        solve; control words and curves (``GadmmLogi`` in include/gadmm_b200.h)."""
        N, d, Nloc, W, dev = self.N, self.d, self.Nloc, self.dist.world, self.dev
        f64, i64, i32 = torch.float64, torch.int64, torch.int32
        sz_tab = (N * d * 8 + 255) // 256 * 256
        sz_fl = (N * 8 + 255) // 256 * 256
        sz_op = (4 * W * 8 + 255) // 256 * 256
        self._sym = self.dist.symm_bytes(sz_tab + sz_fl + 2 * sz_op)
        buf = self._sym.local
        self.table = buf[:N * d * 8].view(f64).view(N, d)
        self.flags = buf[sz_tab:sz_tab + N * 8].view(i64)
        self.objpart = buf[sz_tab + sz_fl:sz_tab + sz_fl + 4 * W * 8].view(f64)
        self.objflag = buf[sz_tab + sz_fl + sz_op:sz_tab + sz_fl + sz_op + 4 * W * 8].view(i64)

        def tab(off):
            return torch.tensor([q + off for q in self._sym.peer_ptrs], dtype=i64, device=dev)

        self._peer = (tab(0), tab(sz_tab), tab(sz_tab + sz_fl), tab(sz_tab + sz_fl + sz_op))
        # lam[k] couples global workers (g0-1+k, g0+k), k = 0..Nloc  (edge -1 / N-1 do not exist)
        self.lam = torch.zeros((Nloc + 1, d), dtype=f64, device=dev)
        first = [(c - self.g0) % 2 for c in (0, 1)]
        self.rows = [((Nloc - f + 1) // 2 if f < Nloc else 0) for f in first]
        self.first = first
        R = max(self.rows)
        self.R = R
        self.G = [_pick_ctas_per_shard(max(r, 1), self.s) for r in self.rows]
        self._vec = {n: torch.zeros((R, d), dtype=f64, device=dev)
                     for n in ("x", "g", "p", "r", "q", "hq", "xtry", "lin")}
        self.vlast = torch.zeros((2, R, d), dtype=f64, device=dev)
        self._sc = {n: torch.zeros((R,), dtype=f64, device=dev)
                    for n in ("mu", "f", "gn", "eta", "rs", "t", "slope", "losstry")}
        self.losslast = torch.zeros((2, R), dtype=f64, device=dev)
        self.live = torch.zeros((R,), dtype=i32, device=dev)
        self.objw = torch.zeros((Nloc,), dtype=f64, device=dev)
        need = max(int(nv.lib().gadmm_logistic_xtwx_scratch_elems(max(r, 1), d, g))
                   for r, g in zip(self.rows, self.G))
        self.pass_scratch = torch.empty((need,), dtype=f64, device=dev)
        self.ctrl = torch.zeros((16,), dtype=i64, device=dev)
        self.ctrl[0] = 1
        self.cond = torch.zeros((4,), dtype=i32, device=dev)
        self.counters = torch.zeros((8,), dtype=i32, device=dev)
        self.obj_hist = torch.zeros((self.max_iter + 2,), dtype=f64, device=dev)
        self.loss_hist = torch.zeros((self.max_iter + 2,), dtype=f64, device=dev)
        self.t_hist = torch.zeros((self.max_iter + 2,), dtype=i64, device=dev)

    # ------------------------------------------------------------------ one pass over shards
    def _xtwx(self, x: torch.Tensor, first: int, count: int, stride: int, hv: int,
              shared_x: bool = False):
        """``v[k] = X' w(X x[k])`` for local shards ``first, first+stride, ...`` (``count``).
        ``hv``: 0 gradient weights (+ loss, curvature state), 1 Hessian product in fp64,
        2 Hessian product with fp32 multiplies (fp32 shards only)."""
        G = _pick_ctas_per_shard(count, self.s)
        need = int(nv.lib().gadmm_logistic_xtwx_scratch_elems(count, self.d, G))
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty((need,), dtype=torch.float64, device=self.dev)
        v = torch.empty((count, self.d), dtype=torch.float64, device=self.dev)
        loss = torch.zeros((count,), dtype=torch.float64, device=self.dev)
        esz = self.s * self.d * self.X.element_size()
        nv.check(nv.lib().gadmm_logistic_xtwx_ex(
            nv.ptr(self.X.data_ptr() + first * esz), int(self.store),
            nv.ptr(self.y.data_ptr() + first * self.s * 8),
            nv.ptr(x), C.c_longlong(0 if shared_x else self.d),
            nv.ptr(self.Dw.data_ptr() + first * self.s * 8), nv.ptr(v), self.d, nv.ptr(loss),
            count, self.s, self.d, G, int(hv), stride, nv.ptr(self._scratch),
            nv.stream_ptr()))
        self.launches += 2
        return v, loss

    # ------------------------------------------------------------------ consensus optimum
    def solve_optimum(self, tol: float = 1e-10):
        """``min_theta sum_n lam/2 |theta|^2 + loss_n(theta)`` by Newton-CG over all shards of all
        ranks (one shared vector, partial sums all-reduced; set-up, outside the GADMM run).
        Returns ``(obj0, theta)``."""
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
        self._primed = False               # the curvature state now belongs to the consensus point
        return float(f.item()), x[0]

    # ------------------------------------------------------------------ descriptor
    def desc(self, obj0: float, acc: float, tol: float, stop: bool, max_newton: int = 40,
             max_cg: int = 250, max_ls: int = 30) -> nv.GadmmLogi:
        p = nv.GadmmLogi()
        p.X, p.y, p.Dw = self.X.data_ptr(), self.y.data_ptr(), self.Dw.data_ptr()
        p.table, p.flags = self.table.data_ptr(), self.flags.data_ptr()
        p.objpart, p.objflag = self.objpart.data_ptr(), self.objflag.data_ptr()
        tp, fp, op, ofp = self._peer
        p.table_peer, p.flags_peer = tp.data_ptr(), fp.data_ptr()
        p.objpart_peer, p.objflag_peer = op.data_ptr(), ofp.data_ptr()
        p.lam = self.lam.data_ptr()
        for n, t in self._vec.items():
            setattr(p, n, t.data_ptr())
        for n, t in self._sc.items():
            setattr(p, n, t.data_ptr())
        p.vlast, p.losslast, p.live = self.vlast.data_ptr(), self.losslast.data_ptr(), self.live.data_ptr()
        p.objw, p.scratch = self.objw.data_ptr(), self.pass_scratch.data_ptr()
        p.ctrl, p.cond, p.counters = self.ctrl.data_ptr(), self.cond.data_ptr(), self.counters.data_ptr()
        p.obj_hist, p.loss_hist, p.t_hist = (self.obj_hist.data_ptr(), self.loss_hist.data_ptr(),
                                             self.t_hist.data_ptr())
        p.rho, p.lam_log, p.obj0, p.acc, p.tol = self.rho, self.lam_log, float(obj0), float(acc), float(tol)
        p.N, p.Nloc, p.g0, p.s, p.d = self.N, self.Nloc, self.g0, self.s, self.d
        p.rank, p.nranks, p.store, p.hv_mode = self.dist.rank, self.dist.world, int(self.store), self.hv_mode
        p.max_iter, p.max_newton, p.max_cg, p.max_ls = self.max_iter, int(max_newton), int(max_cg), int(max_ls)
        p.G[0], p.G[1] = int(self.G[0]), int(self.G[1])
        return p

    # ------------------------------------------------------------------ GADMM
    def run(self, num_iter: int, obj0: float, acc: float, newton_tol: float = 1e-9,
            faithful: bool = False, verbose: bool = False, mode: str = "auto",
            max_newton: int = 40, max_cg: int = 250):
        """``(obj, loss, Iter, time)`` as ``group_ADMM_logistic`` returns them; time is the device
        clock (``%globaltimer``) at the end of every iteration, seconds since the launch, max over
        ranks.  ``mode="graph"``: the whole run is one launch of a CUDA graph with device-side loops;
        ``"host"``: the same kernels with every loop test read back by the host (debugging, and the
        fall-back where conditional graph nodes are unavailable); ``"auto"`` probes once.
        ``num_iter`` is the (absolute) last iteration to run; curves cover iterations 1..done."""
        num_iter = int(num_iter)
        if num_iter > self.max_iter:
            raise ValueError(f"num_iter={num_iter} exceeds the engine's max_iter={self.max_iter}")
        if mode == "auto":
            import os
            mode = "host" if os.environ.get("GADMM_B200_LOGI_HOST") == "1" or \
                not graph_loops_available(self.dev) else "graph"
        L = nv.lib()
        self.ctrl[1] = 0
        self.ctrl[2] = 1
        self.ctrl[3] = num_iter                              # last iteration of this launch
        p = self.desc(obj0, acc, newton_tol, True, max_newton, max_cg)
        self._p = p
        if not self._primed:
            nv.check(L.gadmm_logi_prime(C.byref(p), nv.stream_ptr()))
            self._primed = True
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()
        if mode == "graph":
            key = (float(obj0), float(acc), float(newton_tol), int(max_newton), int(max_cg))
            if self._graph is None or self._graph_key != key:
                if self._graph is not None:
                    L.gadmm_logi_graph_destroy(self._graph)
                h = C.c_void_p()
                nv.check(L.gadmm_logi_graph_create(C.byref(p), C.byref(h)))
                self._graph, self._graph_key = h, key
            nv.check(L.gadmm_logi_graph_launch(self._graph, nv.stream_ptr()))
            self.launches += 1
        elif mode == "host":
            nv.check(L.gadmm_logi_run_host(C.byref(p), nv.stream_ptr()))
        else:
            raise ValueError("mode must be 'auto', 'graph' or 'host'")
        torch.cuda.synchronize(self.dev)
        hit = int(self.ctrl[1].item())
        n_out = int(self.ctrl[0].item()) - 1                 # iterations completed so far (absolute)
        Iter = hit if hit else (100 if faithful else num_iter)
        obj = self.obj_hist[1:n_out + 1].cpu().numpy().copy()
        loss = self.loss_hist[1:n_out + 1].cpu().numpy().copy()
        # iterations restored from a checkpoint carry no clock
        t = torch.clamp((self.t_hist[1:n_out + 1] - self.t_hist[0]).to(torch.float64) * 1e-9, min=0.0)
        if self.dist.world > 1 and n_out:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        self.stats = {"hessvecs": int(self.ctrl[8].item()), "evals": int(self.ctrl[9].item()),
                      "newton_steps": int(self.ctrl[10].item()), "mode": mode,
                      "wait_ms": [round(int(self.ctrl[i].item()) * 1e-6, 3) for i in (11, 12, 13)]}
        if verbose and self.dist.rank == 0:
            for i in range(n_out):
                print(f"  iter {1 + i}: obj {obj[i]:.9e} gap {loss[i]:.3e} t {float(t[i]):.3f}s",
                      flush=True)
        return obj, loss, Iter, t.cpu().numpy()

    def objective(self) -> float:
        """``sum_n lam/2 |theta_n|^2 + loss_n(theta_n)`` over all ranks at the current models (one
        pass over all local shards; leaves the curvature state at those models)."""
        th = self.table[self.g0:self.g0 + self.Nloc].contiguous()
        _, loss = self._xtwx(th, 0, self.Nloc, 1, 0)
        part = (loss.sum() + 0.5 * self.lam_log * (th * th).sum()).reshape(1)
        self.dist.all_reduce_sum_(part)
        return float(part.item())

    def state(self):
        """Checkpointable state: this rank's models and edge duals, the iteration, the curves."""
        torch.cuda.synchronize(self.dev)
        it = int(self.ctrl[0].item()) - 1
        return {"theta": self.table[self.g0:self.g0 + self.Nloc].cpu().numpy().T.copy(),
                "dual": self.lam[1:].cpu().numpy().T.copy(), "dual_left": self.lam[0].cpu().numpy().copy(),
                "g0": self.g0, "iteration": it,
                "obj": self.obj_hist[1:it + 1].cpu().numpy().copy(),
                "loss": self.loss_hist[1:it + 1].cpu().numpy().copy(),
                "launches": self.launches, **self.stats}

    def load_state(self, st):
        """Resume from :meth:`state` of the same partition: the next iteration is
        ``st['iteration'] + 1``.  The halo rows are refreshed from the neighbours."""
        it = int(st["iteration"])
        th = torch.as_tensor(np.asarray(st["theta"], dtype=np.float64).T.copy(), device=self.dev)
        full = torch.zeros((self.N, self.d), dtype=torch.float64, device=self.dev)
        full[self.g0:self.g0 + self.Nloc] = th
        self.dist.all_reduce_sum_(full)
        self.table.copy_(full)
        self.lam[1:] = torch.as_tensor(np.asarray(st["dual"], dtype=np.float64).T.copy(), device=self.dev)
        self.lam[0] = torch.as_tensor(np.asarray(st["dual_left"], dtype=np.float64), device=self.dev)
        self.flags.fill_(it)
        self.objflag.zero_()
        n = min(it, len(st.get("obj", [])))
        if n:
            self.obj_hist[1:n + 1] = torch.as_tensor(np.asarray(st["obj"][:n]), device=self.dev)
            self.loss_hist[1:n + 1] = torch.as_tensor(np.asarray(st["loss"][:n]), device=self.dev)
        self.ctrl.zero_()
        self.ctrl[0] = it + 1
        self._primed = False
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()

    def __del__(self):
        try:
            if self._graph is not None:
                nv.lib().gadmm_logi_graph_destroy(self._graph)
        except Exception:
            pass


def gadmm_logistic_scaled(N: int, d: int, s: int, rho: float, lam_log: float, num_iter: int,
                          acc: float, dist: Optional[Dist] = None, seed: int = 0, data=None,
                          obj0: Optional[float] = None, newton_tol: float = 1e-9,
                          verbose: bool = False, return_state: bool = False,
                          fast_hessvec: bool = True, store: str = "auto", mode: str = "auto"):
    """Build the problem, solve the consensus optimum (unless ``obj0`` is given) and run GADMM."""
    eng = ScaledLogisticGadmm(N, d, s, lam_log, rho, dist=dist, seed=seed, data=data,
                              fast_hessvec=fast_hessvec, store=store, max_iter=int(num_iter))
    if obj0 is None:
        obj0, _ = eng.solve_optimum()
    res = eng.run(num_iter, obj0, acc, newton_tol=newton_tol, verbose=verbose, mode=mode)
    if return_state:
        st = eng.state()
        st["obj0"] = obj0
        return res + (st,)
    return res


def gadmm_logistic_cg(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                      lambda_logistic, acc, dist: Optional[Dist] = None, newton_tol: float = 1e-9,
                      faithful: bool = False, fast_hessvec: bool = True, return_state: bool = False,
                      store: str = "auto", mode: str = "auto", state=None, **_ignored):
    """``group_ADMM_logistic`` (group_ADMM_logistic.m:1) on host data with the Newton-CG engine:
    same signature and outputs as the dense-Hessian :func:`gadmm_logistic_newton`, but it scales to
    wide shards and runs multi-GPU when ``dist`` spans several ranks.  The data stay in double
    unless they are exactly representable in fp32 (``store="auto"``).  Features are zero-padded to
    a multiple of 4 (16-byte rows for the bulk copies; a zero column keeps its model entry at 0)."""
    N, d, s = int(no_workers), int(num_feature), int(per_samples)
    dpad = (d + 3) // 4 * 4
    X = np.zeros((N, s, dpad), dtype=np.float64)
    X[:, :, :d] = np.asarray(XX, dtype=np.float64)[: N * s, :d].reshape(N, s, d)
    y = np.asarray(YY, dtype=np.float64).reshape(-1)[: N * s].reshape(N, s)
    eng = ScaledLogisticGadmm(N, dpad, s, lambda_logistic, rho, dist=dist, data=(X, y),
                              fast_hessvec=fast_hessvec, store=store, max_iter=int(num_iter))
    if state is not None:
        eng.load_state(state)
    res = eng.run(num_iter, obj0, acc, newton_tol=newton_tol, faithful=faithful, mode=mode)
    if return_state:
        st = eng.state()
        st["theta"], st["dual"] = st["theta"][:d], st["dual"][:d]
        return res + (st,)
    return res
