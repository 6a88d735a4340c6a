"""GPU engine for the first-order baselines (GD, DGD, LAG-PS, LAG-WK, IAG, dual averaging).

Each iteration is a gradient-provider launch (packed-symmetric SYMV with the ``-b`` / objective
epilogue for least squares, the fused logistic gradient for logistic regression) followed by one
fused update kernel; the iteration counter, upload counters and stop flag live on the device, so
windows of iterations are replayed as CUDA graphs.  Semantics: GD_DGD_LAG.m, GD_DGD_LAG_logistic.m,
dual_averaging.m, dual_averaging_logisticReg.m (see ``csrc/first_order.cu``).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..ops import linalg as LA
from ..ops import native as nv


class FirstOrder:
    def __init__(self, XX, YY, N, d, s, logistic=False, lam=0.0, device=None, store="f64",
                 dist=None):
        """``dist`` (a :class:`gadmm_b200.parallel.runtime.Dist` with world > 1) splits the workers
        into contiguous id ranges over the GPUs for the shared-iterate methods; ``dist=None`` keeps
        every worker on this GPU."""
        self.dist = dist if (dist is not None and dist.world > 1) else None
        self.dev = device or (dist.device if dist is not None else
                              torch.device("cuda", torch.cuda.current_device()))
        self.Nglob, self.gid0 = int(N), 0
        XX = np.asarray(XX, dtype=np.float64)[: N * s, :d]
        YY = np.asarray(YY, dtype=np.float64).reshape(-1)[: N * s]
        if self.dist is not None:
            from ..parallel.runtime import partition
            own = np.nonzero(partition(int(N), self.dist.world) == self.dist.rank)[0]
            if own.size == 0:
                raise ValueError("every rank must own at least one worker")
            self.gid0, N = int(own[0]), int(own.size)
            XX = XX[self.gid0 * s:(self.gid0 + N) * s]
            YY = YY[self.gid0 * s:(self.gid0 + N) * s]
        self.N, self.d, self.s, self.logistic, self.lam = int(N), int(d), int(s), logistic, float(lam)
        self.dp = nv.padded(self.d)
        self.X = torch.from_numpy(np.ascontiguousarray(XX.reshape(N, s, d))).to(self.dev)
        self.y = torch.from_numpy(np.ascontiguousarray(YY.reshape(N, s))).to(self.dev)
        G, self.b, self.hyy = LA.gram(self.X, self.y, self.dp)
        self.Gsum = G.sum(0)
        if self.dist is not None:
            self.dist.all_reduce_sum_(self.Gsum)      # setup plumbing (NCCL), not on the iteration path
        self.store = nv.STORE_F64 if store in ("f64", nv.STORE_F64) else nv.STORE_F32
        if not logistic:
            self.P = LA.pack_sym(G, self.d, self.store)
            self.ptrs = LA.ptr_table(self.P)
            self.scratch = torch.empty((N, nv.symv_scratch_elems(self.dp)), dtype=torch.float64,
                                       device=self.dev)
            self.cnt = torch.zeros((N,), dtype=torch.int32, device=self.dev)
        else:
            self.wbuf = torch.empty((N, s), dtype=torch.float64, device=self.dev)
        del G
        self._alloc_tables()

    def _alloc_tables(self):
        """Gradient table (and, across GPUs, the chain-halo tables in symmetric memory)."""
        N = self.N
        self._own_sym = None
        if self.dist is not None:
            # chain-halo tables (DGD / dual averaging across GPUs): Ytab | Ztab[2] | rowflag[2] |
            # fx[2][W] | fxflag[2][W], indexed by GLOBAL worker id, in symmetric memory
            NG, dp, W = self.Nglob, self.dp, self.dist.world
            o_z = NG * dp * 8
            o_rf = o_z + 2 * NG * dp * 8
            o_fx = o_rf + (2 * NG * 8 + 255) // 256 * 256
            o_ff = o_fx + 256 * ((2 * W * 8 + 255) // 256)
            tot = o_ff + 256 * ((2 * W * 8 + 255) // 256)
            sym = self.dist.symm_bytes(tot)
            self._own_sym = (sym, o_z, o_rf, o_fx, o_ff)
            Ytab = sym.local[:o_z].view(torch.float64).view(NG, dp)
            self.Y = Ytab[self.gid0:self.gid0 + N]
        else:
            self.Y = torch.zeros((N, self.dp), dtype=torch.float64, device=self.dev)
        self.fw = torch.zeros((N,), dtype=torch.float64, device=self.dev)
        self.launches = 0

    @classmethod
    def from_parts(cls, P, b, hyy, Gsum, N, d, device=None, dist=None, gid0=0, Nglob=None):
        """Least-squares engine from device-resident pieces (packed Grams ``P`` (N, packed), ``b``
        (N, dp), ``0.5 y'y`` (N,), ``sum_n G_n`` (dp, dp) over ALL workers): the scaled configs
        never hold X.  With ``dist`` (world > 1) the pieces are this rank's workers
        ``gid0 .. gid0+N-1`` of ``Nglob``."""
        self = cls.__new__(cls)
        self.dist = dist if (dist is not None and dist.world > 1) else None
        self.Nglob = int(Nglob if Nglob is not None else N)
        self.gid0 = int(gid0) if self.dist is not None else 0
        self.dev = device or P.device
        self.N, self.d, self.s, self.logistic, self.lam = int(N), int(d), 0, False, 0.0
        self.dp = nv.padded(self.d)
        self.P, self.b, self.hyy, self.Gsum = P, b, hyy, Gsum
        self.store = nv.STORE_F64 if P.dtype == torch.float64 else nv.STORE_F32
        self.ptrs = LA.ptr_table(P)
        self.scratch = torch.empty((N, nv.symv_scratch_elems(self.dp)), dtype=torch.float64,
                                   device=self.dev)
        self.cnt = torch.zeros((N,), dtype=torch.int32, device=self.dev)
        self._alloc_tables()
        return self

    def hmax_per_worker(self, iters: int = 60, tol: float = 1e-10, max_it: int = 20000) -> np.ndarray:
        """``Hmax(i) = max(eig(X{i}'*X{i}))`` for every worker at once: batched power iteration on
        the packed Grams (LinearRegression_Synthetic.m:33).  At least ``iters`` iterations, then
        until every worker's Rayleigh quotient has settled to ``tol`` (relative change over 8
        iterations; clustered spectra need far more than a fixed count), at most ``max_it``.
        ``Hmax`` feeds the LAG-PS trigger and the R-IAG sampling (GD_DGD_LAG.m:8,218,357)."""
        mv = LA.BatchedSymv(self.P, self.dp)
        g = torch.Generator(device=self.dev)
        g.manual_seed(1)
        v = torch.zeros((self.N, self.dp), dtype=torch.float64, device=self.dev)
        v[:, :self.d] = torch.rand((self.N, self.d), dtype=torch.float64, device=self.dev, generator=g) + 0.5
        v /= v.norm(dim=1, keepdim=True)
        w = torch.empty_like(v)
        lam = prev = None
        for i in range(max(int(iters), 1) if tol is None else int(max_it)):
            mv(v, out=w)
            lam = (v * w).sum(1)
            v = w / torch.clamp(w.norm(dim=1, keepdim=True), min=1e-300)
            if tol is not None and i + 1 >= iters and (i + 1) % 8 == 0:
                if prev is not None and bool(((lam - prev).abs() <= tol * lam.abs()).all().item()):
                    break
                prev = lam.clone()
        self.hmax_iterations = i + 1
        return lam.cpu().numpy()

    # ---------------------------------------------------------------------------------------
    def lambda_max(self, tol: float = 1e-15, max_it: int = 50000) -> float:
        """Largest eigenvalue of ``X'X = sum_n G_n`` by power iteration on the packed Gram with the
        streaming SYMV (K11 ``eig_extreme``; GD_DGD_LAG.m:19 ``max(eig(X_fede'*X_fede))``)."""
        P = LA.pack_sym(self.Gsum.unsqueeze(0), self.d)
        mv = LA.BatchedSymv(P, self.dp)
        g = torch.Generator(device=self.dev)
        g.manual_seed(0)
        v = torch.zeros((1, self.dp), dtype=torch.float64, device=self.dev)
        v[0, :self.d] = torch.rand(self.d, dtype=torch.float64, device=self.dev, generator=g) + 0.5
        v /= v.norm()
        lam_prev = 0.0
        w = torch.empty_like(v)
        for i in range(max_it):
            mv(v, out=w)
            if i % 16 == 15:
                lam = float((v * w).sum().item())       # Rayleigh quotient (host sync every 16)
                if abs(lam - lam_prev) <= tol * abs(lam):
                    return lam
                lam_prev = lam
            v = w / w.norm()
        return float((v * mv(v)).sum().item())

    def provider(self, x: torch.Tensor, stride: int):
        L = nv.lib()
        if not self.logistic:
            nv.check(L.gadmm_symv_grad(nv.ptr(self.ptrs), nv.ptr(x), C.c_longlong(stride),
                                       nv.ptr(self.b), nv.ptr(self.hyy), nv.ptr(self.Y),
                                       nv.ptr(self.fw), self.N, self.dp, self.store,
                                       nv.ptr(self.scratch), nv.ptr(self.cnt), nv.stream_ptr()))
        else:
            nv.check(L.gadmm_logistic_grad(nv.ptr(self.X), nv.ptr(self.y), nv.ptr(x),
                                           C.c_longlong(stride), C.c_double(self.lam), self.N,
                                           self.s, self.d, self.dp, nv.ptr(self.Y), nv.ptr(self.fw),
                                           nv.ptr(None), nv.ptr(self.wbuf), nv.stream_ptr()))
        self.launches += 1

    # ---------------------------------------------------------------------------------------
    def _hist(self, T):
        f64, i64 = torch.float64, torch.int64
        return (torch.zeros((T + 2,), dtype=f64, device=self.dev),
                torch.zeros((T + 2,), dtype=f64, device=self.dev),
                torch.zeros((T + 2,), dtype=f64, device=self.dev),
                torch.zeros((T + 2,), dtype=i64, device=self.dev))

    def _drive(self, body, ctrl, T, stop, window):
        """Replay ``window`` iterations per CUDA graph until T iterations ran or the stop flag is
        set (host looks at the flag once per window)."""
        done = 0
        graph, glen = None, 0
        while done < T:
            n = min(window, T - done)
            if n >= 8:
                if graph is None or glen != n:
                    torch.cuda.synchronize(self.dev)
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        for _ in range(n):
                            body()
                    glen = n
                graph.replay()
            else:
                for _ in range(n):
                    body()
            done += n
            if stop and int(ctrl[1].item()):
                break
        torch.cuda.synchronize(self.dev)

    @staticmethod
    def _time(t_hist, n):
        if n <= 0:
            return np.zeros(0)
        t = t_hist[1:n + 1].to(torch.float64)
        live = t > 0                                # iterations restored from a checkpoint: no clock
        t0 = t[live][0] if bool(live.any()) else t[0]
        return torch.clamp((t - t0) * 1e-9, min=0.0).cpu().numpy()

    def run_shared(self, kind, step, thr, Hmax, T, obj0, accuracy=None, uni=None, cumprob=None,
                   window=256, state=None, return_state=False):
        """GD / LAG-PS / LAG-WK / IAG.  ``state`` (from an earlier ``return_state=True`` call of the
        same method and partition) resumes after its last iteration; ``T`` is the absolute last
        iteration.  The state is appended to the returned tuple with ``return_state=True``."""
        N, dp, dev = self.N, self.dp, self.dev
        f64 = torch.float64
        theta = torch.zeros((dp,), dtype=f64, device=dev)
        ghat = torch.zeros((N, dp), dtype=f64, device=dev)
        ghat[:, :self.d] = 1.0                                    # grads = ones (GD_DGD_LAG.m:44)
        that = torch.zeros((N, dp), dtype=f64, device=dev)
        dnorm = torch.zeros((10,), dtype=f64, device=dev)
        obj_h, loss_h, comm_h, t_h = self._hist(T)
        ctrl = torch.tensor([1, 0, 1 if accuracy is not None else 0, 1], dtype=torch.int64, device=dev)
        cnt = torch.zeros((2,), dtype=torch.int32, device=dev)
        done0 = 0
        if state is not None:
            done0 = int(state["iteration"])
            theta.copy_(torch.as_tensor(state["theta"], device=dev))
            ghat.copy_(torch.as_tensor(state["ghat"], device=dev))
            that.copy_(torch.as_tensor(state["theta_hat"], device=dev))
            dnorm.copy_(torch.as_tensor(state["dnorm"], device=dev))
            ctrl[0], ctrl[3] = done0 + 1, int(state["comm_iter"])
            for h, key in ((obj_h, "obj"), (loss_h, "loss"), (comm_h, "comm")):
                h[1:done0 + 1] = torch.as_tensor(np.asarray(state[key][:done0]), device=dev)
        hm = torch.as_tensor(np.asarray(Hmax, dtype=np.float64).reshape(-1)[self.gid0:self.gid0 + N]
                             .copy(), device=dev)
        un = torch.as_tensor(np.asarray(uni if uni is not None else np.zeros(T + 2)), dtype=f64,
                             device=dev)
        cp = torch.as_tensor(np.asarray(cumprob if cumprob is not None else np.ones(self.Nglob)),
                             dtype=f64, device=dev)
        p = nv.GadmmShared()
        p.theta, p.Y, p.fw, p.ghat, p.theta_hat = (theta.data_ptr(), self.Y.data_ptr(),
                                                   self.fw.data_ptr(), ghat.data_ptr(), that.data_ptr())
        p.dnorm, p.Hmax, p.uni, p.cumprob = dnorm.data_ptr(), hm.data_ptr(), un.data_ptr(), cp.data_ptr()
        p.obj_hist, p.loss_hist, p.comm_hist, p.t_hist = (obj_h.data_ptr(), loss_h.data_ptr(),
                                                          comm_h.data_ptr(), t_h.data_ptr())
        p.ctrl, p.counters = ctrl.data_ptr(), cnt.data_ptr()
        p.step, p.thr, p.obj0 = float(step), float(thr), float(obj0)
        p.acc = float(accuracy) if accuracy is not None else 0.0
        p.N, p.d, p.dp, p.kind, p.max_iter, p.logistic = N, self.d, dp, kind, T, int(self.logistic)
        W = self.dist.world if self.dist is not None else 1
        rank = self.dist.rank if self.dist is not None else 0
        split = W > 1 or N * dp >= 65536
        row = dp + 64
        if W > 1:
            sym = self.dist.symm_bytes(2 * W * row * 8 + 256)
            xbuf = sym.local[:2 * W * row * 8].view(f64)
            xflag = sym.local[2 * W * row * 8:2 * W * row * 8 + 2 * W * 8].view(torch.int64)
            xb_peer = torch.tensor(sym.peer_ptrs, dtype=torch.int64, device=dev)
            xf_peer = torch.tensor([q + 2 * W * row * 8 for q in sym.peer_ptrs], dtype=torch.int64,
                                   device=dev)
        else:
            sym = None
            xbuf = torch.zeros((2 * row,), dtype=f64, device=dev)
            xflag = torch.zeros((2,), dtype=torch.int64, device=dev)
            xb_peer = torch.tensor([xbuf.data_ptr()], dtype=torch.int64, device=dev)
            xf_peer = torch.tensor([xflag.data_ptr()], dtype=torch.int64, device=dev)
        dnpart = torch.zeros((dp // 64,), dtype=f64, device=dev)
        cnt2 = torch.zeros((2,), dtype=torch.int32, device=dev)
        p.xbuf, p.xflag, p.xbuf_peer, p.xflag_peer = (xbuf.data_ptr(), xflag.data_ptr(),
                                                      xb_peer.data_ptr(), xf_peer.data_ptr())
        p.dnpart, p.counters2 = dnpart.data_ptr(), cnt2.data_ptr()
        p.Nglob, p.gid0, p.rank, p.nranks, p.split_server = self.Nglob, self.gid0, rank, W, int(split)
        L = nv.lib()

        def body():
            self.provider(theta, 0)
            nv.check(L.gadmm_shared_update(C.byref(p), nv.stream_ptr()))
            self.launches += 2 if split else 1

        self._drive(body, ctrl, T - done0, accuracy is not None, window)
        hit = int(ctrl[1].item())
        n = hit if hit else T
        if self.dist is not None:
            self.dist.barrier()
        res = (obj_h[1:n + 1].cpu().numpy(), loss_h[1:n + 1].cpu().numpy(), self._time(t_h, n),
               comm_h[1:n + 1].cpu().numpy(), theta[:self.d].cpu().numpy())
        if return_state:
            it = int(ctrl[0].item()) - 1
            res = res + ({"iteration": it, "theta": theta.cpu().numpy(), "ghat": ghat.cpu().numpy(),
                          "theta_hat": that.cpu().numpy(), "dnorm": dnorm.cpu().numpy(),
                          "comm_iter": int(ctrl[3].item()), "obj": obj_h[1:it + 1].cpu().numpy(),
                          "loss": loss_h[1:it + 1].cpu().numpy(),
                          "comm": comm_h[1:it + 1].cpu().numpy(), "gid0": self.gid0},)
        keep = (theta, ghat, that, dnorm, hm, un, cp, cnt, sym, xbuf, xflag, xb_peer, xf_peer,
                dnpart, cnt2)                               # alive until here
        del keep
        return res

    def _own(self, step, T, obj0, acc, stop, jacobi=0, state=None):
        N, dp, dev = self.N, self.dp, self.dev
        f64 = torch.float64
        Theta = torch.zeros((N, dp), dtype=f64, device=dev)
        Z = torch.zeros((N, dp), dtype=f64, device=dev)
        obj_h, loss_h, _, t_h = self._hist(T)
        ctrl = torch.tensor([1, 0, 1 if stop else 0, 0], dtype=torch.int64, device=dev)
        if state is not None:                       # resume (single-GPU runs)
            it0 = int(state["iteration"])
            Theta.copy_(torch.as_tensor(state["Theta"], device=dev))
            Z.copy_(torch.as_tensor(state["Z"], device=dev))
            ctrl[0] = it0 + 1
            obj_h[1:it0 + 1] = torch.as_tensor(np.asarray(state["obj"][:it0]), device=dev)
            loss_h[1:it0 + 1] = torch.as_tensor(np.asarray(state["loss"][:it0]), device=dev)
        cnt = torch.zeros((1,), dtype=torch.int32, device=dev)
        p = nv.GadmmOwn()
        p.Theta, p.Y, p.fw, p.Z = Theta.data_ptr(), self.Y.data_ptr(), self.fw.data_ptr(), Z.data_ptr()
        p.obj_hist, p.loss_hist, p.t_hist = obj_h.data_ptr(), loss_h.data_ptr(), t_h.data_ptr()
        p.ctrl, p.counters = ctrl.data_ptr(), cnt.data_ptr()
        p.step, p.obj0, p.acc = float(step), float(obj0), float(acc)
        p.N, p.d, p.dp, p.max_iter, p.jacobi = N, self.d, dp, T, int(jacobi)
        p.Nglob, p.gid0, p.rank, p.nranks = self.Nglob, self.gid0, 0, 1
        p.Ytab = self.Y.data_ptr()
        keep = []
        if self._own_sym is not None:
            sym, o_z, o_rf, o_fx, o_ff = self._own_sym
            W = self.dist.world
            torch.cuda.synchronize(dev)
            self.dist.barrier()                     # nobody is still writing into anybody's tables
            sym.local[o_z:].zero_()                 # Z halos, row flags, objective ring: versions restart
            torch.cuda.synchronize(dev)
            self.dist.barrier()
            base = sym.local.data_ptr()
            ftot = torch.zeros((2,), dtype=f64, device=dev)

            def tab(off):
                return torch.tensor([q + off for q in sym.peer_ptrs], dtype=torch.int64, device=dev)

            yp, zp, rp, fp, ffp = tab(0), tab(o_z), tab(o_rf), tab(o_fx), tab(o_ff)
            p.Ytab, p.Ztab, p.rowflag = base, base + o_z, base + o_rf
            p.fx, p.fxflag, p.ftot = base + o_fx, base + o_ff, ftot.data_ptr()
            p.Ytab_peer, p.Ztab_peer, p.rowflag_peer = yp.data_ptr(), zp.data_ptr(), rp.data_ptr()
            p.fx_peer, p.fxflag_peer = fp.data_ptr(), ffp.data_ptr()
            p.rank, p.nranks = self.dist.rank, W
            keep = [ftot, yp, zp, rp, fp, ffp]
        self._own_keep = keep
        return p, Theta, Z, obj_h, loss_h, t_h, ctrl, cnt

    def _own_state(self, Theta, Z, ctrl, obj_h, loss_h):
        it = int(ctrl[0].item()) - 1
        return {"iteration": it, "Theta": Theta.cpu().numpy(), "Z": Z.cpu().numpy(),
                "obj": obj_h[1:it + 1].cpu().numpy(), "loss": loss_h[1:it + 1].cpu().numpy(),
                "gid0": self.gid0}

    def run_dgd(self, step, T, obj0, window=256, state=None, return_state=False):
        if state is not None and self.dist is not None:
            raise ValueError("DGD resume is implemented for single-GPU runs")
        p, Theta, Z, obj_h, loss_h, t_h, ctrl, cnt = self._own(step, T, obj0, 0.0, False, state=state)
        done0 = int(state["iteration"]) if state is not None else 0
        L = nv.lib()

        def body():
            self.provider(Theta, self.dp)
            nv.check(L.gadmm_own_exchange(C.byref(p), 1, nv.stream_ptr()))
            nv.check(L.gadmm_dgd_update(C.byref(p), nv.stream_ptr()))
            self.launches += 2 if p.nranks > 1 else 1

        self._drive(body, ctrl, T - done0, False, window)
        if self.dist is not None:
            self.dist.barrier()
        res = (obj_h[1:T + 1].cpu().numpy(), loss_h[1:T + 1].cpu().numpy(), self._time(t_h, T))
        return res + (self._own_state(Theta, Z, ctrl, obj_h, loss_h),) if return_state else res

    def run_dualavg(self, alpha, T, obj0, acc, jacobi=False, window=256, state=None,
                    return_state=False):
        if state is not None and self.dist is not None:
            raise ValueError("dual-averaging resume is implemented for single-GPU runs")
        p, Theta, Z, obj_h, loss_h, t_h, ctrl, cnt = self._own(alpha, T, obj0, acc, True, jacobi,
                                                                state=state)
        done0 = int(state["iteration"]) if state is not None else 0
        L = nv.lib()

        def body():
            self.provider(Theta, self.dp)
            nv.check(L.gadmm_own_exchange(C.byref(p), 0, nv.stream_ptr()))
            nv.check(L.gadmm_dualavg_update(C.byref(p), 1, nv.stream_ptr()))
            self.launches += 2 if p.nranks > 1 else 1

        self._drive(body, ctrl, T - done0, True, window)
        if not int(ctrl[1].item()):
            # close the last iteration: objective of the final iterate
            self.provider(Theta, self.dp)
            nv.check(L.gadmm_own_exchange(C.byref(p), 0, nv.stream_ptr()))
            nv.check(L.gadmm_dualavg_update(C.byref(p), 0, nv.stream_ptr()))
            torch.cuda.synchronize(self.dev)
        hit = int(ctrl[1].item())
        n = hit if hit else T
        th = Theta[:, :self.d]
        if self.dist is not None:               # assemble the full d x N iterate on every rank
            full = torch.zeros((self.Nglob, self.d), dtype=torch.float64, device=self.dev)
            full[self.gid0:self.gid0 + self.N] = th
            self.dist.all_reduce_sum_(full)
            self.dist.barrier()
            th = full
        res = (obj_h[1:n + 1].cpu().numpy(), loss_h[1:n + 1].cpu().numpy(), n, self._time(t_h, n),
               th.cpu().numpy().T)
        return res + (self._own_state(Theta, Z, ctrl, obj_h, loss_h),) if return_state else res
