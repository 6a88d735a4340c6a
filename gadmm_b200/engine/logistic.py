"""Logistic-regression GADMM on the GPU (K9).

* :func:`gadmm_logistic_gd` -- the reference's inexact local solve (<=100 frozen-penalty gradient
  steps per worker, ``group_ADMM_logistic_GD.m`` / ``logReg_GD.m``): one CTA per worker runs the
  whole inner loop on chip; an outer iteration is 4 launches and windows replay as CUDA graphs.
* :func:`gadmm_logistic_newton` -- exact local solves (``group_ADMM_logistic.m`` calls CVX; here
  damped Newton): per step the fused logistic-gradient kernel yields gradient, loss and the
  Newton weights, the Hessian ``X'DX + (lam + c rho) I`` is a Gram of the row-scaled shard,
  inverted by the blocked Cholesky kernels and applied with the streaming SYMV.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..ops import linalg as LA
from ..ops import native as nv


class _LogisticState:
    def __init__(self, XX, YY, N, d, s, lam_log, device=None):
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.N, self.d, self.s, self.lam_log = int(N), int(d), int(s), float(lam_log)
        self.dp = nv.padded(self.d)
        XX = np.asarray(XX, dtype=np.float64)[: N * s, :d]
        YY = np.asarray(YY, dtype=np.float64).reshape(-1)[: N * s]
        self.X = torch.from_numpy(np.ascontiguousarray(XX.reshape(N, s, d))).to(self.dev)
        self.y = torch.from_numpy(np.ascontiguousarray(YY.reshape(N, s))).to(self.dev)
        f64 = torch.float64
        self.theta = torch.zeros((N, self.dp), dtype=f64, device=self.dev)
        self.lam = torch.zeros((N, self.dp), dtype=f64, device=self.dev)
        self.Y = torch.zeros((N, self.dp), dtype=f64, device=self.dev)
        self.fw = torch.zeros((N,), dtype=f64, device=self.dev)
        self.wout = torch.zeros((N, s), dtype=f64, device=self.dev)
        self.wbuf = torch.zeros((N, s), dtype=f64, device=self.dev)
        self.launches = 0

    def grad(self, theta, want_w=False):
        nv.check(nv.lib().gadmm_logistic_grad(
            nv.ptr(self.X), nv.ptr(self.y), nv.ptr(theta), C.c_longlong(self.dp),
            C.c_double(self.lam_log), self.N, self.s, self.d, self.dp, nv.ptr(self.Y),
            nv.ptr(self.fw), nv.ptr(self.wout if want_w else None), nv.ptr(self.wbuf),
            nv.stream_ptr()))
        self.launches += 1


def gadmm_logistic_gd(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                      lambda_logistic, acc, gdStep, naive=True, window=16, return_state=False,
                      dist=None):
    """``group_ADMM_logistic_GD`` on the GPU: ``(obj, loss, Iter, time)``.  With a multi-rank
    ``dist`` the chain is split over the GPUs (:func:`_gadmm_logistic_gd_dist`)."""
    if dist is not None and getattr(dist, "world", 1) > 1:
        return _gadmm_logistic_gd_dist(XX, YY, rho, no_workers, num_feature, per_samples, num_iter,
                                       obj0, lambda_logistic, acc, gdStep, dist, return_state)
    st = _LogisticState(XX, YY, no_workers, num_feature, per_samples, lambda_logistic)
    dev, T = st.dev, int(num_iter)
    f64 = torch.float64
    obj_h = torch.zeros((T + 2,), dtype=f64, device=dev)
    loss_h = torch.zeros((T + 2,), dtype=f64, device=dev)
    t_h = torch.zeros((T + 2,), dtype=torch.int64, device=dev)
    ctrl = torch.tensor([1, 0, 1, 0], dtype=torch.int64, device=dev)
    cnt = torch.zeros((1,), dtype=torch.int32, device=dev)
    inner = torch.zeros((1,), dtype=torch.int64, device=dev)
    p = nv.GadmmLogistic()
    p.X, p.y, p.theta, p.lam, p.fw = (st.X.data_ptr(), st.y.data_ptr(), st.theta.data_ptr(),
                                      st.lam.data_ptr(), st.fw.data_ptr())
    p.inner_steps, p.ctrl, p.counters = inner.data_ptr(), ctrl.data_ptr(), cnt.data_ptr()
    p.obj_hist, p.loss_hist, p.t_hist = obj_h.data_ptr(), loss_h.data_ptr(), t_h.data_ptr()
    p.rho, p.lam_log, p.gd_step, p.inner_tol = float(rho), float(lambda_logistic), float(gdStep), 1e-4
    p.obj0, p.acc = float(obj0), float(acc)
    p.N, p.s, p.d, p.dp, p.max_iter, p.max_inner = st.N, st.s, st.d, st.dp, T, 100
    L = nv.lib()

    def body():
        nv.check(L.gadmm_logistic_half_gd(C.byref(p), 0, nv.stream_ptr()))
        nv.check(L.gadmm_logistic_half_gd(C.byref(p), 1, nv.stream_ptr()))
        st.grad(st.theta)
        nv.check(L.gadmm_logistic_dual_obj(C.byref(p), nv.stream_ptr()))
        st.launches += 3

    body()                                   # first iteration outside capture (attribute set-up)
    done, graph = 1, None
    while done < T and not int(ctrl[1].item()):
        n = min(window, T - done)
        if n == window:
            if graph is None:
                torch.cuda.synchronize(dev)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    for _ in range(n):
                        body()
            graph.replay()
        else:
            for _ in range(n):
                body()
        done += n
    torch.cuda.synchronize(dev)
    hit = int(ctrl[1].item())
    it = hit if hit else T
    t = t_h[1:it + 1].to(f64)
    res = (obj_h[1:it + 1].cpu().numpy(), loss_h[1:it + 1].cpu().numpy(), it,
           ((t - t[0]) * 1e-9).cpu().numpy())
    if return_state:
        return res + ({"theta": st.theta[:, :st.d].cpu().numpy().T, "dual": st.lam[:, :st.d].cpu().numpy().T,
                       "iteration": it, "inner_steps": int(inner.item()), "launches": st.launches},)
    return res


def _gadmm_logistic_gd_dist(XX, YY, rho, N, d, s, num_iter, obj0, lam_log, acc, gdStep, dist,
                            return_state=False):
    """Inexact-solve logistic GADMM with contiguous chain segments per GPU.  Same kernels as the
    single-GPU path (``logistic_half_gd_kernel`` on the rank's workers, the fused loss kernel); the
    model table is a symmetric buffer replicated on every rank, a rank stores its boundary models
    into the neighbouring rank's replica over NVLink after each colour, and the ranks meet at a
    barrier (an outer iteration is up to 200 inner gradient steps per worker, so the exchange is
    not worth fusing).  Duals of the one edge that crosses into a rank are recomputed there."""
    from ..parallel.runtime import partition
    dev = dist.device
    N, d, s, T = int(N), int(d), int(s), int(num_iter)
    dp = nv.padded(d)
    owner = partition(N, dist.world)
    mine = np.nonzero(owner == dist.rank)[0]
    if mine.size == 0:
        raise ValueError("more ranks than workers")
    g0, nloc = int(mine[0]), int(mine.size)
    f64 = torch.float64
    XXa = np.asarray(XX, dtype=np.float64)[: N * s, :d].reshape(N, s, d)[g0:g0 + nloc]
    YYa = np.asarray(YY, dtype=np.float64).reshape(-1)[: N * s].reshape(N, s)[g0:g0 + nloc]
    X = torch.from_numpy(np.ascontiguousarray(XXa)).to(dev)
    y = torch.from_numpy(np.ascontiguousarray(YYa)).to(dev)
    sym = dist.symm_bytes(N * dp * 8)
    theta = sym.local[: N * dp * 8].view(f64).view(N, dp)
    lam = torch.zeros((N, dp), dtype=f64, device=dev)
    fw = torch.zeros((nloc,), dtype=f64, device=dev)
    Yg = torch.zeros((nloc, dp), dtype=f64, device=dev)
    wbuf = torch.zeros((nloc, s), dtype=f64, device=dev)
    ctrl = torch.tensor([1, 0, 0, 0], dtype=torch.int64, device=dev)      # stop decided on the host
    inner = torch.zeros((1,), dtype=torch.int64, device=dev)
    p = nv.GadmmLogistic()
    p.X, p.y, p.theta, p.lam, p.fw = X.data_ptr(), y.data_ptr(), theta.data_ptr(), lam.data_ptr(), fw.data_ptr()
    p.inner_steps, p.ctrl = inner.data_ptr(), ctrl.data_ptr()
    p.rho, p.lam_log, p.gd_step, p.inner_tol = float(rho), float(lam_log), float(gdStep), 1e-4
    p.obj0, p.acc = float(obj0), float(acc)
    p.N, p.s, p.d, p.dp, p.max_iter, p.max_inner = N, s, d, dp, T + 1, 100
    p.n0, p.nloc = g0, nloc
    L = nv.lib()

    def push(colour):
        for gid, nb in ((g0, g0 - 1), (g0 + nloc - 1, g0 + nloc)):
            if gid % 2 == colour and 0 <= nb < N:
                view = sym.handle.get_buffer(int(owner[nb]), (N, dp), f64)
                view[gid].copy_(theta[gid])

    lo, hi = max(g0 - 1, 0), min(g0 + nloc, N - 1)       # edges lo .. hi-1 touch this rank
    obj, loss, tv = [], [], []
    it, hit = T, False
    dist.barrier()
    ev0 = torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(1, T + 1):
        for colour in (0, 1):
            nv.check(L.gadmm_logistic_half_gd(C.byref(p), colour, nv.stream_ptr()))
            push(colour)
            dist.barrier()
        lam[lo:hi] += rho * (theta[lo:hi] - theta[lo + 1:hi + 1])
        nv.check(L.gadmm_logistic_grad(nv.ptr(X), nv.ptr(y), nv.ptr(theta[g0:]), C.c_longlong(dp),
                                       C.c_double(lam_log), nloc, s, d, dp, nv.ptr(Yg), nv.ptr(fw),
                                       nv.ptr(None), nv.ptr(wbuf), nv.stream_ptr()))
        tot = fw.sum().reshape(1)
        dist.all_reduce_sum_(tot)
        f = float(tot.item())
        obj.append(f)
        loss.append(abs(f - obj0))
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        ev.synchronize()
        tv.append(ev0.elapsed_time(ev) * 1e-3)
        if loss[-1] < acc:
            it, hit = i, True
            break
    n_out = it if hit else len(obj)
    tvec = torch.tensor(tv[:n_out], dtype=f64, device=dev)
    if n_out:
        tvec = tvec - tvec[0].clone()
        torch.distributed.all_reduce(tvec, op=torch.distributed.ReduceOp.MAX)
    res = (np.asarray(obj[:n_out]), np.asarray(loss[:n_out]), it, tvec.cpu().numpy())
    if return_state:
        dist.all_reduce_sum_(inner)
        return res + ({"theta": theta[g0:g0 + nloc, :d].cpu().numpy().T,
                       "dual": lam[g0:g0 + nloc, :d].cpu().numpy().T, "g0": g0, "iteration": it,
                       "inner_steps": int(inner.item())},)
    return res


def logreg_gd_single(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep, return_steps=False,
                     **_kw):
    """``logReg_GD`` (logReg_GD.m:1-25) for one worker with caller-supplied linear terms, on the GPU:
    one launch of ``logistic_half_gd_kernel`` (the whole <= 100-step inner loop on chip).  The kernel
    assembles its frozen term from chain state, ``-lam_{n-1} + lam_n + rho(th_n - th_{n-1}) +
    rho(th_n - th_{n+1})``; a three-worker chain with ``rho = 1``, ``lam_0 = C1``, ``lam_1 = C2``,
    ``th_0 = x - term1``, ``th_2 = x - term2`` and the worker in the middle reproduces
    ``-C1 + C2 + term1 + term2`` exactly as the caller passed it."""
    dev = torch.device("cuda", torch.cuda.current_device())
    f64 = torch.float64
    Xh = np.asarray(X, dtype=np.float64)
    s, d = Xh.shape
    dp = nv.padded(d)
    Xd = torch.from_numpy(np.ascontiguousarray(Xh)).to(dev)
    yd = torch.from_numpy(np.asarray(Y, dtype=np.float64).reshape(-1).copy()).to(dev)

    def vec(v):
        out = torch.zeros((dp,), dtype=f64, device=dev)
        out[:d] = torch.from_numpy(np.asarray(v, dtype=np.float64).reshape(-1).copy()).to(dev)
        return out

    x0, c1, c2, t1, t2 = vec(x), vec(C1), vec(C2), vec(term1), vec(term2)
    theta = torch.stack([x0 - t1, x0, x0 - t2]).contiguous()
    lam = torch.stack([c1, c2, torch.zeros_like(c1)]).contiguous()
    ctrl = torch.tensor([1, 0, 0, 0], dtype=torch.int64, device=dev)
    inner = torch.zeros((1,), dtype=torch.int64, device=dev)
    p = nv.GadmmLogistic()
    p.X, p.y, p.theta, p.lam = Xd.data_ptr(), yd.data_ptr(), theta.data_ptr(), lam.data_ptr()
    p.inner_steps, p.ctrl = inner.data_ptr(), ctrl.data_ptr()
    p.rho, p.lam_log, p.gd_step, p.inner_tol = 1.0, float(lambda_logistic), float(gdStep), 1e-4
    p.N, p.s, p.d, p.dp, p.max_iter, p.max_inner = 3, int(s), int(d), dp, 1, 100
    p.n0, p.nloc = 1, 1
    nv.check(nv.lib().gadmm_logistic_half_gd(C.byref(p), 1, nv.stream_ptr()))
    torch.cuda.synchronize(dev)
    out = theta[1, :d].cpu().numpy()
    return (out, int(inner.item())) if return_steps else out


def opt_sol_logistic_gpu(XX, YY, dim, lam, num_workers, tol: float = 1e-12, return_theta=False):
    """``opt_sol_logistic`` (opt_sol_logistic.m:13-24) on the GPU: the consensus problem
    ``N lam/2 |x|^2 + sum log(1+exp(-y .* Xx))`` solved by Newton-CG on the one-pass products
    (all rows as one fp64 shard, so nothing is rounded).  Returns the optimal objective value."""
    from .logistic_scaled import ScaledLogisticGadmm
    from ..parallel.runtime import SoloDist
    XX = np.asarray(XX, dtype=np.float64)[:, : int(dim)]
    YY = np.asarray(YY, dtype=np.float64).reshape(-1)
    rows, d = XX.shape
    dpad = (d + 3) // 4 * 4
    X = np.zeros((1, rows, dpad), dtype=np.float64)
    X[0, :, :d] = XX
    eng = ScaledLogisticGadmm(1, dpad, rows, float(num_workers) * float(lam), 1.0, dist=SoloDist(),
                              data=(X, YY.reshape(1, rows)), store="f64", max_iter=1)
    obj0, th = eng.solve_optimum(tol=tol)
    return (obj0, th[:d].cpu().numpy()) if return_theta else obj0


def gadmm_logistic_newton(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                          lambda_logistic, acc, newton_tol=1e-10, max_newton=30, faithful=False,
                          naive=True, return_state=False, hessian_tc_min_d=128):
    """``group_ADMM_logistic`` with exact local solves by damped Newton on the GPU."""
    st = _LogisticState(XX, YY, no_workers, num_feature, per_samples, lambda_logistic)
    N, d, dp, dev = st.N, st.d, st.dp, st.dev
    f64 = torch.float64
    idx_all = torch.arange(N, device=dev)
    hasL = (idx_all > 0).to(f64).unsqueeze(1)
    hasR = (idx_all < N - 1).to(f64).unsqueeze(1)
    cvec = (hasL + hasR).squeeze(1)
    eye = torch.eye(dp, dtype=f64, device=dev)
    obj, loss, tv = [], [], []
    Iter = 100 if faithful else int(num_iter)
    hit = False
    ev0 = torch.cuda.Event(enable_timing=True)
    ev0.record()
    trial = torch.zeros_like(st.theta)
    for i in range(1, int(num_iter) + 1):
        for col in (0, 1):
            act = idx_all[col::2]
            if act.numel() == 0:
                continue
            th = st.theta
            left = torch.roll(th, 1, 0) * hasL
            right = torch.roll(th, -1, 0) * hasR
            lam_left = torch.roll(st.lam, 1, 0) * hasL
            lin = (-lam_left + st.lam * hasR)[act]
            centre = (rho * (left + right))[act]
            cr = (cvec[act] * rho).unsqueeze(1)
            x = th[act].clone()
            Xa = st.X[act]

            def phi_and_grad(xa, want_w):
                trial.copy_(st.theta)
                trial[act] = xa
                st.grad(trial, want_w)
                f = st.fw[act] + (lin * xa).sum(1) + 0.5 * cr.squeeze(1) * (xa * xa).sum(1) \
                    - (centre * xa).sum(1)
                g = st.Y[act] + lin + cr * xa - centre
                return f, g

            for _ in range(max_newton):
                f, g = phi_and_grad(x, True)
                gn = g.norm(dim=1)
                live = gn > newton_tol * torch.clamp(x.norm(dim=1), min=1.0)
                if not bool(live.any().item()):
                    break
                w = st.wout[act].contiguous()
                if d >= hessian_tc_min_d:
                    # X'DX on the tensor cores: the row scaling by sqrt(D) is fused into the
                    # split/transpose pass of the tcgen05 Gram (fp32-grade Hessian; Newton with an
                    # inexact Hessian still converges, the gradient stays fp64)
                    H = LA.gram_tc(Xa, dp, w=w)
                else:
                    Xw = Xa * torch.sqrt(w).unsqueeze(2)
                    H, _, _ = LA.gram(Xw, torch.zeros_like(w), dp)
                H += cr.unsqueeze(2) * eye
                Hinv = LA.shifted_inverse(H, d, st.lam_log)
                step = LA.BatchedSymv(LA.pack_sym(Hinv, d), dp)(g.contiguous())
                step = step * live.unsqueeze(1).to(f64)
                slope = (g * step).sum(1)
                t = torch.ones_like(gn)
                for _ls in range(40):
                    f_try, _ = phi_and_grad(x - t.unsqueeze(1) * step, False)
                    bad = (f_try > f - 1e-4 * t * slope) & live
                    if not bool(bad.any().item()):
                        break
                    t = torch.where(bad, 0.5 * t, t)
                x = x - t.unsqueeze(1) * step
            st.theta[act] = x
        st.lam[:N - 1] += rho * (st.theta[:N - 1] - st.theta[1:])
        st.grad(st.theta)
        fobj = float(st.fw.sum().item())
        obj.append(fobj)
        loss.append(abs(fobj - obj0))
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        ev.synchronize()
        tv.append(ev0.elapsed_time(ev) * 1e-3)
        if loss[-1] < acc:
            Iter, hit = i, True
            break
    n_out = Iter if hit else len(obj)
    tvec = np.asarray(tv[:n_out])
    res = (np.asarray(obj[:n_out]), np.asarray(loss[:n_out]), Iter, tvec - (tvec[0] if n_out else 0))
    if return_state:
        return res + ({"theta": st.theta[:, :d].cpu().numpy().T, "dual": st.lam[:, :d].cpu().numpy().T,
                       "iteration": Iter, "launches": st.launches},)
    return res
