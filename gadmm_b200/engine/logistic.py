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
                      lambda_logistic, acc, gdStep, naive=True, window=16, return_state=False):
    """``group_ADMM_logistic_GD`` on the GPU: ``(obj, loss, Iter, time)``."""
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


def logreg_gd_single(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep, **kw):
    """``logReg_GD`` for one worker with caller-supplied linear terms.  The GPU kernel
    (``logistic_half_gd_kernel``) derives those terms from the chain state itself, so this
    stand-alone helper (a single 50x50 solve, launch-bound on a GPU) evaluates on the host with the
    oracle's implementation; the batched in-chain solve is :func:`gadmm_logistic_gd`."""
    from ..oracle import logReg_GD
    return logReg_GD(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep, **kw)


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
