"""CPU float64 oracle: distributed dual averaging on the chain (dual_averaging*.m).

The reference overwrites ``Z_prev(:,n)`` inside the worker loop (dual_averaging.m:44), so worker
n sees the *new* Z of n-1 and the *old* Z of n+1 (Gauss-Seidel).  ``mode='faithful'`` keeps that;
``mode='jacobi'`` is the parallel-friendly variant (flagged as deviating).
"""
from __future__ import annotations

import time

import numpy as np

from ..models.problems import LinearProblem, LogisticProblem, logistic_gradient
from .gadmm import _Timer, _finish


def _dual_avg(N, d, grad_fn, obj_fn, num_iter, obj0, acc, alpha, mode, grads_all=None):
    """``grads_all(theta) -> (N, d)``, when given, supplies every worker's gradient at the start of
    the sweep instead of ``grad_fn(n, theta_n)`` inside it -- the same values, since worker n's
    model is not touched before its own step (used by the distributed backends)."""
    theta = np.zeros((d, N))
    Zp = np.zeros((d, N))
    obj, loss = [], []
    timer = _Timer()
    Iter = num_iter
    for i in range(1, num_iter + 1):
        src = Zp if mode == "faithful" else Zp.copy()
        Gall = grads_all(theta) if grads_all is not None else None
        for n in range(N):
            t0 = time.perf_counter()
            g = Gall[n] if Gall is not None else grad_fn(n, theta[:, n])
            if N == 1:
                z = g.copy()
            elif n == 0:
                z = src[:, 1] + g
            elif n == N - 1:
                z = src[:, n - 1] + g
            else:
                z = 0.5 * src[:, n + 1] + 0.5 * src[:, n - 1] + g
            theta[:, n] = -alpha * z
            Zp[:, n] = z
            if n == 0:
                timer.add(i - 1, time.perf_counter() - t0)
        f = obj_fn(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter = i
            break
    return _finish(obj, loss, Iter, timer.vector(Iter)), {"theta": theta, "dual": Zp,
                                                          "iteration": Iter}


def dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc, alpha,
                   mode: str = "faithful", return_state: bool = False):
    """``[obj, loss, Iter, time] = dual_averaging(...)`` (dual_averaging.m:1-75)."""
    prob = LinearProblem(XX, YY, no_workers, num_feature, noSamples)
    res, st = _dual_avg(prob.N, prob.d, lambda n, x: prob.G[n] @ x - prob.b[n], prob.objective,
                        num_iter, obj0, acc, alpha, mode)
    return res + (st,) if return_state else res


def dual_averaging_logisticReg(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                               alpha, eta, mode: str = "faithful", naive: bool = True,
                               return_state: bool = False):
    """Logistic variant (dual_averaging_logisticReg.m:1-80): gradient
    ``-(H'*(Y./(1+exp(Y.*(H*x))))) + eta*x``, objective ``eta/2||x||^2 + sum log(1+exp(.))``."""
    prob = LogisticProblem(XX, YY, no_workers, num_feature, noSamples, eta, naive)
    Xs, ys = prob.Xs, prob.ys
    res, st = _dual_avg(prob.N, prob.d,
                        lambda n, x: logistic_gradient(Xs[n], ys[n], x, eta, naive),
                        prob.objective, num_iter, obj0, acc, alpha, mode)
    return res + (st,) if return_state else res
