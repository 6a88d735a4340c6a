"""CPU float64 oracle: standard (star / parameter-server) ADMM, closed form.

Semantics of standared_ADMM.m:17-107 -- workers 1..N-1 solve against the server's previous
iterate, worker N is the server *and* owns a data shard, duals are per worker.
"""
from __future__ import annotations

import time

import numpy as np

from ..models.problems import LinearProblem
from .gadmm import _Timer, _finish, _spd_inverse


def standared_ADMM(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                   return_state: bool = False):
    """``[obj, loss, Iter, time] = standared_ADMM(...)`` (standared_ADMM.m:1-2)."""
    prob = LinearProblem(XX, YY, no_workers, num_feature, noSamples)
    N, d = prob.N, prob.d
    eye = np.eye(d)
    Mw = _spd_inverse(prob.G[: N - 1] + rho * eye) if N > 1 else np.zeros((0, d, d))
    Ms = _spd_inverse(prob.G[N - 1] + (N - 1) * rho * eye) if N > 1 else \
        np.linalg.pinv(prob.G[N - 1])
    lam = np.zeros((d, N))
    theta = np.zeros((d, N))
    obj, loss = [], []
    timer = _Timer()
    Iter = num_iter
    for i in range(1, num_iter + 1):
        z_prev = theta[:, N - 1].copy()
        for n in range(N - 1):
            t0 = time.perf_counter()
            theta[:, n] = Mw[n] @ (prob.b[n] - lam[:, n] + rho * z_prev)   # :42
            if n == 0:
                timer.add(i - 1, time.perf_counter() - t0)
        u = lam[:, : N - 1].sum(axis=1) + rho * theta[:, : N - 1].sum(axis=1)  # :66-71
        theta[:, N - 1] = Ms @ (prob.b[N - 1] + u)                           # :73
        lam[:, : N - 1] += rho * (theta[:, : N - 1] - theta[:, [N - 1]])     # :84-88
        f = prob.objective(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter = i
            break
    res = _finish(obj, loss, Iter, timer.vector(Iter))
    if return_state:
        return res + ({"theta": theta, "dual": lam, "iteration": Iter},)
    return res
