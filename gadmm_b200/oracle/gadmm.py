"""CPU float64 oracle for the GADMM family (chain topology, head/tail groups).

Straight restatement of SURVEY.md section 2.2, *not* of the MATLAB text:
  * edge-dual GADMM, closed form      -- group_ADMM_closedForm.m:13-108
  * node-dual GADMM on chain 1..N     -- static_group_ADMM_closedForm.m:28-160
  * D-GADMM (live / pre-generated)    -- dynamic_group_ADMM_closedForm.m:18-190, ..._v0.m:16-26
  * exact-solve GADMM (CVX semantics) -- group_ADMM.m:16-78, group_ADMM_logistic.m:18-105
  * inexact logistic GADMM            -- group_ADMM_logistic_GD.m:18-124, logReg_GD.m:3-25

The Gram matrices and the two shifted inverses ``(G_n + c*rho*I)^-1, c in {1,2}`` are cached (the
reference rebuilds them every iteration; the iterates are the same).  This module is both the
correctness oracle for every GPU kernel and the timed CPU baseline.
"""
from __future__ import annotations

import time

import numpy as np

from ..models.problems import LinearProblem, LogisticProblem, logistic_gradient
from ..topology import chain_tables, findPath2


def _spd_inverse(A: np.ndarray) -> np.ndarray:
    """Batched SPD inverse through Cholesky (MATLAB's ``\\`` picks Cholesky for SPD input)."""
    L = np.linalg.cholesky(A)
    Linv = np.linalg.inv(L)
    return np.swapaxes(Linv, -1, -2) @ Linv


class _Timer:
    """The reference's clock: ``time(i) = time(i-1) + 2*toc`` around worker 1's solve only
    (group_ADMM_closedForm.m:40-42,53-55); iteration 1 is never timed."""

    def __init__(self):
        self.t = [0.0]

    def add(self, i: int, dt: float):
        if i > 0:
            self.t.append(self.t[-1] + 2.0 * dt)

    def vector(self, n: int) -> np.ndarray:
        v = np.asarray(self.t[:n], dtype=np.float64)
        if v.shape[0] < n:  # worker 1 was not timed (e.g. N==0); pad flat
            v = np.concatenate([v, np.full(n - v.shape[0], v[-1])])
        return v


def _finish(obj, loss, it, *extra):
    obj = np.asarray(obj[:it], dtype=np.float64)
    loss = np.asarray(loss[:it], dtype=np.float64)
    return (obj, loss, it) + tuple(extra)


def _chain_solver(prob: LinearProblem, rho: float, num_iter: int, obj0: float, acc: float,
                  path_fn, cost_fn=None, state=None, theta0=None):
    """Node-dual GADMM on an arbitrary (possibly changing) chain.

    ``path_fn(i)`` returns the 0-based chain to use in iteration ``i`` (1-based, as the reference
    counts) or ``None`` to keep the previous one; ``cost_fn(i)`` returns the per-iteration energy.
    """
    N, d = prob.N, prob.d
    eye = np.eye(d)
    Minv = {c: _spd_inverse(prob.G + c * rho * eye) for c in (1, 2)}
    theta = np.zeros((d, N)) if theta0 is None else np.array(theta0, dtype=np.float64, copy=True)
    nu = np.zeros((d, N))
    if state is not None:
        theta = np.array(state["theta"], dtype=np.float64, copy=True)
        nu = np.array(state["dual"], dtype=np.float64, copy=True)
    obj, loss, com = [], [], []
    timer = _Timer()
    Iter = num_iter
    path = None
    for i in range(1, num_iter + 1):
        newp = path_fn(i)
        if newp is not None:
            path = np.asarray(newp, dtype=np.int64)
            left, right, color, cmul = chain_tables(path)
        if cost_fn is not None:
            c_i = float(cost_fn(i))
            com.append(c_i if i == 1 else com[-1] + c_i)
        for col in (0, 1):
            for n in path[col::2]:
                t0 = time.perf_counter()
                r = prob.b[n] - nu[:, n]
                if left[n] >= 0:
                    r = r + rho * theta[:, left[n]]
                if right[n] >= 0:
                    r = r + rho * theta[:, right[n]]
                theta[:, n] = Minv[int(cmul[n])][n] @ r
                if n == 0 and col == 0:
                    timer.add(i - 1, time.perf_counter() - t0)
        for n in range(N):
            upd = 0.0
            if right[n] >= 0:
                upd = upd + rho * (theta[:, n] - theta[:, right[n]])
            if left[n] >= 0:
                upd = upd - rho * (theta[:, left[n]] - theta[:, n])
            nu[:, n] = nu[:, n] + upd
        f = prob.objective(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter = i
            break
    out_state = {"theta": theta, "dual": nu, "iteration": Iter, "path": path + 1}
    return obj, loss, Iter, timer.vector(Iter), np.asarray(com[:Iter]), out_state


def group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                          return_state: bool = False):
    """Edge-dual closed-form GADMM on the natural chain (group_ADMM_closedForm.m:1-117).

    Returns ``(obj_GADMM, loss_GADMM, Iter, gadmm_time)``.
    """
    prob = LinearProblem(XX, YY, no_workers, num_feature, noSamples)
    N, d = prob.N, prob.d
    eye = np.eye(d)
    Minv = {c: _spd_inverse(prob.G + c * rho * eye) for c in (1, 2)}
    lam = np.zeros((d, N))       # lam[:, n] couples workers n and n+1 (0-based)
    theta = np.zeros((d, N))
    obj, loss = [], []
    timer = _Timer()
    Iter = num_iter
    for i in range(1, num_iter + 1):
        for start in (0, 1):
            for n in range(start, N, 2):
                t0 = time.perf_counter()
                r = prob.b[n].copy()
                if n > 0:
                    r += lam[:, n - 1] + rho * theta[:, n - 1]
                if n < N - 1:
                    r += -lam[:, n] + rho * theta[:, n + 1]
                c = 1 if (n == 0 or n == N - 1) else 2
                theta[:, n] = Minv[c][n] @ r
                if n == 0:
                    timer.add(i - 1, time.perf_counter() - t0)
        lam[:, : N - 1] += rho * (theta[:, : N - 1] - theta[:, 1:])
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


def group_ADMM(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc):
    """GADMM with the local problem solved *exactly* (CVX in the reference, group_ADMM.m:16-78).

    For a quadratic local problem the exact minimiser is the closed-form update, so the iterates
    coincide with :func:`group_ADMM_closedForm` to solver precision.  Returns
    3 outputs -- the reference has **no** time vector here (group_ADMM.m:1).
    """
    obj, loss, it, _ = group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples,
                                             num_iter, obj0, acc)
    return obj, loss, it


def static_group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                                 acc, coherence_Time, pathCost_static_matrix, state=None,
                                 return_state: bool = False):
    """Node-dual GADMM on chain 1..N plus the energy meter (static_group_ADMM_closedForm.m)."""
    prob = LinearProblem(XX, YY, no_workers, num_feature, noSamples)
    costs = np.atleast_2d(np.asarray(pathCost_static_matrix, dtype=np.float64))
    kk = [0]
    cur = [float(np.sum(costs[0]))]

    def path_fn(i):
        return np.arange(no_workers) if i == 1 else None

    def cost_fn(i):
        if i > 1 and i % coherence_Time == 0:
            kk[0] += 1
            cur[0] = float(np.sum(costs[kk[0]]))
        return cur[0]

    obj, loss, it, tvec, com, st = _chain_solver(prob, rho, num_iter, obj0, acc, path_fn, cost_fn,
                                                 state)
    res = _finish(obj, loss, it, tvec, com)
    return res + (st,) if return_state else res


def dynamic_group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                                  acc, path, pathCost, coherence_Time, rng=None, state=None,
                                  return_state: bool = False, rechain=None):
    """D-GADMM, chain regenerated by ``findPath2`` every ``coherence_Time`` iterations
    (dynamic_group_ADMM_closedForm.m:18-21).  ``rechain(i) -> (path1, pathCost)`` overrides the
    generator (tests feed identical chains to the oracle and to the GPU engine)."""
    prob = LinearProblem(XX, YY, no_workers, num_feature, noSamples)
    g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng
    cur_cost = [float(np.sum(pathCost))]
    first = np.asarray(path, dtype=np.int64).ravel() - 1

    def path_fn(i):
        if i == 1:
            return first
        if i % coherence_Time == 0:
            if rechain is not None:
                p1, pc = rechain(i)
            else:
                p1, pc, _, _, _ = findPath2(no_workers, g)
            cur_cost[0] = float(np.sum(pc))
            return np.asarray(p1, dtype=np.int64) - 1
        return None

    obj, loss, it, tvec, com, st = _chain_solver(prob, rho, num_iter, obj0, acc, path_fn,
                                                 lambda i: cur_cost[0], state)
    res = _finish(obj, loss, it, tvec, com)
    return res + (st,) if return_state else res


def dynamic_group_ADMM_closedForm_v0(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                     obj0, acc, path_matrix, pathCost_matrix, coherence_Time,
                                     faithful: bool = True, state=None, return_state: bool = False):
    """D-GADMM with pre-generated chains (dynamic_group_ADMM_closedForm_v0.m:16-26).

    ``faithful=True`` keeps quirk 4: the first window's cost is the matrix *column*
    ``pathCost_matrix(:,1)`` while refreshes use rows.
    """
    prob = LinearProblem(XX, YY, no_workers, num_feature, noSamples)
    paths = np.atleast_2d(np.asarray(path_matrix, dtype=np.int64))
    costs = np.atleast_2d(np.asarray(pathCost_matrix, dtype=np.float64))
    kk = [0]
    cur = [float(np.sum(costs[:, 0])) if faithful else float(np.sum(costs[0]))]

    def path_fn(i):
        if i == 1:
            return paths[0] - 1
        if i % coherence_Time == 0:
            kk[0] += 1
            cur[0] = float(np.sum(costs[kk[0]]))
            return paths[kk[0]] - 1
        return None

    obj, loss, it, tvec, com, st = _chain_solver(prob, rho, num_iter, obj0, acc, path_fn,
                                                 lambda i: cur[0], state)
    res = _finish(obj, loss, it, tvec, com)
    return res + (st,) if return_state else res


# --------------------------------------------------------------------------------------------
# logistic regression
# --------------------------------------------------------------------------------------------

def logReg_GD(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep, naive: bool = True,
              max_steps: int = 100, tol: float = 1e-4, return_steps: bool = False):
    """``x = logReg_GD(X,Y,x,C1,C2,term1,term2,lambda_logistic,gdStep)`` (logReg_GD.m:1-25):
    at most 100 fixed-step gradient steps with the penalty terms frozen; exit when **all**
    ``abs(dx) < 1e-4``."""
    x = np.array(x, dtype=np.float64, copy=True).reshape(-1)
    const = -np.asarray(C1).reshape(-1) + np.asarray(C2).reshape(-1) \
        + np.asarray(term1).reshape(-1) + np.asarray(term2).reshape(-1)
    steps = 0
    for _ in range(max_steps):
        g = logistic_gradient(X, Y, x, lambda_logistic, naive) + const
        x_prev = x
        x = x - gdStep * g
        steps += 1
        if np.all(np.abs(x - x_prev) < tol):
            break
    return (x, steps) if return_steps else x


def group_ADMM_logistic_GD(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                           lambda_logistic, acc, gdStep, naive: bool = True,
                           return_state: bool = False):
    """Inexact logistic GADMM (group_ADMM_logistic_GD.m:1-132)."""
    prob = LogisticProblem(XX, YY, no_workers, num_feature, per_samples, lambda_logistic, naive)
    N, d = prob.N, prob.d
    Xs, ys = prob.Xs, prob.ys
    lam = np.zeros((d, N))
    theta = np.zeros((d, N))
    z = np.zeros(d)
    obj, loss = [], []
    timer = _Timer()
    Iter = num_iter
    inner = 0
    for i in range(1, num_iter + 1):
        for start in (0, 1):
            for n in range(start, N, 2):
                C1 = lam[:, n - 1] if n > 0 else z
                t1 = rho * (theta[:, n] - theta[:, n - 1]) if n > 0 else z
                C2 = lam[:, n] if n < N - 1 else z
                t2 = rho * (theta[:, n] - theta[:, n + 1]) if n < N - 1 else z
                t0 = time.perf_counter()
                theta[:, n], st = logReg_GD(Xs[n], ys[n], theta[:, n], C1, C2, t1, t2,
                                            lambda_logistic, gdStep, naive, return_steps=True)
                inner += st
                if n == 0:
                    timer.add(i - 1, time.perf_counter() - t0)
        lam[:, : N - 1] += rho * (theta[:, : N - 1] - theta[:, 1:])
        f = prob.objective(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter = i
            break
    res = _finish(obj, loss, Iter, timer.vector(Iter))
    if return_state:
        return res + ({"theta": theta, "dual": lam, "iteration": Iter, "inner_steps": inner},)
    return res


def _logistic_local_newton(X, y, x0, lin, rho_c, centre, lam, naive, tol=1e-10, max_newton=50):
    """Exact minimiser of ``lam/2||x||^2 + sum log(1+exp(-y.*(Xx))) + lin'x + rho_c/2||x||^2
    - rho*centre'x`` by damped Newton; this is what CVX solves in group_ADMM_logistic.m:48-55."""
    x = x0.copy()
    d = x.shape[0]

    def fval(v):
        zz = -y * (X @ v)
        return (0.5 * (lam + rho_c) * float(v @ v) + float(np.sum(np.logaddexp(0.0, zz)))
                + float(lin @ v) - float(centre @ v))

    for _ in range(max_newton):
        z = y * (X @ x)
        sig = 0.5 * (1.0 - np.tanh(0.5 * z))          # 1/(1+exp(z))
        g = -(X.T @ (y * sig)) + (lam + rho_c) * x + lin - centre
        if np.linalg.norm(g) <= tol * max(1.0, np.linalg.norm(x)):
            break
        w = sig * (1.0 - sig)
        H = (X.T * w) @ X + (lam + rho_c) * np.eye(d)
        step = np.linalg.solve(H, g)
        t, f0, slope = 1.0, fval(x), float(g @ step)
        while t > 1e-8 and fval(x - t * step) > f0 - 1e-4 * t * slope:
            t *= 0.5
        x = x - t * step
    return x


def group_ADMM_logistic(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                        lambda_logistic, acc, faithful: bool = False, naive: bool = True,
                        return_state: bool = False):
    """GADMM with exact logistic local solves (group_ADMM_logistic.m:1-123).

    The reference calls CVX; here the local strongly-convex problem is solved by damped Newton to
    1e-10 (CVX default precision is ~1.5e-8).  ``faithful=True`` keeps
    quirk 2: ``Iter`` defaults to 100 instead of ``num_iter`` when the tolerance is not reached
    (group_ADMM_logistic.m:5).
    """
    prob = LogisticProblem(XX, YY, no_workers, num_feature, per_samples, lambda_logistic, naive)
    N, d = prob.N, prob.d
    Xs, ys = prob.Xs, prob.ys
    lam = np.zeros((d, N))
    theta = np.zeros((d, N))
    obj, loss = [], []
    timer = _Timer()
    Iter = 100 if faithful else num_iter
    hit = False
    for i in range(1, num_iter + 1):
        for start in (0, 1):
            for n in range(start, N, 2):
                lin = np.zeros(d)
                centre = np.zeros(d)
                c = 0
                if n > 0:
                    lin -= lam[:, n - 1]
                    centre += rho * theta[:, n - 1]
                    c += 1
                if n < N - 1:
                    lin += lam[:, n]
                    centre += rho * theta[:, n + 1]
                    c += 1
                t0 = time.perf_counter()
                theta[:, n] = _logistic_local_newton(Xs[n], ys[n], theta[:, n], lin, c * rho,
                                                     centre, lambda_logistic, naive)
                if n == 0:
                    timer.add(i - 1, time.perf_counter() - t0)
        lam[:, : N - 1] += rho * (theta[:, : N - 1] - theta[:, 1:])
        f = prob.objective(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        if loss[-1] < acc:
            Iter = i
            hit = True
            break
    n_out = len(obj) if not hit else Iter
    res = (np.asarray(obj[:n_out]), np.asarray(loss[:n_out]), Iter, timer.vector(n_out))
    if return_state:
        return res + ({"theta": theta, "dual": lam, "iteration": Iter},)
    return res
