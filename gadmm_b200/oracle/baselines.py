"""CPU float64 oracle: the six first-order baselines of GD_DGD_LAG.m / GD_DGD_LAG_logistic.m.

GD (parameter server), DGD (chain, gradient averaging), LAG-PS, LAG-WK, cyclic IAG and
non-uniform randomised IAG, for least squares and l2-regularised logistic regression.
Reference line numbers are given per block; the quirks of SURVEY.md section 2.2 (initial
gradients of ones, iteration-1 guards, worker-1 timing refresh in LAG-PS, idle +1 of ``comm_iter``
every 1000 iterations) are reproduced.  The per-worker loops are vectorised over workers -- each
worker's LAG decision depends only on the shared iterate and its own state, so the results are
identical to the sequential loops.
"""
from __future__ import annotations

import time

import numpy as np

from ..models.problems import LinearProblem, LogisticProblem, _as_f64

TRIGGERSLOT = 10  # GD_DGD_LAG.m:18


class _Family:
    """Gradient/objective callbacks of one problem family."""

    def __init__(self, prob, logistic: bool):
        self.prob = prob
        self.logistic = logistic
        self.N, self.d = prob.N, prob.d
        if logistic:
            self.Xs, self.ys, self.lam = prob.Xs, prob.ys, prob.lam
        else:
            self.Gsum = prob.G.sum(axis=0)
            self.bsum = prob.b.sum(axis=0)

    # gradient of every worker at a *shared* iterate -> (N, d)
    def grads_shared(self, th):
        if not self.logistic:
            return self.prob.G @ th - self.prob.b
        z = self.ys * (self.Xs @ th)
        with np.errstate(over="ignore"):
            w = self.ys / (1.0 + np.exp(z)) if self.prob.naive else \
                self.ys * 0.5 * (1.0 - np.tanh(0.5 * z))
        return -np.einsum("nsd,ns->nd", self.Xs, w) + self.lam * th

    # gradient of every worker at its *own* iterate (Theta is d x N) -> (N, d)
    def grads_own(self, Th):
        if not self.logistic:
            return np.einsum("nij,jn->ni", self.prob.G, Th) - self.prob.b
        z = self.ys * np.einsum("nsd,dn->ns", self.Xs, Th)
        with np.errstate(over="ignore"):
            w = self.ys / (1.0 + np.exp(z)) if self.prob.naive else \
                self.ys * 0.5 * (1.0 - np.tanh(0.5 * z))
        return -np.einsum("nsd,ns->nd", self.Xs, w) + self.lam * Th.T

    def grad_one(self, n, th):
        if not self.logistic:
            return self.prob.G[n] @ th - self.prob.b[n]
        z = self.ys[n] * (self.Xs[n] @ th)
        with np.errstate(over="ignore"):
            w = self.ys[n] / (1.0 + np.exp(z)) if self.prob.naive else \
                self.ys[n] * 0.5 * (1.0 - np.tanh(0.5 * z))
        return -(self.Xs[n].T @ w) + self.lam * th

    def full_grad(self, th):
        if not self.logistic:
            return self.Gsum @ th - self.bsum
        return self.grads_shared(th).sum(axis=0)

    def obj_shared(self, th):
        return self.prob.consensus_objective(th)

    def obj_own(self, Th):
        return self.prob.objective(Th)


def _lag_like(fam, kind, alpha, thr, Hmax, num_iter, obj0, accuracy, rng=None, nonprob=None):
    """LAG-PS / LAG-WK / cyclic IAG / R-IAG share the server loop ``theta -= step*sum(grads)``."""
    N, d = fam.N, fam.d
    theta = np.zeros(d)
    grads = np.ones((N, d))
    theta_hat = np.zeros((N, d))
    dnorm = np.zeros(TRIGGERSLOT)       # ring of ||theta^{k-j+1}-theta^{k-j}||^2
    comm_iter = 1
    obj, loss, comm, tvec = [], [], [], [0.0]
    gt = tt = 0.0
    step = alpha if kind in ("lag_ps", "lag_wk") else alpha / N
    for it in range(1, num_iter + 1):
        flag = False
        if kind == "lag_ps":
            if it > 1:          # timing refresh of worker 1 (GD_DGD_LAG.m:204-209)
                t0 = time.perf_counter()
                grads[0] = fam.grad_one(0, theta)
                gt += time.perf_counter() - t0
            if it > TRIGGERSLOT:
                trig = dnorm.sum()
                lhs = Hmax ** 2 * np.sum((theta_hat - theta) ** 2, axis=1)
                up = lhs > thr * trig
                if up.any():
                    fresh = fam.grads_shared(theta)
                    grads[up] = fresh[up]
                    theta_hat[up] = theta
                    comm_iter += int(up.sum())
                    flag = True
        elif kind == "lag_wk":
            t0 = time.perf_counter()
            fresh = fam.grads_shared(theta)
            if it > 1:
                gt += (time.perf_counter() - t0) / N
            if it > TRIGGERSLOT:
                trig = dnorm.sum()
                up = np.sum((fresh - grads) ** 2, axis=1) > thr * trig
                if up.any():
                    grads[up] = fresh[up]
                    comm_iter += int(up.sum())
                    flag = True
        elif kind == "iag_cyclic":
            if it > 1:
                i = it % N      # i=mod(iter,N)+1, 1-based (GD_DGD_LAG.m:337)
                grads[i] = fam.grad_one(i, theta)
        elif kind == "iag_random":
            workprob = rng.random()
            i = int(np.searchsorted(np.cumsum(nonprob), workprob, side="left"))
            i = min(i, N - 1)
            if it > 1:
                grads[i] = fam.grad_one(i, theta)
        f = fam.obj_shared(theta)
        obj.append(f)
        loss.append(abs(f - obj0))
        t0 = time.perf_counter()
        new = theta - step * grads.sum(axis=0)
        if it > 1:
            tt += time.perf_counter() - t0
        dnorm[it % TRIGGERSLOT] = float(np.sum((new - theta) ** 2))
        theta = new
        if kind in ("lag_ps", "lag_wk"):
            if not flag and it % 1000 == 0:
                comm_iter += 1   # idle bump (GD_DGD_LAG.m:246-251)
            comm.append(comm_iter)
            if it > 1:
                tvec.append(tt + gt)
        if accuracy is not None and loss[-1] < accuracy:
            break
    out = [np.asarray(obj), np.asarray(loss)]
    if kind in ("lag_ps", "lag_wk"):
        out += [np.asarray(tvec[: len(obj)]), np.asarray(comm, dtype=np.float64)]
    return out


def _gd(fam, alpha, num_iter, logistic):
    d = fam.d
    theta = np.zeros(d)
    g = np.ones(d)
    obj, tvec = [], [0.0]
    gt = tt = 0.0
    for it in range(1, num_iter + 1):
        t0 = time.perf_counter()
        g1 = fam.grad_one(0, theta)
        if it > 1:
            gt += time.perf_counter() - t0
        if logistic:
            g = g1           # iteration 1 uses worker 1's gradient only (GD_DGD_LAG_logistic.m:98)
        if it > 1:
            g = fam.full_grad(theta)
            if logistic:
                # server adds N*lam*theta; grads_shared already carries lam*theta per worker
                pass
        obj.append(fam.obj_shared(theta))
        t0 = time.perf_counter()
        theta = theta - alpha * g
        if it > 1:
            tt += time.perf_counter() - t0
            tvec.append(tt + gt)
    return np.asarray(obj), np.asarray(tvec[: len(obj)]), theta


def _dgd(fam, alpha_d, num_iter, obj0):
    N, d = fam.N, fam.d
    Th = np.zeros((d, N))
    G = np.ones((N, d))
    obj, loss, tvec = [], [], []
    gt = tt = 0.0
    for it in range(1, num_iter + 1):
        if it > 1:
            t0 = time.perf_counter()
            G = fam.grads_own(Th)
            gt += (time.perf_counter() - t0) / N
        f = fam.obj_own(Th)
        obj.append(f)
        loss.append(abs(f - obj0))
        t0 = time.perf_counter()
        if N == 1:
            avg = G
        else:
            avg = np.empty_like(G)
            avg[0] = 0.5 * (G[0] + G[1])
            avg[-1] = 0.5 * (G[-1] + G[-2])
            avg[1:-1] = (G[1:-1] + G[2:] + G[:-2]) / 3.0
        Th = Th - alpha_d * avg.T
        tt += (time.perf_counter() - t0) / N
        tvec.append(tt + gt)
    return np.asarray(obj), np.asarray(loss), np.asarray(tvec)


def _split(X, y):
    Xl = [_as_f64(x) for x in X]
    yl = [_as_f64(v).reshape(-1) for v in y]
    return Xl, yl


def GD_DGD_LAG(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede, y_fede, num_sample,
               rng=None):
    """19-output linear-regression baselines (GD_DGD_LAG.m:1-4).

    ``X``/``y`` are per-worker lists (MATLAB cell arrays); only the first ``num_workers`` entries
    are used (quirk 1: the reference allocates an N x N cell).
    """
    prob = LinearProblem(X_fede, y_fede, num_workers, num_feature, num_sample)
    fam = _Family(prob, logistic=False)
    Hmax = _as_f64(Hmax).reshape(-1)[:num_workers]
    Hmaxall = float(np.linalg.eigvalsh(fam.Gsum)[-1])          # :19
    alpha = 1.0 / Hmaxall                                        # :46
    thrd = 10.0 / (alpha ** 2 * num_workers ** 2) / TRIGGERSLOT  # :47
    thrd5 = 1.0 / (alpha ** 2 * num_workers ** 2) / TRIGGERSLOT  # :65
    nonprob = Hmax / Hmax.sum()                                  # :8
    g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng

    obj_GD, time_GD, _ = _gd(fam, alpha, num_iter, logistic=False)
    loss_GD = np.abs(obj_GD - obj0)
    obj_DGD, loss_DGD, time_DGD = _dgd(fam, alpha / 100.0, num_iter, obj0)       # :127
    o_ps, l_ps, t_ps, c_ps = _lag_like(fam, "lag_ps", alpha, thrd, Hmax, num_iter, obj0, None)
    o_wk, l_wk, t_wk, c_wk = _lag_like(fam, "lag_wk", alpha, thrd5, Hmax, num_iter, obj0, None)
    o_c, l_c = _lag_like(fam, "iag_cyclic", alpha, 0.0, Hmax, num_iter, obj0, None)
    o_r, l_r = _lag_like(fam, "iag_random", alpha, 0.0, Hmax, num_iter, obj0, None, g, nonprob)
    return (obj_GD, loss_GD, time_GD, obj_DGD, loss_DGD, time_DGD,
            o_ps, l_ps, t_ps, c_ps, o_wk, l_wk, t_wk, c_wk, o_c, l_c, o_r, l_r, alpha)


def GD_DGD_LAG_logistic(X, y, num_workers, num_feature, Hmax, num_iter, X_fede, y_fede,
                        num_sample, lam, accuracy, rng=None, naive: bool = True):
    """20-output logistic baselines (GD_DGD_LAG_logistic.m:1-5).  There is **no** ``obj0`` input:
    the optimum is *defined* as GD's last objective and returned as ``obj1`` (``:131-133``)."""
    prob = LogisticProblem(X_fede, y_fede, num_workers, num_feature, num_sample, lam, naive)
    fam = _Family(prob, logistic=True)
    Hmax = _as_f64(Hmax).reshape(-1)[:num_workers]
    XX = prob.XX
    Hmaxall = 0.25 * float(np.linalg.eigvalsh(XX.T @ XX)[-1]) + lam   # :24
    alpha = 1.0 / Hmaxall
    thrd = 10.0 / (alpha ** 2 * num_workers ** 2) / TRIGGERSLOT
    thrd5 = 1.0 / (alpha ** 2 * num_workers ** 2) / TRIGGERSLOT
    nonprob = Hmax / Hmax.sum()
    g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng

    obj_GD, time_GD, _ = _gd(fam, alpha, num_iter, logistic=True)
    obj1 = float(obj_GD[-1])
    loss_GD = np.abs(obj_GD - obj1)
    obj_DGD, loss_DGD, time_DGD = _dgd(fam, alpha / 100.0, num_iter, obj1)
    o_ps, l_ps, t_ps, c_ps = _lag_like(fam, "lag_ps", alpha, thrd, Hmax, num_iter, obj1, accuracy)
    o_wk, l_wk, t_wk, c_wk = _lag_like(fam, "lag_wk", alpha, thrd5, Hmax, num_iter, obj1, accuracy)
    o_c, l_c = _lag_like(fam, "iag_cyclic", alpha, 0.0, Hmax, num_iter, obj1, accuracy)
    o_r, l_r = _lag_like(fam, "iag_random", alpha, 0.0, Hmax, num_iter, obj1, accuracy, g, nonprob)
    return (obj_GD, loss_GD, time_GD, obj_DGD, loss_DGD, time_DGD,
            o_ps, l_ps, t_ps, c_ps, o_wk, l_wk, t_wk, c_wk, o_c, l_c, o_r, l_r, alpha, obj1)
