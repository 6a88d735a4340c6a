"""Problem families: sharded least squares and l2-regularised logistic regression.

The reference has exactly two "models" (SURVEY.md section 2.2): per-worker least squares
``f_n(x) = 0.5*||X_n x - y_n||^2`` (group_ADMM_closedForm.m:96-101) and per-worker regularised
logistic loss ``f_n(x) = lam/2*||x||^2 + sum(log(1+exp(-y.*(X_n x))))``
(group_ADMM_logistic_GD.m:109-115).  Everything here is float64 NumPy; the GPU engine keeps its
own device-resident representation (:mod:`gadmm_b200.engine`).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


def _as_f64(a) -> np.ndarray:
    """Accept NumPy / torch / lists; MAT-files deliver int16 labels."""
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(a), dtype=np.float64)


def shard_rows(n: int, s: int) -> slice:
    """Rows of worker ``n`` (0-based) in the stacked matrix: ``first=(ii-1)*s2+1`` in the reference."""
    return slice(n * s, (n + 1) * s)


def stack_shards(X_list, y_list):
    """``X_fede=[X_fede;X{i}]`` (LinearRegression_Synthetic.m:44-49)."""
    XX = np.concatenate([_as_f64(x) for x in X_list], axis=0)
    YY = np.concatenate([_as_f64(y).reshape(-1) for y in y_list], axis=0)
    return XX, YY


def linear_objective(X: np.ndarray, y: np.ndarray, theta: np.ndarray) -> float:
    r = X @ theta - y
    return 0.5 * float(r @ r)


def logistic_objective(X, y, theta, lam, naive: bool = True) -> float:
    """``lam/2*||theta||^2 + sum(log(1+exp(-y.*(X*theta))))``.

    ``naive=True`` reproduces the reference's overflow-prone form (quirk 8);
    ``naive=False`` uses ``logaddexp`` and is identical wherever the naive form is finite.
    """
    z = -y * (X @ theta)
    if naive:
        with np.errstate(over="ignore"):
            v = np.log(1.0 + np.exp(z))
    else:
        v = np.logaddexp(0.0, z)
    return 0.5 * lam * float(theta @ theta) + float(np.sum(v))


def logistic_gradient(X, y, theta, lam, naive: bool = True) -> np.ndarray:
    """``-(X'*(y./(1+exp(y.*(X*theta))))) + lam*theta`` (logReg_GD.m:9)."""
    z = y * (X @ theta)
    if naive:
        with np.errstate(over="ignore"):
            w = y / (1.0 + np.exp(z))
    else:
        w = y * 0.5 * (1.0 - np.tanh(0.5 * z))
    return -(X.T @ w) + lam * theta


@dataclass
class LinearProblem:
    """Row-stacked least-squares data split over ``N`` workers with ``s`` samples each."""

    XX: np.ndarray
    YY: np.ndarray
    N: int
    d: int
    s: int
    G: np.ndarray = field(init=False, repr=False)   # (N, d, d) Gram matrices X_n' X_n
    b: np.ndarray = field(init=False, repr=False)   # (N, d)    X_n' y_n
    yy: np.ndarray = field(init=False, repr=False)  # (N,)      y_n' y_n

    def __post_init__(self):
        self.XX = _as_f64(self.XX)[: self.N * self.s, : self.d]
        self.YY = _as_f64(self.YY).reshape(-1)[: self.N * self.s]
        Xs = self.XX.reshape(self.N, self.s, self.d)
        ys = self.YY.reshape(self.N, self.s)
        self.G = np.einsum("nsi,nsj->nij", Xs, Xs)
        self.b = np.einsum("nsi,ns->ni", Xs, ys)
        self.yy = np.einsum("ns,ns->n", ys, ys)

    @property
    def Xs(self):
        return self.XX.reshape(self.N, self.s, self.d)

    @property
    def ys(self):
        return self.YY.reshape(self.N, self.s)

    def objective(self, theta: np.ndarray) -> float:
        """sum_n 0.5*||X_n theta_n - y_n||^2 with each worker's own column (theta is d x N)."""
        r = np.einsum("nsd,dn->ns", self.Xs, theta) - self.ys
        return 0.5 * float(np.einsum("ns,ns->", r, r))

    def consensus_objective(self, theta: np.ndarray) -> float:
        return linear_objective(self.XX, self.YY, theta)


@dataclass
class LogisticProblem:
    XX: np.ndarray
    YY: np.ndarray
    N: int
    d: int
    s: int
    lam: float
    naive: bool = True

    def __post_init__(self):
        self.XX = _as_f64(self.XX)[: self.N * self.s, : self.d]
        self.YY = _as_f64(self.YY).reshape(-1)[: self.N * self.s]

    @property
    def Xs(self):
        return self.XX.reshape(self.N, self.s, self.d)

    @property
    def ys(self):
        return self.YY.reshape(self.N, self.s)

    def objective(self, theta: np.ndarray) -> float:
        """sum_n lam/2||theta_n||^2 + sum log(1+exp(-y.*(X_n theta_n)))  (d x N input)."""
        z = -self.ys * np.einsum("nsd,dn->ns", self.Xs, theta)
        if self.naive:
            with np.errstate(over="ignore"):
                v = np.log(1.0 + np.exp(z))
        else:
            v = np.logaddexp(0.0, z)
        return 0.5 * self.lam * float(np.sum(theta * theta)) + float(np.sum(v))

    def consensus_objective(self, theta: np.ndarray) -> float:
        """``N*lam/2*||theta||^2 + sum(log(1+exp(-y.*(X theta))))`` (GD_DGD_LAG_logistic.m:108)."""
        return logistic_objective(self.XX, self.YY, theta, self.N * self.lam, self.naive)
