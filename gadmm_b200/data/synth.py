"""Synthetic data of the reference's shapes (host side, float64).

``X_i = c^(i-1) * q_i q_i' + I`` with ``q_i`` the i-th column of the Q factor of a Gaussian
matrix (LinearRegression_Synthetic.m:23-35 with c=1.3; LogisticRegression_Synthetic.m:32 with
c=1).  The reference takes its targets/labels from data files that are not shipped
(``data29/y.txt``, ``data11/y.txt``); :func:`surrogate_targets` / :func:`surrogate_labels` stand
in for them.  Device-side generation for the scaled configs lives in
``csrc/synth.cu`` (Philox) and :mod:`gadmm_b200.engine.scaled`.
"""
from __future__ import annotations

import numpy as np


def _rng(rng):
    return rng if isinstance(rng, np.random.Generator) else np.random.default_rng(rng)


def surrogate_targets(s: int, rng=None) -> np.ndarray:
    """Body-Fat-like targets (percent body fat: mean ~19, sd ~8)."""
    return 19.0 + 8.0 * _rng(rng).standard_normal(s)


def surrogate_labels(s: int, rng=None, p_pos: float = 0.24) -> np.ndarray:
    """+-1 labels with the class balance of ``inputData.mat`` (288/1200 positives)."""
    return np.where(_rng(rng).random(s) < p_pos, 1.0, -1.0)


def hmax_linear(X_list) -> np.ndarray:
    """``Hmax(i)=max(eig(X{i}'*X{i}))`` (LinearRegression_Synthetic.m:33)."""
    return np.array([np.linalg.eigvalsh(x.T @ x)[-1] for x in X_list])


def hmax_logistic(X_list, lam: float) -> np.ndarray:
    """``Hmax(i)=0.25*max(abs(eig(X{i}'*X{i})))+lambda`` (LogisticRegression_Synthetic.m:43-46)."""
    return 0.25 * hmax_linear(X_list) + lam


def _rank_one_family(N, s, growth, rng):
    g = _rng(rng)
    if N > s:
        raise ValueError("the reference construction needs num_workers <= num_sample (Q columns)")
    Q, _ = np.linalg.qr(g.standard_normal((s, s)))
    eye = np.eye(s)
    return [growth ** i * np.outer(Q[:, i], Q[:, i]) + eye for i in range(N)]


def linreg_synthetic(N: int = 24, s: int = 50, growth: float = 1.3, rng=None, y=None):
    """Returns ``X (list), y (list), Hmax, X_fede, y_fede`` for the linear synthetic driver.
    Feature count equals ``s`` (the construction is square)."""
    g = _rng(rng)
    X = _rank_one_family(N, s, growth, g)
    ydata = surrogate_targets(s, g) if y is None else np.asarray(y, dtype=np.float64).reshape(-1)[:s]
    ys = [ydata.copy() for _ in range(N)]
    return X, ys, hmax_linear(X), np.concatenate(X, 0), np.concatenate(ys, 0)


def logreg_synthetic(N: int = 24, s: int = 50, lam: float = 1e-5, rng=None, y=None):
    """Logistic synthetic driver data (LogisticRegression_Synthetic.m:17-46)."""
    g = _rng(rng)
    X = _rank_one_family(N, s, 1.0, g)
    ydata = surrogate_labels(s, g) if y is None else np.asarray(y, dtype=np.float64).reshape(-1)[:s]
    ys = [ydata.copy() for _ in range(N)]
    return X, ys, hmax_logistic(X, lam), np.concatenate(X, 0), np.concatenate(ys, 0)


def split_dataset(Xdata, ydata, per_split: int = None, num_workers: int = None):
    """Contiguous equal splits: ``per_split`` samples per worker with ``N=floor(total/per_split)``
    (LinearRegression_Real.m:13-39) or ``num_workers`` workers with ``floor(n/N)`` samples each
    (LogisticRegression_real.m:17-34).  Trailing samples are dropped as in the reference."""
    Xdata = np.asarray(Xdata, dtype=np.float64)
    ydata = np.asarray(ydata, dtype=np.float64).reshape(-1)
    total = Xdata.shape[0]
    if per_split is None:
        per_split = total // int(num_workers)
    N = total // per_split if num_workers is None else int(num_workers)
    X = [Xdata[n * per_split:(n + 1) * per_split] for n in range(N)]
    y = [ydata[n * per_split:(n + 1) * per_split] for n in range(N)]
    return X, y, np.concatenate(X, 0), np.concatenate(y, 0), per_split
