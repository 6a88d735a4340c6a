"""Global optima (objective *values*, like the reference's opt_sol*.m)."""
from __future__ import annotations

import numpy as np

from ..models.problems import _as_f64, logistic_objective


def opt_sol_closedForm(XX, YY, return_theta: bool = False):
    """``x=(XX'*XX)\\(XX'*YY); out=0.5*sum_square(XX*x-YY)`` (opt_sol_closedForm.m:2-3)."""
    XX = _as_f64(XX)
    YY = _as_f64(YY).reshape(-1)
    G = XX.T @ XX
    b = XX.T @ YY
    try:
        x = np.linalg.solve(G, b)
    except np.linalg.LinAlgError:
        x = np.linalg.lstsq(XX, YY, rcond=None)[0]
    r = XX @ x - YY
    out = 0.5 * float(r @ r)
    return (out, x) if return_theta else out


def opt_sol(XX, YY, size=None, return_theta: bool = False):
    """CVX least-squares optimum (opt_sol.m:3-9): the same minimiser as the normal equations; here
    obtained by an orthogonal (QR/SVD) least-squares solve, which is what an interior-point
    solver converges to at tight precision."""
    XX = _as_f64(XX)
    YY = _as_f64(YY).reshape(-1)
    if size is not None:
        XX = XX[:, : int(size)]
    x = np.linalg.lstsq(XX, YY, rcond=None)[0]
    r = XX @ x - YY
    out = 0.5 * float(r @ r)
    return (out, x) if return_theta else out


def opt_sol_logistic(XX, YY, dim, lam, num_workers, tol: float = 1e-12, max_newton: int = 100,
                     return_theta: bool = False):
    """Optimum of ``N*lam/2||x||^2 + sum(log(1+exp(-y.*(Xx))))`` (opt_sol_logistic.m:13-24) by
    damped Newton (CVX in the reference; "exact" = <=1e-8 relative)."""
    XX = _as_f64(XX)[:, : int(dim)]
    YY = _as_f64(YY).reshape(-1)
    mu = num_workers * lam
    x = np.zeros(XX.shape[1])

    def fval(v):
        return logistic_objective(XX, YY, v, mu, naive=False)

    for _ in range(max_newton):
        z = YY * (XX @ x)
        sig = 0.5 * (1.0 - np.tanh(0.5 * z))
        g = -(XX.T @ (YY * sig)) + mu * x
        if np.linalg.norm(g) <= tol * max(1.0, np.linalg.norm(x)):
            break
        w = sig * (1.0 - sig)
        H = (XX.T * w) @ XX + mu * np.eye(XX.shape[1])
        step = np.linalg.solve(H, g)
        t, f0, slope = 1.0, fval(x), float(g @ step)
        while t > 1e-10 and fval(x - t * step) > f0 - 1e-4 * t * slope:
            t *= 0.5
        x = x - t * step
    out = fval(x)
    return (out, x) if return_theta else out
