#!/usr/bin/env python
"""Quick start: the reference's LinearRegression_Synthetic flow in a dozen lines.

Builds the 24-worker synthetic least-squares problem of LinearRegression_Synthetic.m:17-35
(shards ``X_i = 1.3^(i-1) q_i q_i' + I``), computes the optimum, and runs chain GADMM, D-GADMM
and standard (star) ADMM through the public API.  ``--backend auto`` uses the sm_100a kernels when
a GPU is present and the CPU oracle otherwise, so the script runs anywhere.

    python examples/quickstart.py
    python examples/quickstart.py --backend b200 --workers 24 --rho 5
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import gadmm_b200.api as G  # noqa: E402


def synthetic(N, d, s, seed=0):
    rng = np.random.default_rng(seed)
    y = rng.standard_normal(s)
    X = []
    for i in range(N):
        q, _ = np.linalg.qr(rng.standard_normal((d, 1)))
        X.append(1.3 ** i * (q @ q.T) + np.eye(d)[:s])
    return np.vstack(X), np.tile(y, N)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="auto", choices=["auto", "oracle", "b200"])
    ap.add_argument("--workers", type=int, default=24)
    ap.add_argument("--dim", type=int, default=50)
    ap.add_argument("--rho", type=float, default=5.0)
    ap.add_argument("--iters", type=int, default=1000)
    ap.add_argument("--acc", type=float, default=1e-4)
    a = ap.parse_args()
    N, d, s = a.workers, a.dim, a.dim
    XX, YY = synthetic(N, d, s)
    obj0 = G.opt_sol_closedForm(XX, YY, backend=a.backend)
    print(f"optimum of the stacked problem: {obj0:.10f}")

    obj, loss, it, t = G.group_ADMM_closedForm(XX, YY, a.rho, N, d, s, a.iters, obj0, a.acc,
                                               backend=a.backend)
    print(f"GADMM (chain 1..N)     : {it:5d} iterations to |obj - obj0| < {a.acc:g}, "
          f"last gap {loss[-1]:.3e}, clock {t[-1]:.4f} s")

    path, cost, *_ = G.findPath2(N)
    obj, loss, it, t, comm = G.dynamic_group_ADMM_closedForm(XX, YY, a.rho, N, d, s, a.iters, obj0,
                                                             a.acc, path, cost, 10,
                                                             backend=a.backend)
    print(f"D-GADMM (tau = 10)     : {it:5d} iterations, last gap {loss[-1]:.3e}, "
          f"link energy {comm[-1]:.3e}")

    obj, loss, it, t = G.standared_ADMM(XX, YY, 1.0, N, d, s, 4 * a.iters, obj0, a.acc,
                                        backend=a.backend)
    print(f"standard ADMM (star)   : {it:5d} iterations, last gap {loss[-1]:.3e}")


if __name__ == "__main__":
    main()
