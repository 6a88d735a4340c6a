"""``dynamic_LinearRegression_Real.m``: Body-Fat data, rho=0.1, cap 20 000, tau=50.  The shipped
script calls a function that does not exist (``dynamic_group_ADMM_closedForm_v2``, :109); this
driver runs the live D-GADMM (``dynamic_group_ADMM_closedForm``) in its place and GADMM on the
natural chain for comparison."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import split_dataset
from ..topology import findPath
from .common import add_common_args, apply_overrides, dataset_or_surrogate, finish, optimum_backend


@dataclass
class Config:
    per_split: int = 25
    rho: float = 0.1                  # dynamic_LinearRegression_Real.m:85
    num_iter: int = 20000             # :84
    tau: int = 50                     # :102
    acc: float = 1e-4
    backend: str = "auto"
    seed: int = 0
    data_dir: str = None


def run(cfg: Config):
    rng = np.random.default_rng(cfg.seed)
    Xd, yd, surrogate = dataset_or_surrogate(cfg.data_dir, "data29", 252, 14, rng)
    X, y, XX, YY, s = split_dataset(Xd, yd, per_split=cfg.per_split)
    N, d = len(X), XX.shape[1]
    obj0 = api.opt_sol_closedForm(XX, YY, backend=optimum_backend(cfg.backend))
    base = (XX, YY, cfg.rho, N, d, s, cfg.num_iter, obj0, cfg.acc)
    g = api.group_ADMM_closedForm(*base, backend=cfg.backend)
    path, cost, _ = findPath(N, rng)
    r = api.dynamic_group_ADMM_closedForm(*base, path, cost, cfg.tau, backend=cfg.backend,
                                          rng=np.random.default_rng(cfg.seed + 1))
    curves = {"obj0": np.array([obj0]), "loss_GADMM": g[1], "gadmm_time": g[3],
              "loss_GADMM_d": r[1], "gadmm_time_d": r[3], "com_cost_d": r[4]}
    return curves, {"driver": "dgadmm_real", "N": N, "d": d, "obj0": obj0, "Iter_gadmm": int(g[2]),
                    "Iter_dgadmm": int(r[2]), "surrogate_data": surrogate}


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_iter = 3000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
