"""``Dynamic_LinearRegression_Synthetic.m``: N=50, rho=3, cap 5000.  Part A: static GADMM vs
D-GADMM on 1+1000 pre-generated ``findPath`` chains at tau=10, with the 5/15 re-chaining overhead
on D-GADMM's cost and time (:123,125).  Part B: live D-GADMM for tau in {1e9, 1, 10, 50, 100}.
``faithful=True`` keeps quirks 4/6/7 of SURVEY.md section 2.2 (matrix-column first cost, findPath
seed chain with findPath2 refreshes, static cost measured along path0)."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import linreg_synthetic
from ..results import with_rechain_overhead
from ..topology import calc_cost, findPath
from .common import add_common_args, apply_overrides, dataset_or_surrogate, finish, optimum_backend


@dataclass
class Config:
    num_workers: int = 50                 # Dynamic_LinearRegression_Synthetic.m:17
    num_sample: int = 50
    growth: float = 1.3
    rho: float = 3.0                      # :79
    num_iter: int = 5000                  # :78
    acc: float = 1e-4
    n_topologies: int = 1000              # :95
    tau_a: int = 10                       # :83
    taus_b: tuple = (1e9, 1, 10, 50, 100)  # :142-173
    faithful: bool = True
    backend: str = "auto"
    seed: int = 0
    data_dir: str = None


def run(cfg: Config):
    rng = np.random.default_rng(cfg.seed)
    _, ydata, surrogate = dataset_or_surrogate(cfg.data_dir, "data29", cfg.num_sample, 1, rng)
    X, y, Hmax, XX, YY = linreg_synthetic(cfg.num_workers, cfg.num_sample, cfg.growth, rng,
                                          y=ydata[:cfg.num_sample])
    N, s, d = cfg.num_workers, cfg.num_sample, XX.shape[1]
    obj0 = api.opt_sol_closedForm(XX, YY, backend=optimum_backend(cfg.backend))
    base = (XX, YY, cfg.rho, N, d, s, cfg.num_iter, obj0, cfg.acc)
    # ---- part A
    path0, cost0, grid = findPath(N, rng)
    paths, costs, static = [path0], [cost0], [calc_cost(grid, path0)]
    for _ in range(cfg.n_topologies):
        p, c, grid = findPath(N, rng)
        paths.append(p)
        costs.append(c)
        static.append(calc_cost(grid, path0))      # quirk 7: static cost walks path0 on the new grid
    paths, costs, static = np.array(paths), np.array(costs), np.array(static)
    rs = api.static_group_ADMM_closedForm(*base, cfg.tau_a, static, backend=cfg.backend)
    rd = api.dynamic_group_ADMM_closedForm_v0(*base, paths, costs, cfg.tau_a, backend=cfg.backend,
                                              faithful=cfg.faithful)
    curves = {"obj0": np.array([obj0]),
              "loss_GADMM_rho3s": rs[1], "Iter_3s": np.array([float(rs[2])]), "gadmm_time3s": rs[3],
              "cumulative_com_GADMM_rho3s": rs[4],
              "loss_GADMM_rho3_d0": rd[1], "Iter_3_d0": np.array([float(rd[2])]),
              "gadmm_time3_d0": with_rechain_overhead(rd[3]),
              "cumulative_com_GADMM_rho3_d0": with_rechain_overhead(rd[4])}
    summary = {"driver": "dgadmm_synth", "N": N, "obj0": obj0, "Iter_static": int(rs[2]),
               "Iter_dgadmm_v0": int(rd[2]), "surrogate_targets": surrogate}
    # ---- part B
    path, cost, _ = findPath(N, rng)               # quirk 6: findPath seed, findPath2 refreshes
    for tau in cfg.taus_b:
        r = api.dynamic_group_ADMM_closedForm(*base, path, cost, tau, backend=cfg.backend,
                                              rng=np.random.default_rng(cfg.seed + 17))
        tag = "inf" if tau >= 1e8 else f"{int(tau)}"
        curves[f"loss_GADMM_d_tau{tag}"] = r[1]
        curves[f"com_cost_d_tau{tag}"] = r[4]
        curves[f"gadmm_time_d_tau{tag}"] = r[3]
        summary[f"Iter_tau{tag}"] = int(r[2])
    return curves, summary


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_workers, cfg.n_topologies, cfg.num_iter = 24, 600, 3000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
