"""``LinearRegression_gadmm_vs_admm.m``: N=24, rho=1 (the script's variables say "rho5"), cap
20 000.  One ``findPath2`` topology gives the star cost ``sum(P_central)+max(P_central)`` and the
chain cost ``sum(pathCost)`` per iteration; runs star ADMM, GADMM and D-GADMM tau in {1,10,50}
and emits the energy-weighted cumulative communication curves (:128-179)."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import linreg_synthetic
from ..results import energy_cumulative, star_cost
from ..topology import findPath2
from .common import add_common_args, apply_overrides, dataset_or_surrogate, finish, optimum_backend


@dataclass
class Config:
    num_workers: int = 24             # LinearRegression_gadmm_vs_admm.m:17
    num_sample: int = 50
    growth: float = 1.3
    rho: float = 1.0                  # :91
    num_iter: int = 20000             # :78
    taus: tuple = (1, 10, 50)         # :101-119
    acc: float = 1e-4
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
    path, pathCost, grid, P_central, center = findPath2(N, rng)
    cost_star, cost_chain = star_cost(P_central), float(np.sum(pathCost))      # :84-87
    base = (XX, YY, cfg.rho, N, d, s, cfg.num_iter, obj0, cfg.acc)
    a = api.standared_ADMM(*base, backend=cfg.backend)
    g = api.group_ADMM_closedForm(*base, backend=cfg.backend)
    curves = {"obj0": np.array([obj0]), "cost_star_topology": np.array([cost_star]),
              "cost_decentralized": np.array([cost_chain]), "center": np.array([float(center)]),
              "loss_ADMM": a[1], "admm_time": a[3],
              "cumulative_com_ADMM": energy_cumulative(cost_star, a[2]),
              "loss_GADMM": g[1], "gadmm_time": g[3],
              "cumulative_com_GADMM": energy_cumulative(cost_chain, g[2])}
    summary = {"driver": "gadmm_vs_admm", "N": N, "obj0": obj0, "Iter_admm": int(a[2]),
               "Iter_gadmm": int(g[2]), "cost_star": cost_star, "cost_chain": cost_chain,
               "surrogate_targets": surrogate}
    for tau in cfg.taus:
        r = api.dynamic_group_ADMM_closedForm(*base, path, pathCost, tau, backend=cfg.backend,
                                              rng=np.random.default_rng(cfg.seed + 3))
        curves[f"loss_GADMM_d{tau}"] = r[1]
        # the script charges sum(pathCost) of the *initial* topology per iteration (:158-179)
        curves[f"cumulative_com_GADMM_d{tau}"] = energy_cumulative(cost_chain, r[2])
        curves[f"com_cost_d{tau}"] = r[4]
        summary[f"Iter_d{tau}"] = int(r[2])
    return curves, summary


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_iter = 6000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
