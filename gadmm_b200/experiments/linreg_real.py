"""``LinearRegression_Real.m``: Body-Fat data split 25 samples per worker, baselines 40 000 it,
GADMM rho in {3,5,7} with cap 800."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import hmax_linear, split_dataset
from ..results import cumulative_com
from .common import add_common_args, apply_overrides, baseline_curves, dataset_or_surrogate, finish, gadmm_comm, optimum_backend


@dataclass
class Config:
    per_split: int = 25              # LinearRegression_Real.m:15
    num_iter_baselines: int = 40000  # :20
    num_iter_gadmm: int = 800        # :84
    rhos: tuple = (3.0, 5.0, 7.0)
    acc: float = 1e-4
    backend: str = "auto"
    seed: int = 0
    data_dir: str = None


def run(cfg: Config):
    rng = np.random.default_rng(cfg.seed)
    Xd, yd, surrogate = dataset_or_surrogate(cfg.data_dir, "data29", 252, 14, rng)
    X, y, XX, YY, s = split_dataset(Xd, yd, per_split=cfg.per_split)
    N, d = len(X), XX.shape[1]
    Hmax = hmax_linear(X)
    obj0 = api.opt_sol_closedForm(XX, YY, backend=optimum_backend(cfg.backend))
    curves = {"obj0": np.array([obj0])}
    out = api.GD_DGD_LAG(X, y, N, d, Hmax, cfg.num_iter_baselines, obj0, XX, YY, s,
                         backend=cfg.backend, rng=rng)
    baseline_curves(curves, out, N, obj0)
    da = api.dual_averaging(XX, YY, N, d, s, cfg.num_iter_baselines, obj0, cfg.acc, out[18],
                            backend=cfg.backend)
    curves.update(obj_dualAvg=da[0], loss_dualAvg=da[1], dualAvg_time=da[3],
                  cumulative_com_dualAvg=cumulative_com("dualAvg", da[2], N))
    summary = {"driver": "linreg_real", "N": N, "d": d, "obj0": obj0, "surrogate_data": surrogate}
    for rho in cfg.rhos:
        r = api.group_ADMM_closedForm(XX, YY, rho, N, d, s, cfg.num_iter_gadmm, obj0, cfg.acc,
                                      backend=cfg.backend)
        gadmm_comm(f"GADMM_rho{rho:g}", curves, r[0], r[1], r[2], N, obj0)
        curves[f"gadmm_time_rho{rho:g}"] = r[3]
        summary[f"Iter_rho{rho:g}"] = int(r[2])
    return curves, summary


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_iter_baselines = 2000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
