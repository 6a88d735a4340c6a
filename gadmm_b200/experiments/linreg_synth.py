"""``LinearRegression_Synthetic.m``: N=24, 50x50 shards ``1.3^(i-1) q q' + I``, GD/DGD/LAG/IAG
baselines (60 000 it), dual averaging, GADMM with rho in {3,5,7} (cap 1000, acc 1e-4)."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import linreg_synthetic
from ..results import cumulative_com
from .common import add_common_args, apply_overrides, baseline_curves, dataset_or_surrogate, finish, gadmm_comm, optimum_backend


@dataclass
class Config:
    num_workers: int = 24            # LinearRegression_Synthetic.m:17
    num_sample: int = 50             # :21-22 (first 50 rows of the Body-Fat set)
    growth: float = 1.3              # :32
    num_iter_baselines: int = 60000  # :14
    num_iter_gadmm: int = 1000       # :78
    rhos: tuple = (3.0, 5.0, 7.0)    # :79,88,93
    acc: float = 1e-4                # :73,80
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
    curves = {"obj0": np.array([obj0]), "opt_obj": obj0 * np.ones(cfg.num_iter_baselines)}
    out = api.GD_DGD_LAG(X, y, N, d, Hmax, cfg.num_iter_baselines, obj0, XX, YY, s,
                         backend=cfg.backend, rng=rng)
    baseline_curves(curves, out, N, obj0)
    da = api.dual_averaging(XX, YY, N, d, s, cfg.num_iter_baselines, obj0, cfg.acc, out[18],
                            backend=cfg.backend)
    curves.update(obj_dualAvg=da[0], loss_dualAvg=da[1], Iter_dualAvg=np.array([float(da[2])]),
                  dualAvg_time=da[3], cumulative_com_dualAvg=cumulative_com("dualAvg", da[2], N))
    summary = {"driver": "linreg_synth", "N": N, "d": d, "obj0": obj0, "surrogate_targets": surrogate,
               "Iter_dualAvg": int(da[2])}
    for rho in cfg.rhos:
        r = api.group_ADMM_closedForm(XX, YY, rho, N, d, s, cfg.num_iter_gadmm, obj0, cfg.acc,
                                      backend=cfg.backend)
        tag = f"GADMM_rho{rho:g}"
        gadmm_comm(tag, curves, r[0], r[1], r[2], N, obj0)
        curves[f"gadmm_time_rho{rho:g}"] = r[3]
        summary[f"Iter_rho{rho:g}"] = int(r[2])
    return curves, summary


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_iter_baselines = 3000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
