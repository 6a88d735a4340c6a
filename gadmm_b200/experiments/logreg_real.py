"""``LogisticRegression_real.m``: derm data, N=10, lambda=1e-5, dual averaging cap 500 000,
GADMM-GD rho in {0.03, 0.02}, gdStep 0.08, cap 1000."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import hmax_logistic, split_dataset
from ..results import cumulative_com
from .common import add_common_args, apply_overrides, baseline_curves, dataset_or_surrogate, finish, gadmm_comm


@dataclass
class Config:
    num_workers: int = 10               # LogisticRegression_real.m:17
    lam: float = 1e-5                   # :38-39
    num_iter_baselines: int = 100000    # :13
    num_iter_dualavg: int = 500000      # :65
    num_iter_gadmm: int = 1000          # :73
    rhos: tuple = (0.03, 0.02)          # :77,82
    gd_step: float = 0.08               # :74
    acc: float = 1e-4
    backend: str = "auto"
    seed: int = 0
    data_dir: str = None


def run(cfg: Config):
    rng = np.random.default_rng(cfg.seed)
    Xd, yd, surrogate = dataset_or_surrogate(cfg.data_dir, "data11", 358, 34, rng, labels=True)
    if surrogate:
        Xd = Xd * 0.3      # keep the surrogate well inside the naive exp() range
    X, y, XX, YY, s = split_dataset(Xd, yd, num_workers=cfg.num_workers)
    N, d = cfg.num_workers, XX.shape[1]
    Hmax = hmax_logistic(X, cfg.lam)
    out = api.GD_DGD_LAG_logistic(X, y, N, d, Hmax, cfg.num_iter_baselines, XX, YY, s, cfg.lam,
                                  cfg.acc, backend=cfg.backend, rng=rng)
    obj0 = out[19]
    curves = {"obj0": np.array([obj0])}
    baseline_curves(curves, out, N, obj0)
    da = api.dual_averaging_logisticReg(XX, YY, N, d, s, cfg.num_iter_dualavg, obj0, cfg.acc, out[18],
                                        cfg.lam, backend=cfg.backend)
    curves.update(obj_dualAvg=da[0], loss_dualAvg=da[1], dualAvg_time=da[3],
                  cumulative_com_dualAvg=cumulative_com("dualAvg", da[2], N))
    summary = {"driver": "logreg_real", "N": N, "d": d, "obj0": obj0, "surrogate_data": surrogate}
    for rho in cfg.rhos:
        r = api.group_ADMM_logistic_GD(XX, YY, rho, N, d, s, cfg.num_iter_gadmm, obj0, cfg.lam,
                                       cfg.acc, cfg.gd_step, backend=cfg.backend)
        gadmm_comm(f"GADMM_rho{rho:g}", curves, r[0], r[1], r[2], N)
        summary[f"Iter_rho{rho:g}"] = int(r[2])
    return curves, summary


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_iter_baselines, cfg.num_iter_dualavg = 3000, 3000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
