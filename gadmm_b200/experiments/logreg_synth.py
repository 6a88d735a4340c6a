"""``LogisticRegression_Synthetic.m``: N=24, shards ``q q' + I``, lambda=1e-5, baselines 100 000 it
(GD's last objective defines obj0), GADMM-GD rho in {3e-4, 2e-4}, gdStep 2.2, cap 400."""
from __future__ import annotations

import argparse
from dataclasses import dataclass

import numpy as np

from .. import api
from ..data import logreg_synthetic
from ..io import load_input_data
from ..results import cumulative_com
from .common import add_common_args, apply_overrides, baseline_curves, finish, gadmm_comm


@dataclass
class Config:
    num_workers: int = 24               # LogisticRegression_Synthetic.m:17
    num_sample: int = 50
    lam: float = 1e-5                   # :40-41
    num_iter_baselines: int = 100000    # :14
    num_iter_dualavg: int = 100000
    num_iter_gadmm: int = 400           # :88
    rhos: tuple = (3e-4, 2e-4)          # :91,98
    gd_step: float = 2.2                # :89
    acc: float = 1e-4
    backend: str = "auto"
    seed: int = 0
    input_mat: str = None               # e.g. /root/reference/inputData.mat


def run(cfg: Config):
    rng = np.random.default_rng(cfg.seed)
    N, s = cfg.num_workers, cfg.num_sample
    if cfg.input_mat:
        XX, YY = load_input_data(cfg.input_mat)
        d = XX.shape[1]
        X = [XX[i * s:(i + 1) * s] for i in range(N)]
        y = [YY[i * s:(i + 1) * s] for i in range(N)]
        from ..data import hmax_logistic
        Hmax = hmax_logistic(X, cfg.lam)
    else:
        X, y, Hmax, XX, YY = logreg_synthetic(N, s, cfg.lam, rng)
        d = XX.shape[1]
    out = api.GD_DGD_LAG_logistic(X, y, N, d, Hmax, cfg.num_iter_baselines, XX, YY, s, cfg.lam,
                                  cfg.acc, backend=cfg.backend, rng=rng)
    obj0 = out[19]
    curves = {"obj0": np.array([obj0])}
    baseline_curves(curves, out, N, obj0)
    da = api.dual_averaging_logisticReg(XX, YY, N, d, s, cfg.num_iter_dualavg, obj0, cfg.acc, out[18],
                                        cfg.lam, backend=cfg.backend)
    curves.update(obj_dualAvg=da[0], loss_dualAvg=da[1], dualAvg_time=da[3],
                  cumulative_com_dualAvg=cumulative_com("dualAvg", da[2], N))
    summary = {"driver": "logreg_synth", "N": N, "d": d, "obj0": obj0}
    for rho in cfg.rhos:
        r = api.group_ADMM_logistic_GD(XX, YY, rho, N, d, s, cfg.num_iter_gadmm, obj0, cfg.lam,
                                       cfg.acc, cfg.gd_step, backend=cfg.backend)
        gadmm_comm(f"GADMM_rho{rho:g}", curves, r[0], r[1], r[2], N)
        curves[f"gadmm_time_rho{rho:g}"] = r[3]
        summary[f"Iter_rho{rho:g}"] = int(r[2])
    return curves, summary


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    add_common_args(ap)
    ap.add_argument("--input-mat", dest="input_mat", default=None)
    a = ap.parse_args(argv)
    cfg = apply_overrides(Config(), a)
    if a.quick:
        cfg.num_iter_baselines, cfg.num_iter_dualavg = 5000, 2000
    curves, summary = run(cfg)
    return finish(curves, a.out, summary, a.fig)


if __name__ == "__main__":
    main()
