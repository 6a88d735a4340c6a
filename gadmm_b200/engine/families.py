"""Reference-signature GPU entry points for the baseline families and star ADMM."""
from __future__ import annotations

import numpy as np

from ..ops import native as nv
from .first_order import FirstOrder

TRIGGERSLOT = 10


def _as_rng(rng):
    return rng if isinstance(rng, np.random.Generator) else np.random.default_rng(rng)


def _six(fo: FirstOrder, alpha, Hmax, N, T, obj0, accuracy, rng, fo_own=None):
    thrd = 10.0 / (alpha ** 2 * N ** 2) / TRIGGERSLOT          # GD_DGD_LAG.m:47
    thrd5 = 1.0 / (alpha ** 2 * N ** 2) / TRIGGERSLOT          # :65
    nonprob = Hmax / Hmax.sum()
    uni = np.concatenate([[0.0], _as_rng(rng).random(T), [0.0]])
    if fo.dist is not None:
        uni = fo.dist.broadcast_object(uni)      # every GPU must sample the same worker
    o_dgd = (fo_own or fo).run_dgd(alpha / 100.0, T, obj0)     # :127 (chain stencil: one GPU)
    ps = fo.run_shared(nv.FO_LAG_PS, alpha, thrd, Hmax, T, obj0, accuracy)
    wk = fo.run_shared(nv.FO_LAG_WK, alpha, thrd5, Hmax, T, obj0, accuracy)
    cy = fo.run_shared(nv.FO_IAG_CYCLIC, alpha / N, 0.0, Hmax, T, obj0, accuracy)
    ra = fo.run_shared(nv.FO_IAG_RANDOM, alpha / N, 0.0, Hmax, T, obj0, accuracy, uni=uni,
                       cumprob=np.cumsum(nonprob))
    return o_dgd, ps, wk, cy, ra


def _dist_or_none(dist):
    if dist is not None:
        return dist if dist.world > 1 else None
    import os
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        from ..parallel.runtime import Dist
        return Dist()
    return None


def gd_dgd_lag(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede, y_fede, num_sample,
               rng=None, store="f64", dist=None):
    """GPU ``GD_DGD_LAG`` (19 outputs, GD_DGD_LAG.m:1-4).  Under torchrun the parameter-server
    methods (GD, LAG-PS, LAG-WK, IAG) split the workers over the GPUs and reduce inside the server
    kernel through NVLink peer stores; DGD exchanges the boundary gradient rows of its chain
    segments the same way."""
    N, d, T = int(num_workers), int(num_feature), int(num_iter)
    dist = _dist_or_none(dist)
    fo = FirstOrder(X_fede, y_fede, N, d, num_sample, logistic=False, store=store, dist=dist)
    fo_own = None       # DGD shares the distributed engine (chain halos over NVLink)
    Hmax = np.asarray(Hmax, dtype=np.float64).reshape(-1)[:N]
    alpha = 1.0 / fo.lambda_max()                              # :19,46
    gd = fo.run_shared(nv.FO_GD, alpha, 0.0, Hmax, T, obj0, None)
    dgd, ps, wk, cy, ra = _six(fo, alpha, Hmax, N, T, obj0, None, rng, fo_own)
    return (gd[0], gd[1], gd[2], dgd[0], dgd[1], dgd[2], ps[0], ps[1], ps[2], ps[3],
            wk[0], wk[1], wk[2], wk[3], cy[0], cy[1], ra[0], ra[1], alpha)


def gd_dgd_lag_logistic(X, y, num_workers, num_feature, Hmax, num_iter, X_fede, y_fede, num_sample,
                        lam, accuracy, rng=None, naive=True, dist=None):
    """GPU ``GD_DGD_LAG_logistic`` (20 outputs; ``obj1`` = GD's last objective, :131-133)."""
    N, d, T = int(num_workers), int(num_feature), int(num_iter)
    dist = _dist_or_none(dist)
    fo = FirstOrder(X_fede, y_fede, N, d, num_sample, logistic=True, lam=lam, dist=dist)
    fo_own = None
    Hmax = np.asarray(Hmax, dtype=np.float64).reshape(-1)[:N]
    alpha = 1.0 / (0.25 * fo.lambda_max() + lam)               # :24
    gd = fo.run_shared(nv.FO_GD, alpha, 0.0, Hmax, T, 0.0, None)
    obj1 = float(gd[0][-1])
    loss_gd = np.abs(gd[0] - obj1)
    dgd, ps, wk, cy, ra = _six(fo, alpha, Hmax, N, T, obj1, accuracy, rng, fo_own)
    return (gd[0], loss_gd, gd[2], dgd[0], dgd[1], dgd[2], ps[0], ps[1], ps[2], ps[3],
            wk[0], wk[1], wk[2], wk[3], cy[0], cy[1], ra[0], ra[1], alpha, obj1)


def dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc, alpha,
                   mode="faithful", return_state=False, dist=None):
    fo = FirstOrder(XX, YY, no_workers, num_feature, noSamples, logistic=False,
                    dist=_dist_or_none(dist))
    obj, loss, it, tv, Theta = fo.run_dualavg(alpha, int(num_iter), obj0, acc, jacobi=(mode == "jacobi"))
    res = (obj, loss, it, tv)
    return res + ({"theta": Theta, "iteration": it},) if return_state else res


def dual_averaging_logistic(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc, alpha,
                            eta, mode="faithful", naive=True, return_state=False, dist=None):
    fo = FirstOrder(XX, YY, no_workers, num_feature, noSamples, logistic=True, lam=eta,
                    dist=_dist_or_none(dist))
    obj, loss, it, tv, Theta = fo.run_dualavg(alpha, int(num_iter), obj0, acc, jacobi=(mode == "jacobi"))
    res = (obj, loss, it, tv)
    return res + ({"theta": Theta, "iteration": it},) if return_state else res
