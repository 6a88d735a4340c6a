"""Exercise every kernel family once at a moderate, memory-resident size -- the workload for one
`ncu --set full` tour (profiles/kernel_tour.md).  Also runs un-profiled as a smoke check."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import gadmm_b200 as g
from gadmm_b200 import oracle as O
from gadmm_b200.engine.logistic_scaled import ScaledLogisticGadmm
from gadmm_b200.engine.scaled import build_scaled_first_order
from gadmm_b200.ops import native as nv


def main():
    torch.cuda.set_device(0)
    # logistic Newton-CG building block: 148 shards x 512 x 8192 fp32 (2.5 GB), one GADMM iteration
    eng = ScaledLogisticGadmm(296, 8192, 512, 1e-3, 0.3)
    x = torch.randn((148, 8192), dtype=torch.float64, device=eng.dev)
    for mode in (0, 1, 2):                       # gradient / Hessian product fp64 / fp32 multiplies
        eng._xtwx(x, 0, 148, 2, mode)
    eng.run(1, 0.0, 0.0)                         # one device-resident GADMM iteration (logi_* kernels)
    del eng, x
    torch.cuda.empty_cache()
    # first-order baselines at 128 workers x d = 2048 (packed Grams 2.1 GB)
    fo, obj0 = build_scaled_first_order(128, 2048, 512)
    alpha = 1.0 / fo.lambda_max(tol=1e-6, max_it=200)
    Hmax = fo.hmax_per_worker(10, tol=None)
    thr = 10.0 / (alpha ** 2 * 128 ** 2) / 10
    fo.run_shared(nv.FO_GD, alpha, 0.0, Hmax, 3, obj0, None)
    fo.run_shared(nv.FO_LAG_WK, alpha, thr / 10, Hmax, 3, obj0, None)
    fo.run_dgd(alpha / 100, 3, obj0)
    fo.run_dualavg(alpha, 3, obj0, 1e-4)
    del fo
    torch.cuda.empty_cache()
    # star ADMM, D-GADMM with device-side re-chaining, persistent kernel: host data, small shapes
    rng = np.random.default_rng(0)
    N, s, d = 64, 96, 512
    XX = rng.standard_normal((N * s, d)) / np.sqrt(s)
    YY = XX @ rng.standard_normal(d) + 0.1 * rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    g.standared_ADMM(XX, YY, 1.0, N, d, s, 4, obj0, 0.0, backend="b200")
    g.dynamic_group_ADMM_closedForm(XX, YY, 2.0, N, d, s, 6, obj0, 0.0, np.arange(1, N + 1),
                                    np.ones(N - 1), 2, backend="b200", chain_gen="device", rng=3)
    N2, s2, d2 = 24, 50, 50
    X2 = rng.standard_normal((N2 * s2, d2))
    Y2 = rng.standard_normal(N2 * s2)
    o2 = O.opt_sol_closedForm(X2, Y2)
    g.group_ADMM_closedForm(X2, Y2, 3.0, N2, d2, s2, 200, o2, 0.0, backend="b200")
    lam = 1e-3
    Yl = np.where(Y2 > 0, 1.0, -1.0)
    ol = O.opt_sol_logistic(X2, Yl, d2, lam, N2)
    g.group_ADMM_logistic_GD(X2, Yl, 0.5, N2, d2, s2, 2, ol, lam, 0.0, 0.005, backend="b200")
    torch.cuda.synchronize()
    print("TOUR_OK")


if __name__ == "__main__":
    main()
