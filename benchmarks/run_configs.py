#!/usr/bin/env python
"""Run the BASELINE.json configurations (2-5) on the GPUs of one node and print one JSON line per
measurement (rank 0).  Launch with torchrun for N > 1 GPUs:

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        benchmarks/run_configs.py --which all

Sizes default to what fits a short run; every line states the shape it used.  Timing: CUDA events
/ device clocks inside the engines, max over ranks; "time_to_acc_s" is the device time from the end
of iteration 1 to the iteration where |obj - obj0| < 1e-4 first held.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def emit(D, **kw):
    if D.rank == 0:
        print(json.dumps(kw), flush=True)


def reference_shape(D, a):
    """Config 2 at the reference's own shape (N=24, s=d=50): GADMM vs star ADMM, time-to-1e-4."""
    import gadmm_b200 as g
    from gadmm_b200 import oracle as O
    from gadmm_b200.data import linreg_synthetic
    rng = np.random.default_rng(0)
    X, y, Hmax, XX, YY = linreg_synthetic(24, 50, 1.3, rng, y=rng.standard_normal(50))
    obj0 = O.opt_sol_closedForm(XX, YY)
    for rho in (3.0, 5.0, 7.0):
        for backend in ("b200", "nccl"):
            r = g.group_ADMM_closedForm(XX, YY, rho, 24, 50, 50, 4000, obj0, 1e-4, backend=backend)
            emit(D, config="reference shape N=24 s=d=50", algo="GADMM", backend=backend, rho=rho,
                 gpus=D.world, Iter=int(r[2]), comm_rounds=int(r[2]) * 24,
                 time_to_acc_s=float(r[3][-1]), us_per_iter=1e6 * float(r[3][-1]) / max(r[2] - 1, 1))
        if D.rank == 0:
            t0 = time.perf_counter()
            ro = O.group_ADMM_closedForm(XX, YY, rho, 24, 50, 50, 4000, obj0, 1e-4)
            emit(D, config="reference shape N=24 s=d=50", algo="GADMM", backend="oracle (CPU fp64)",
                 rho=rho, Iter=int(ro[2]), wall_s=time.perf_counter() - t0)
        D.barrier()
    for backend in ("b200", "nccl"):
        r = g.standared_ADMM(XX, YY, 1.0, 24, 50, 50, 4000, obj0, 1e-4, backend=backend)
        emit(D, config="reference shape N=24 s=d=50", algo="star ADMM", backend=backend, rho=1.0,
             gpus=D.world, Iter=int(r[2]), time_to_acc_s=float(r[3][-1]))


def dgadmm(D, a):
    """Config 4: D-GADMM with findPath2 re-chaining every tau iterations across the GPUs."""
    import gadmm_b200 as g
    from gadmm_b200 import oracle as O
    from gadmm_b200.data import linreg_synthetic
    N, s = a.dg_workers, a.dg_workers
    rng = np.random.default_rng(1)
    X, y, Hmax, XX, YY = linreg_synthetic(N, s, 1.0 + 3.0 / N, rng, y=rng.standard_normal(s))
    obj0 = O.opt_sol_closedForm(XX, YY)
    path, cost, _ = g.findPath(N, rng=2)
    for tau in (1e9, 100, 10, 1):
        r = g.dynamic_group_ADMM_closedForm(XX, YY, 3.0, N, s, s, a.dg_iters, obj0, 1e-4, path, cost,
                                            tau, backend="b200", rng=5)
        emit(D, config=f"D-GADMM N={N} s=d={s} rho=3", algo="D-GADMM", tau=("inf" if tau > 1e8 else tau),
             gpus=D.world, Iter=int(r[2]), reached=bool(r[1][-1] < 1e-4),
             time_s=float(r[3][-1]), us_per_iter=1e6 * float(r[3][-1]) / max(r[2] - 1, 1),
             final_loss=float(r[1][-1]))


def one_worker_per_gpu(D, a):
    """Config 5: one worker per GPU, s samples x d features each (Gram GEMM + neighbour exchange)."""
    import torch
    from gadmm_b200.engine.chain import ChainEngine
    from gadmm_b200.engine.scaled import build_scaled_linreg
    N, d, s = D.world, a.t3_d, a.t3_s
    t0 = time.time()
    eng = ChainEngine(N, d, 1.0, 0.0, 1e-4, a.t3_iters + 8, dist=D, stop=True)
    obj0 = build_scaled_linreg(eng, s, seed=3, gram="tc", noise=0.1)
    eng.set_path(np.arange(N))
    torch.cuda.synchronize()
    D.barrier()
    setup = time.time() - t0
    it = eng.run(a.t3_iters, window=64)
    obj, loss, it, tv = eng.results(it)
    emit(D, config=f"one worker per GPU: {N} workers x s={s} x d={d}", algo="GADMM", gpus=N,
         setup_s=setup, gram_useful_flop_per_gpu=float(s) * d * d, Iter=int(it),
         reached=bool(loss[-1] < 1e-4), time_to_acc_s=float(tv[-1]),
         us_per_iter=1e6 * float(tv[-1]) / max(it - 1, 1), final_loss=float(loss[-1]), obj0=obj0)


def logistic(D, a):
    """Config 3 (scaled down): logistic GADMM with exact Newton local solves vs the inexact GD one."""
    import gadmm_b200 as g
    from gadmm_b200 import oracle as O
    if D.rank != 0:
        return
    rng = np.random.default_rng(4)
    N, s, d, lam = a.lg_workers, a.lg_s, a.lg_d, 1e-3
    XX = rng.standard_normal((N * s, d)) / np.sqrt(d)
    th = rng.standard_normal(d)
    YY = np.where(XX @ th + 0.3 * rng.standard_normal(N * s) > 0, 1.0, -1.0)
    star = O.opt_sol_logistic(XX, YY, d, lam, N)
    t0 = time.perf_counter()
    r = g.group_ADMM_logistic(XX, YY, 0.5, N, d, s, 400, star, lam, 1e-4, backend="b200")
    emit(D, config=f"logistic N={N} s={s} d={d}", algo="GADMM logistic (Newton)", gpus=1,
         Iter=int(r[2]), wall_s=time.perf_counter() - t0, final_loss=float(r[1][-1]))
    t0 = time.perf_counter()
    r = g.group_ADMM_logistic_GD(XX, YY, 0.5, N, d, s, 3000, star, lam, 1e-4, 0.005, backend="b200")
    emit(D, config=f"logistic N={N} s={s} d={d}", algo="GADMM logistic (inexact GD)", gpus=1,
         Iter=int(r[2]), wall_s=time.perf_counter() - t0, final_loss=float(r[1][-1]))


def logistic_scaled(D, a):
    """Config 3 at full size: logistic GADMM, exact local solves (Newton-CG), chain split over the
    GPUs of the job (--ls-workers x --ls-d x --ls-s, default 512 x 8192 x 1024)."""
    import torch
    from gadmm_b200.engine.logistic_scaled import ScaledLogisticGadmm
    t0 = time.time()
    eng = ScaledLogisticGadmm(a.ls_workers, a.ls_d, a.ls_s, 1e-3, a.ls_rho, dist=D, max_iter=a.ls_iters)
    obj0, _ = eng.solve_optimum()
    torch.cuda.synchronize()
    setup = time.time() - t0
    acc = a.ls_rel_acc * abs(obj0)
    obj, loss, it, tv = eng.run(a.ls_iters, obj0, acc)
    st = eng.state()
    emit(D, config=f"logistic {a.ls_workers} workers x d={a.ls_d} x s={a.ls_s}, rho={a.ls_rho}",
         algo="GADMM logistic (Newton-CG local solves)", gpus=D.world, setup_s=setup, Iter=int(it),
         reached=bool(loss[-1] < acc), rel_gap=float(loss[-1] / abs(obj0)), time_s=float(tv[-1]),
         ms_per_iter=1e3 * float(tv[-1]) / max(len(tv), 1), hessvecs=st.get("hessvecs"),
         evals=st.get("evals"), launches=st["launches"])


def baselines_scaled(D, a):
    """Config 2, the "vs" list at the scaled shape on ONE GPU: iteration rate of every first-order
    baseline (gradient provider = streaming SYMV over all packed Grams + one fused update kernel)."""
    if D.rank != 0:
        return
    import torch
    from gadmm_b200.engine.scaled import build_scaled_first_order
    from gadmm_b200.ops import native as nv
    N, d, s, T = a.bl_workers, a.bl_d, a.bl_s, a.bl_iters
    t0 = time.time()
    fo, obj0 = build_scaled_first_order(N, d, s)
    alpha = 1.0 / fo.lambda_max(tol=1e-9, max_it=400)
    Hmax = fo.hmax_per_worker(30, tol=None)
    setup = time.time() - t0
    thr = 10.0 / (alpha ** 2 * N ** 2) / 10
    runs = (("GD", lambda: fo.run_shared(nv.FO_GD, alpha, 0.0, Hmax, T, obj0, None)),
            ("LAG-PS", lambda: fo.run_shared(nv.FO_LAG_PS, alpha, thr, Hmax, T, obj0, None)),
            ("LAG-WK", lambda: fo.run_shared(nv.FO_LAG_WK, alpha, thr / 10, Hmax, T, obj0, None)),
            ("cyclic IAG", lambda: fo.run_shared(nv.FO_IAG_CYCLIC, alpha / N, 0.0, Hmax, T, obj0, None)),
            ("DGD", lambda: fo.run_dgd(alpha / 100, T, obj0)),
            ("dual averaging", lambda: fo.run_dualavg(alpha, T, obj0, 1e-4)))
    for name, fn in runs:
        r = fn()
        tv = r[3] if name == "dual averaging" else r[2]
        loss = r[1]
        emit(D, config=f"baselines N={N} d={d} s={s} (1 GPU)", algo=name, iters=len(loss),
             ms_per_iter=1e3 * float(tv[-1]) / max(len(tv) - 1, 1), loss_first=float(loss[0]),
             loss_last=float(loss[-1]), setup_s=setup,
             uploads=(float(r[3][-1]) if name.startswith("LAG") else None))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bl-workers", type=int, default=1024)
    ap.add_argument("--bl-d", type=int, default=4096)
    ap.add_argument("--bl-s", type=int, default=1024)
    ap.add_argument("--bl-iters", type=int, default=60)
    ap.add_argument("--which", default="all")
    ap.add_argument("--dg-workers", type=int, default=64)
    ap.add_argument("--dg-iters", type=int, default=6000)
    ap.add_argument("--t3-d", type=int, default=16384)
    ap.add_argument("--t3-s", type=int, default=1048576)
    ap.add_argument("--t3-iters", type=int, default=2000)
    ap.add_argument("--lg-workers", type=int, default=16)
    ap.add_argument("--lg-s", type=int, default=512)
    ap.add_argument("--lg-d", type=int, default=256)
    ap.add_argument("--ls-workers", type=int, default=512)
    ap.add_argument("--ls-d", type=int, default=8192)
    ap.add_argument("--ls-s", type=int, default=1024)
    ap.add_argument("--ls-rho", type=float, default=0.3)
    ap.add_argument("--ls-iters", type=int, default=30)
    ap.add_argument("--ls-rel-acc", type=float, default=1e-4)
    a = ap.parse_args()
    import torch
    from gadmm_b200.parallel.runtime import Dist, init_process_group
    rank, world, local = init_process_group("nccl")
    torch.cuda.set_device(local)
    D = Dist()
    todo = a.which.split(",")
    for name, fn in (("reference", reference_shape), ("dgadmm", dgadmm),
                     ("t3", one_worker_per_gpu), ("logistic", logistic),
                     ("logistic_scaled", logistic_scaled),
                     ("baselines", baselines_scaled)):
        if "all" in todo or name in todo:
            fn(D, a)
            D.barrier()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
