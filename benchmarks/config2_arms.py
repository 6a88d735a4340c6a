#!/usr/bin/env python
"""BASELINE.json config 2: GADMM vs D-GADMM vs standard (star) ADMM vs GD / DGD / LAG-PS / LAG-WK /
IAG / dual averaging on ONE synthetic least-squares problem (default 1024 workers x d = 4096,
s = 1024 samples per worker), every arm on the same data (worker g's shard = Philox stream g, so
chain segments, star ranges and parameter-server ranges over any number of GPUs see the same
problem), `--impl b200` (hand-written kernels, NVLink flags) or `--impl nccl` (torch ops + NCCL).

Per arm: device-timed ms per iteration (max over ranks), iterations / time to the 1e-4 objective
gap where the gap stays below 1e-4 to the end of the run ("reached"), otherwise the gap after the
iteration budget, HBM bytes per iteration per GPU and the achieved fraction of the measured copy
bandwidth, uploads for the LAG variants.  Reference drivers: LinearRegression_Synthetic.m:63-94,
LinearRegression_gadmm_vs_admm.m:94-119, standared_ADMM.m:57-88.

    python -m torch.distributed.run --nproc-per-node 8 ... benchmarks/config2_arms.py --impl b200
"""
import argparse
import gc
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def first_stay_below(loss, acc):
    bad = np.nonzero(np.asarray(loss) >= acc)[0]
    first = (int(bad[-1]) + 1 if bad.size else 0) + 1
    return first if first <= len(loss) else 0


def record(D, out, algo, impl, loss, tv, bytes_per_iter_gpu, hbm, extra=None):
    n = len(loss)
    hit = first_stay_below(loss, 1e-4)
    ms = 1e3 * float(tv[-1] - tv[0]) / max(n - 1, 1)
    rec = {"algo": algo, "impl": impl, "gpus": D.world, "iterations_run": n,
           "ms_per_iter": ms, "reached": bool(hit), "iters_to_1e-4": hit or None,
           "time_to_1e-4_s": float(tv[hit - 1]) if hit else None, "gap_first": float(loss[0]),
           "gap_at_end": float(loss[-1]),
           "GB_per_iter_per_gpu": bytes_per_iter_gpu / 1e9,
           "frac_of_hbm_copy": bytes_per_iter_gpu / (ms * 1e-3) / 1e9 / hbm if ms > 0 else None}
    rec.update(extra or {})
    if D.rank == 0:
        print(json.dumps(rec), flush=True)
        if out:
            with open(out, "a") as f:
                f.write(json.dumps(rec) + "\n")


def free():
    import torch
    gc.collect()
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=1024)
    ap.add_argument("--dim", dest="d", type=int, default=4096)      # (torchrun's parser trips over --d)
    ap.add_argument("--samples", dest="s", type=int, default=1024)
    ap.add_argument("--rho", type=float, default=10.0)
    ap.add_argument("--star-rho", type=float, default=1.0)
    ap.add_argument("--tau", type=int, default=10)
    ap.add_argument("--cap", type=int, default=320, help="iteration budget of the ADMM family")
    ap.add_argument("--fo-iters", type=int, default=200, help="iteration budget of the first-order family")
    ap.add_argument("--impl", default="b200", choices=["b200", "nccl"])
    ap.add_argument("--algos", default="gadmm,dgadmm,star,gd,lagps,lagwk,iag,dgd,dualavg,dualavg_jacobi")
    ap.add_argument("--store", default="f64", choices=["f64", "mixed"],
                    help="chain arms: fp64 inverses, or fp32 inverses until the gap is small, then fp64")
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    import torch
    from gadmm_b200.parallel.runtime import Dist, init_process_group, partition
    from gadmm_b200.engine import scaled as SC
    from gadmm_b200.ops import native as nv
    from gadmm_b200.ops import linalg as LA
    rank, world, local = init_process_group("nccl")
    torch.cuda.set_device(local)
    D = Dist()
    dev = D.device
    N, d, s = a.workers, a.d, a.s
    dp = nv.padded(d)
    algos = a.algos.split(",")
    try:
        hbm = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                          "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        hbm = 6650.0
    t0 = time.time()
    obj0, Gsum, bsum, hsum = SC.scaled_optimum(N, d, s, dp, D)
    if rank == 0:
        print(json.dumps({"setup": "optimum", "obj0": obj0, "s": round(time.time() - t0, 2)}), flush=True)
    packed_bytes = nv.packed_elems(dp) * 8
    own = np.nonzero(partition(N, world) == rank)[0]

    if a.impl == "nccl":
        from gadmm_b200.parallel import nccl_scaled as NS
        NS.run_arms(a, D, obj0, Gsum, algos, record, hbm)
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    # ------------------------------------------------------------------ chain GADMM / D-GADMM
    if "gadmm" in algos or "dgadmm" in algos:
        from gadmm_b200.engine.chain import ChainEngine
        from gadmm_b200.ops.topology import find_path_batch_device
        eng = ChainEngine(N, d, a.rho, obj0, 1e-4, a.cap + 8, dist=D, dynamic="dgadmm" in algos, stop=False,
                          store=a.store)
        for k0, G, b, hyy in SC.worker_grams(eng.gids, d, s, dp, dev=dev):
            eng.ingest_gram(G, b, hyy, k0)
        eng._desc = None
        for name in ("gadmm", "dgadmm"):
            if name not in algos:
                continue
            eng.reset(stop=False, acc=1e-4, obj0=obj0)
            eng.set_path(np.arange(N))
            if name == "dgadmm":
                def chunk_src(base, K):
                    return find_path_batch_device(K, N, seed=11, stream_id=base, device=dev)[0]
                it = eng.run_device_rechain(a.cap, None, a.tau, window=32, chunk=64, chunk_source=chunk_src)
            else:
                it = eng.run(a.cap, window=32)
            obj, loss, it, tv = eng.results(it)
            record(D, a.out, "GADMM (static chain)" if name == "gadmm" else f"D-GADMM tau={a.tau}",
                   "b200", loss, tv, eng.Nloc * packed_bytes, hbm,
                   {"rho": a.rho, "store": a.store, "switched_to_fp64_at": getattr(eng, "switched_at", None)})
        del eng
        free()

    # ------------------------------------------------------------------ star ADMM
    if "star" in algos:
        from gadmm_b200.engine.star import StarEngine
        eng = StarEngine(None, None, a.star_rho, N, d, s, a.cap, obj0, 1e-4, dist=D, stop=False)
        for k0, G, b, hyy in SC.worker_grams(eng.gids, d, s, dp, dev=dev):
            eng.ingest_gram(G, b, hyy, k0)
        for _, G, b, hyy in SC.worker_grams([N - 1], d, s, dp, dev=dev):
            eng.ingest_server(G, b[0], hyy)
        D.barrier()
        it = eng.run(a.cap, window=32)
        obj, loss, it, tv = eng.results(it)
        record(D, a.out, "standard (star) ADMM", "b200", loss, tv, (eng.Nw + 1) * packed_bytes, hbm,
               {"rho": a.star_rho, "nvls": bool(eng.nvls)})
        del eng
        free()

    # ------------------------------------------------------------------ first-order family
    fo_list = [x for x in ("gd", "lagps", "lagwk", "iag", "dgd", "dualavg", "dualavg_jacobi") if x in algos]
    if fo_list:
        from gadmm_b200.engine.first_order import FirstOrder
        Nl = len(own)
        P = torch.empty((Nl, nv.packed_elems(dp)), dtype=torch.float64, device=dev)
        b = torch.zeros((Nl, dp), dtype=torch.float64, device=dev)
        hyy = torch.zeros((Nl,), dtype=torch.float64, device=dev)
        for k0, G, bb, hh in SC.worker_grams(own, d, s, dp, dev=dev):
            nb = G.shape[0]
            LA.pack_sym(G, d, nv.STORE_F64, out=P[k0:k0 + nb])
            b[k0:k0 + nb], hyy[k0:k0 + nb] = bb, hh
        fo = FirstOrder.from_parts(P, b, hyy, Gsum, Nl, d, dev, dist=D, gid0=int(own[0]), Nglob=N)
        alpha = 1.0 / fo.lambda_max(tol=1e-9, max_it=2000)
        hm_loc = torch.as_tensor(fo.hmax_per_worker(30, tol=1e-6, max_it=400), device=dev)
        Hmax = torch.zeros((N,), dtype=torch.float64, device=dev)
        Hmax[int(own[0]):int(own[0]) + Nl] = hm_loc
        D.all_reduce_sum_(Hmax)
        Hmax = Hmax.cpu().numpy()
        thr = 10.0 / (alpha ** 2 * N ** 2) / 10
        T = a.fo_iters
        runs = {"gd": ("GD", lambda: fo.run_shared(nv.FO_GD, alpha, 0.0, Hmax, T, obj0, None)),
                "lagps": ("LAG-PS", lambda: fo.run_shared(nv.FO_LAG_PS, alpha, thr, Hmax, T, obj0, None)),
                "lagwk": ("LAG-WK", lambda: fo.run_shared(nv.FO_LAG_WK, alpha, thr / 10, Hmax, T, obj0, None)),
                "iag": ("cyclic IAG", lambda: fo.run_shared(nv.FO_IAG_CYCLIC, alpha / N, 0.0, Hmax, T, obj0, None)),
                "dgd": ("DGD", lambda: fo.run_dgd(alpha / 100, T, obj0)),
                "dualavg": ("dual averaging (faithful Gauss-Seidel order)",
                            lambda: fo.run_dualavg(alpha, T, obj0, 0.0)),
                "dualavg_jacobi": ("dual averaging (Jacobi order)",
                                   lambda: fo.run_dualavg(alpha, T, obj0, 0.0, jacobi=True))}
        for key in fo_list:
            name, fn = runs[key]
            r = fn()
            tv = r[3] if key.startswith("dualavg") else r[2]
            tvt = torch.as_tensor(np.asarray(tv), device=dev)
            if world > 1:
                torch.distributed.all_reduce(tvt, op=torch.distributed.ReduceOp.MAX)
            extra = {"step": alpha}
            if key.startswith("lag"):
                extra["uploads"] = float(r[3][-1])
                extra["uploads_if_every_worker_every_iter"] = float(N * len(r[1]))
            record(D, a.out, name, "b200", r[1], tvt.cpu().numpy(), Nl * packed_bytes, hbm, extra)
        del fo, P
        free()
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
