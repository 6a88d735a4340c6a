"""BASELINE.json config 3 (logistic GADMM, local Newton solves, 512 workers x d = 8192) and the
throughput of its building block.  Under torchrun the chain is split over the GPUs.

    python benchmarks/logistic_scaled_bench.py --workers 512 --dim 8192 --samples 1024 --iters 20
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=512)
    ap.add_argument("--dim", dest="d", type=int, default=8192)
    ap.add_argument("--samples", dest="s", type=int, default=1024)
    ap.add_argument("--rho", type=float, default=0.05)
    ap.add_argument("--lam", type=float, default=1e-3)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--acc", type=float, default=1e-4)
    ap.add_argument("--rel-acc", type=float, default=0.0, help="stop at gap < rel_acc * obj0 instead of --acc")
    ap.add_argument("--quiet", action="store_true")
    ap.add_argument("--kernel-only", action="store_true")
    ap.add_argument("--mode", default="auto", choices=["auto", "graph", "host"])
    a = ap.parse_args()
    import torch
    from gadmm_b200.engine.logistic_scaled import ScaledLogisticGadmm
    from gadmm_b200.parallel.runtime import Dist, init_process_group
    init_process_group()
    D = Dist()
    t0 = time.time()
    eng = ScaledLogisticGadmm(a.workers, a.d, a.s, a.lam, a.rho, dist=D, max_iter=a.iters)
    torch.cuda.synchronize()
    out = {"workers": a.workers, "d": a.d, "s": a.s, "n_gpus": D.world, "rho": a.rho, "lam": a.lam,
           "setup_s": round(time.time() - t0, 3)}
    # ---- building block: one Hessian product over one colour / over all local shards
    for label, first, count, stride in (("colour", 0, (eng.Nloc + 1) // 2, 2), ("all", 0, eng.Nloc, 1)):
        x = torch.randn((count, a.d), dtype=torch.float64, device=eng.dev)
        eng._xtwx(x, first, count, stride, 0)
        for hv in (0, 1, 2):
            for _ in range(2):
                eng._xtwx(x, first, count, stride, hv)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            e0.record()
            for _ in range(reps):
                eng._xtwx(x, first, count, stride, hv)
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1) / reps
            gb = count * a.s * a.d * 4 / 1e9
            out[f"xtwx_{label}_{('grad', 'hv', 'hvfast')[hv]}_ms"] = round(ms, 4)
            out[f"xtwx_{label}_{('grad', 'hv', 'hvfast')[hv]}_GBps"] = round(gb / ms * 1e3, 1)
    if not a.kernel_only:
        t0 = time.time()
        obj0, _ = eng.solve_optimum()
        torch.cuda.synchronize()
        out["optimum_s"] = round(time.time() - t0, 3)
        out["obj0"] = obj0
        acc = a.rel_acc * abs(obj0) if a.rel_acc > 0 else a.acc
        obj, loss, it, tv = eng.run(a.iters, obj0, acc, verbose=not a.quiet, mode=a.mode)
        st = eng.state()
        out.update({"iterations": int(it), "acc": acc, "reached": bool(loss[-1] < acc), "final_gap": float(loss[-1]),
                    "time_s": float(tv[-1]), "s_per_iter": float(tv[-1] / max(len(tv), 1)),
                    "first_iter_s": float(tv[0]),
                    "s_per_iter_steady": float((tv[-1] - tv[0]) / max(len(tv) - 1, 1)),
                    "hessvecs": st.get("hessvecs"), "evals": st.get("evals"), "newton_steps": st.get("newton_steps"),
                    "mode": st.get("mode"), "launches": st["launches"]})
        per_rank = [None] * D.world
        mine = {"rank": D.rank, "hessvecs": st.get("hessvecs"), "evals": st.get("evals"),
                "newton_steps": st.get("newton_steps"), "wait_ms": st.get("wait_ms")}
        if D.world > 1:
            torch.distributed.all_gather_object(per_rank, mine)
        else:
            per_rank = [mine]
        out["per_rank"] = per_rank
    if D.rank == 0:
        print(json.dumps(out))
    if D.world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
