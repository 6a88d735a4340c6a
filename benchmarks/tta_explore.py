#!/usr/bin/env python
"""Time-to-1e-4 exploration at the scaled shape: static GADMM vs D-GADMM (device re-chaining)."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=128)
    ap.add_argument("--d", type=int, default=4096)
    ap.add_argument("--s", type=int, default=1024)
    ap.add_argument("--cap", type=int, default=8000)
    ap.add_argument("--rhos", default="0.3,1,3")
    ap.add_argument("--taus", default="0,10")
    ap.add_argument("--noise", type=float, default=0.1)
    ap.add_argument("--impl", default="b200", choices=["b200", "nccl"])
    ap.add_argument("--gram", default="f64")
    ap.add_argument("--device-chains", action="store_true")
    a = ap.parse_args()
    import torch
    from gadmm_b200.parallel.runtime import Dist, init_process_group
    from gadmm_b200.engine.chain import ChainEngine
    from gadmm_b200.engine.scaled import build_scaled_linreg
    from gadmm_b200.topology import findPath2
    rank, world, local = init_process_group("nccl")
    torch.cuda.set_device(local)
    D = Dist()
    N = a.workers
    for rho in [float(x) for x in a.rhos.split(",")]:
        for tau in [int(x) for x in a.taus.split(",")]:
            t0 = time.time()
            if a.impl == "nccl":
                from gadmm_b200.parallel import nccl_baseline as NB
                eng = NB.NcclChainGadmm(N, a.d, rho, 0.0, 1e-4, a.cap, dist=D, dynamic=(tau > 0))
                obj0 = NB.build_scaled_linreg(eng, a.s, seed=0, noise=a.noise)
            else:
                eng = ChainEngine(N, a.d, rho, 0.0, 1e-4, a.cap, dist=D, dynamic=(tau > 0), stop=True)
                obj0 = build_scaled_linreg(eng, a.s, seed=0, noise=a.noise, gram=a.gram)
            eng.set_path(np.arange(N))
            torch.cuda.synchronize(); D.barrier()
            setup = time.time() - t0
            if tau > 0 and a.impl == "nccl":
                g = np.random.default_rng(11)
                it = eng.run(a.cap, rechain=lambda i: np.asarray(findPath2(N, g)[0]) - 1, coherence=tau)
            elif tau > 0:
                g = np.random.default_rng(11)
                def src(kk):
                    out = findPath2(N, g) if D.rank == 0 else None
                    out = D.broadcast_object(out)
                    return np.asarray(out[0]) - 1
                from gadmm_b200.ops.topology import find_path_batch_device
                def chunk_src(base, K):      # same seed on every rank -> same chains everywhere
                    return find_path_batch_device(K, N, seed=11, stream_id=base, device=D.device)[0]
                it = eng.run_device_rechain(a.cap, src, tau, window=64, chunk=128,
                                            chunk_source=chunk_src if a.device_chains else None)
            else:
                it = eng.run(a.cap, window=64)
            obj, loss, it, tv = eng.results(it)
            if D.rank == 0:
                print(json.dumps({"impl": a.impl, "N": N, "d": a.d, "s": a.s, "gpus": world, "rho": rho,
                                  "tau": tau or "static", "Iter": int(it), "reached": bool(loss[-1] < 1e-4),
                                  "time_s": float(tv[-1]), "ms_per_iter": 1e3 * float(tv[-1]) / max(it - 1, 1),
                                  "final_loss": float(loss[-1]), "obj0": obj0, "setup_s": setup}), flush=True)
            del eng
            torch.cuda.empty_cache()
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
