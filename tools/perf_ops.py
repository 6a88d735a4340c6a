#!/usr/bin/env python
"""Device-timed micro-benchmarks of the native ops (CUDA events, warm-up, L2-exceeding inputs).
Writes one JSON document; copy the numbers you want judged into profiles/."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gadmm_b200.ops import linalg as LA  # noqa: E402
from gadmm_b200.ops import native as nv  # noqa: E402


def timed(fn, warm=2, rep=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(rep):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    only = sys.argv[1] if len(sys.argv) > 1 else "all"
    dev = torch.device("cuda", 0)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                            "MEASURED_PEAKS.json")))
    except Exception:
        pass
    out = {"peaks": peaks, "results": []}
    g = torch.Generator(device=dev)
    g.manual_seed(0)
    if only in ("all", "gram"):
        for (N, s, d) in ((8, 4096, 4096), (2, 32768, 4096), (1, 131072, 2048)):
            X32 = torch.randn((N, s, d), dtype=torch.float32, device=dev, generator=g)
            G = torch.empty((N, d, d), dtype=torch.float64, device=dev)
            ms = timed(lambda: LA.gram_tc(X32, out=G))
            flop = N * s * d * (d + 128) * 1.0     # lower triangle of 128-tiles: ~s*d^2 useful flop
            full = 2.0 * N * s * d * d
            out["results"].append({"op": "gram_tcgen05 (split-tf32 x3, incl. split/transpose prep)",
                                   "N": N, "s": s, "d": d, "ms": ms,
                                   "useful_TFLOPs_fp32_equiv": flop / ms / 1e9,
                                   "full_matrix_equiv_TFLOPs": full / ms / 1e9,
                                   "tf32_mma_TFLOPs_issued": 3 * flop / ms / 1e9})
            if N * s * d <= 8 * 4096 * 4096:
                Xd = X32.double()
                y = torch.zeros((N, s), dtype=torch.float64, device=dev)
                ms64 = timed(lambda: LA.gram(Xd, y, d), warm=1, rep=3)
                ref = Xd[0].T @ Xd[0]
                err = float(((G[0] - ref).abs().max() / ref.abs().max()).item())
                out["results"].append({"op": "gram_f64 (SIMT fp64)", "N": N, "s": s, "d": d,
                                       "ms": ms64, "useful_TFLOPs": flop / ms64 / 1e9,
                                       "gram_tc_rel_err_vs_fp64": err})
                del Xd
            del X32, G
    if only in ("all", "inverse"):
        for (N, d) in ((16, 4096), (64, 1024)):
            A = torch.randn((N, d, d), dtype=torch.float64, device=dev, generator=g)
            Gm = A @ A.transpose(1, 2) / d
            del A
            ms = timed(lambda: LA.shifted_inverse(Gm, d, 1.0), warm=1, rep=3)
            out["results"].append({"op": "spd_inverse_f64 (blocked Cholesky + trtri + lauum)", "N": N,
                                   "d": d, "ms": ms, "TFLOPs_fp64": N * float(d) ** 3 / ms / 1e9})
            del Gm
    if only in ("all", "symv"):
        for store, name in ((nv.STORE_F64, "f64"), (nv.STORE_F32, "f32")):
            N, d = 256, 4096
            P = torch.randn((N, nv.packed_elems(d)), dtype=torch.float64 if store == nv.STORE_F64
                            else torch.float32, device=dev, generator=g)
            x = torch.randn((N, d), dtype=torch.float64, device=dev, generator=g)
            mv = LA.BatchedSymv(P, d)
            y = torch.empty_like(x)
            ms = timed(lambda: mv(x, out=y))
            by = P.numel() * P.element_size()
            out["results"].append({"op": f"symv_batched packed {name}", "N": N, "d": d, "ms": ms,
                                   "GBps": by / ms / 1e6,
                                   "frac_of_measured_hbm": by / ms / 1e6 / peaks.get("hbm_gbs", 6650.0)})
            del P
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
