"""Sensitivity of the tcgen05 Gram to the K-segment length (epilogue share) -- diagnostic."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gadmm_b200.ops import linalg as LA


def main():
    dev = torch.device("cuda")
    out = []
    for (N, s, d) in ((2, 32768, 4096), (8, 4096, 4096)):
        X = torch.randn((N, s, d), dtype=torch.float32, device=dev)
        G = torch.empty((N, d, d), dtype=torch.float64, device=dev)
        for kseg in (2048, 4096, 8192):
            if kseg > s and kseg != 32768:
                continue
            for _ in range(2):
                LA.gram_tc(X, d, out=G, kseg=kseg)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                LA.gram_tc(X, d, out=G, kseg=kseg)
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1) / 3
            nb = d // 128
            issued = 3 * 2.0 * N * s * 128 * 128 * (nb * (nb + 1) / 2)
            out.append({"N": N, "s": s, "d": d, "kseg": kseg, "ms": round(ms, 3),
                        "issued_TFLOPs": round(issued / ms / 1e9, 1)})
            print(out[-1], flush=True)
    json.dump(out, open("gpurun_out/gram_kseg_probe.json", "w"))


if __name__ == "__main__":
    main()
