"""Scaled synthetic problems generated directly on the owning GPU (BASELINE.json configs 2-5).

``X_n`` never exists as a whole: each batch of workers draws its shard, reduces it to the Gram
``G_n`` / ``b_n`` / ``0.5 y'y`` and is discarded (T1 factors are 68.7 GB,
the raw data would not fit beside them).  Data model: ``X_n = Z_n / sqrt(s)`` with Gaussian ``Z``,
``y_n = X_n theta_true + noise`` -- random-init, synthetic, same shapes as the named config.
"""
from __future__ import annotations

import numpy as np
import torch

from ..ops import linalg as LA
from ..ops import native as nv
from .chain import ChainEngine


def build_scaled_linreg(eng: ChainEngine, s: int, seed: int = 0, noise: float = 0.1,
                        batch: int = 0, verbose: bool = False):
    """Fill ``eng`` with a synthetic least-squares problem; returns ``obj0`` (global optimum).

    The optimum needs ``sum_n G_n`` and ``sum_n b_n``; they are accumulated per rank and summed
    with one NCCL all-reduce (plumbing, outside any timed region)."""
    dev, d, dp = eng.dev, eng.d, eng.dp
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)
    theta_true = torch.randn(d, dtype=torch.float64, device=dev, generator=gen)
    if batch <= 0:
        batch = 16       # fixed: the per-batch generator seed defines the data set
    Gsum = torch.zeros((dp, dp), dtype=torch.float64, device=dev)
    bsum = torch.zeros((dp,), dtype=torch.float64, device=dev)
    hsum = torch.zeros((), dtype=torch.float64, device=dev)
    scale = 1.0 / np.sqrt(s)
    for k0 in range(0, eng.Nloc, batch):
        nb = min(batch, eng.Nloc - k0)
        gen.manual_seed(seed * 1000003 + int(eng.gids[k0]))
        X = torch.randn((nb, s, d), dtype=torch.float32, device=dev, generator=gen).double()
        X *= scale
        y = torch.einsum("nsd,d->ns", X, theta_true)
        y += noise * torch.randn((nb, s), dtype=torch.float64, device=dev, generator=gen)
        G, b, hyy = LA.gram(X, y, dp)
        del X
        Gsum += G.sum(0)
        bsum += b.sum(0)
        hsum += hyy.sum()
        eng.ingest_gram(G, b, hyy, k0)
        del G
        if verbose and eng.dist.rank == 0:
            print(f"  setup: workers {k0 + nb}/{eng.Nloc}", flush=True)
    eng.dist.all_reduce_sum_(Gsum)
    eng.dist.all_reduce_sum_(bsum)
    eng.dist.all_reduce_sum_(hsum)
    A = LA.shifted_inverse(Gsum.unsqueeze(0), d, 0.0)
    P = LA.pack_sym(A, d)
    th = LA.BatchedSymv(P, dp)(bsum.unsqueeze(0))[0]
    obj0 = 0.5 * float(th @ (Gsum @ th)) - float(bsum @ th) + float(hsum)
    eng.obj0 = obj0
    eng._desc = None
    return obj0
