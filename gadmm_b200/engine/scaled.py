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
                        batch: int = 0, verbose: bool = False, gram: str = "f64",
                        s_chunk: int = 32768):
    """Fill ``eng`` with a synthetic least-squares problem; returns ``obj0`` (global optimum).

    The optimum needs ``sum_n G_n`` and ``sum_n b_n``; they are accumulated per rank and summed
    with one NCCL all-reduce (plumbing, outside any timed region)."""
    if gram == "tc":
        return _build_scaled_tc(eng, s, seed, noise, verbose, s_chunk)
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


def _build_scaled_tc(eng: ChainEngine, s: int, seed: int, noise: float, verbose: bool,
                     s_chunk: int):
    """Tensor-core path for sample-heavy shards (BASELINE.json config 5: 1M samples x d=16384 per
    worker): the shard is drawn chunk by chunk in fp32 with the Philox kernel, every chunk is
    folded into ``G_n`` by the tcgen05 split-tf32 Gram (``accumulate=True``) and dropped."""
    from ..ops.topology import philox_normal
    dev, d, dp = eng.dev, eng.d, eng.dp
    theta_true = philox_normal((d,), seed=1234, stream_id=7, dtype=torch.float32, device=dev)
    Gsum = torch.zeros((dp, dp), dtype=torch.float64, device=dev)
    bsum = torch.zeros((dp,), dtype=torch.float64, device=dev)
    hsum = torch.zeros((), dtype=torch.float64, device=dev)
    sd = 1.0 / np.sqrt(s)
    nb_max = max(1, min(eng.Nloc, int(2e9 // (min(s, s_chunk) * d * 4))))
    for k0 in range(0, eng.Nloc, nb_max):
        nb = min(nb_max, eng.Nloc - k0)
        G = torch.zeros((nb, dp, dp), dtype=torch.float64, device=dev)
        b = torch.zeros((nb, dp), dtype=torch.float64, device=dev)
        hyy = torch.zeros((nb,), dtype=torch.float64, device=dev)
        for c0 in range(0, s, s_chunk):
            sc = min(s_chunk, s - c0)
            X = philox_normal((nb, sc, d), seed=seed * 1000003 + int(eng.gids[k0]),
                              stream_id=1 + c0 // s_chunk, sd=sd, dtype=torch.float32, device=dev)
            y = torch.matmul(X, theta_true)
            y += noise * philox_normal((nb, sc), seed=seed * 1000003 + int(eng.gids[k0]),
                                       stream_id=100000 + c0 // s_chunk, dtype=torch.float32,
                                       device=dev)
            LA.gram_tc(X, dp, out=G, accumulate=True)
            b[:, :d] += torch.einsum("nsd,ns->nd", X, y).double()
            hyy += 0.5 * (y.double() ** 2).sum(1)
            del X, y
        Gsum += G.sum(0)
        bsum += b.sum(0)
        hsum += hyy.sum()
        eng.ingest_gram(G, b, hyy, k0)
        del G
        if verbose and eng.dist.rank == 0:
            print(f"  setup(tc): workers {k0 + nb}/{eng.Nloc}", flush=True)
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


def build_scaled_first_order(N: int, d: int, s: int, seed: int = 0, noise: float = 0.1, dev=None,
                             store: str = "f64", batch: int = 16):
    """The data set of :func:`build_scaled_linreg` (same per-batch generator seeds) reduced to packed
    Gram matrices for the first-order baselines (GD / DGD / LAG / IAG / dual averaging at the scaled
    shape, BASELINE.json config 2).  Single GPU: the packed Grams of 1024 x 4096 take 69.8 GB.
    Returns ``(FirstOrder engine, obj0)``."""
    from .first_order import FirstOrder
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    dp = nv.padded(d)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)
    theta_true = torch.randn(d, dtype=torch.float64, device=dev, generator=gen)
    st = nv.STORE_F64 if store == "f64" else nv.STORE_F32
    P = torch.empty((N, nv.packed_elems(dp)), dtype=torch.float64 if store == "f64" else torch.float32,
                    device=dev)
    b = torch.zeros((N, dp), dtype=torch.float64, device=dev)
    hyy = torch.zeros((N,), dtype=torch.float64, device=dev)
    Gsum = torch.zeros((dp, dp), dtype=torch.float64, device=dev)
    scale = 1.0 / np.sqrt(s)
    for k0 in range(0, N, batch):
        nb = min(batch, N - k0)
        gen.manual_seed(seed * 1000003 + k0)
        X = torch.randn((nb, s, d), dtype=torch.float32, device=dev, generator=gen).double()
        X *= scale
        y = torch.einsum("nsd,d->ns", X, theta_true)
        y += noise * torch.randn((nb, s), dtype=torch.float64, device=dev, generator=gen)
        G, bb, hh = LA.gram(X, y, dp)
        del X
        LA.pack_sym(G, d, st, out=P[k0:k0 + nb])
        b[k0:k0 + nb], hyy[k0:k0 + nb] = bb, hh
        Gsum += G.sum(0)
        del G
    A = LA.shifted_inverse(Gsum.unsqueeze(0), d, 0.0)
    th = LA.BatchedSymv(LA.pack_sym(A, d), dp)(b.sum(0).unsqueeze(0))[0]
    bs = b.sum(0)
    obj0 = 0.5 * float(th @ (Gsum @ th)) - float(bs @ th) + float(hyy.sum())
    return FirstOrder.from_parts(P, b, hyy, Gsum, N, d, dev), obj0


# ------------------------------------------------------------------------------------------
# One data set for every algorithm and every partition (BASELINE.json config 2: GADMM vs star ADMM
# vs GD / DGD / LAG / dual averaging on the same 1024 x 4096 problem): worker g's shard comes from
# Philox stream g, so chain segments, star worker ranges and parameter-server ranges over any
# number of GPUs all see the same X_g, y_g.
def scaled_theta_true(d: int, seed: int, dev):
    from ..ops.topology import philox_normal
    return philox_normal((d,), seed * 7919 + 17, 0xFFFF, 0.0, 1.0, device=dev)


def worker_grams(gids, d: int, s: int, dp: int, seed: int = 0, noise: float = 0.1, dev=None,
                 batch: int = 16):
    """Yield ``(k0, G, b, hyy)`` for consecutive batches of the workers ``gids`` (global ids):
    ``X_g = Z_g / sqrt(s)`` (fp32 draws widened to fp64), ``y_g = X_g theta* + noise``."""
    from ..ops.topology import philox_normal
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    th = scaled_theta_true(d, seed, dev)
    gids = [int(g) for g in gids]
    sd = 1.0 / np.sqrt(s)
    X32 = torch.empty((s, d), dtype=torch.float32, device=dev)
    for k0 in range(0, len(gids), batch):
        ids = gids[k0:k0 + batch]
        X = torch.empty((len(ids), s, d), dtype=torch.float64, device=dev)
        y = torch.empty((len(ids), s), dtype=torch.float64, device=dev)
        for i, g in enumerate(ids):
            philox_normal(None, seed, g, 0.0, sd, out=X32)
            X[i].copy_(X32)
            y[i] = X[i] @ th + philox_normal((s,), seed + 1, g, 0.0, noise, device=dev)
        G, b, hyy = LA.gram(X, y, dp)
        del X
        yield k0, G, b, hyy


def scaled_optimum(N: int, d: int, s: int, dp: int, dist, seed: int = 0, noise: float = 0.1,
                   batch: int = 16):
    """``(obj0, Gsum, bsum, hsum)`` of the data set above: summed Gram over all workers (each rank
    sums its share, one all-reduce), normal equations in fp64 (K12)."""
    from ..parallel.runtime import partition
    dev = dist.device
    own = np.nonzero(partition(N, dist.world) == dist.rank)[0]
    Gsum = torch.zeros((dp, dp), dtype=torch.float64, device=dev)
    bsum = torch.zeros((dp,), dtype=torch.float64, device=dev)
    hsum = torch.zeros((), dtype=torch.float64, device=dev)
    for _, G, b, hyy in worker_grams(own, d, s, dp, seed, noise, dev, batch):
        Gsum += G.sum(0)
        bsum += b.sum(0)
        hsum += hyy.sum()
    dist.all_reduce_sum_(Gsum)
    dist.all_reduce_sum_(bsum)
    dist.all_reduce_sum_(hsum)
    A = LA.shifted_inverse(Gsum.unsqueeze(0).clone(), d, 0.0)
    th = LA.BatchedSymv(LA.pack_sym(A, d), dp)(bsum.unsqueeze(0))[0]
    obj0 = 0.5 * float(th @ (Gsum @ th)) - float(bsum @ th) + float(hsum)
    return obj0, Gsum, bsum, hsum
