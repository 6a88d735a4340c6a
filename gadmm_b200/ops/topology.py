"""Device-side topology / random fills (``csrc/topology.cu``)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import native as nv


def philox_uniform(n: int, seed: int, stream_id: int = 0, scale: float = 1.0, device=None):
    device = device or torch.device("cuda", torch.cuda.current_device())
    out = torch.empty((n,), dtype=torch.float64, device=device)
    nv.check(nv.lib().gadmm_philox_uniform(nv.ptr(out), C.c_size_t(n), C.c_ulonglong(seed),
                                           C.c_uint(stream_id), C.c_double(scale), nv.stream_ptr()))
    return out


def philox_normal(shape, seed: int, stream_id: int = 0, mean: float = 0.0, sd: float = 1.0,
                  dtype=torch.float64, device=None, out=None):
    device = device or torch.device("cuda", torch.cuda.current_device())
    if out is None:
        out = torch.empty(shape, dtype=dtype, device=device)
    nv.check(nv.lib().gadmm_philox_normal(nv.ptr(out), C.c_size_t(out.numel()),
                                          1 if out.dtype == torch.float32 else 0,
                                          C.c_ulonglong(seed), C.c_uint(stream_id), C.c_double(mean),
                                          C.c_double(sd), nv.stream_ptr()))
    return out


def find_path_device(xy: torch.Tensor, model: int = 0, want_grid: bool = True):
    """Greedy chain on the GPU.  ``xy``: (N,2) float64 CUDA.  Returns the reference's tuples with a
    1-based path: model 0 -> ``(path, pathCost, d_square)``; model 1 -> ``(path, pathCost,
    d_square, P_central, center)`` (findPath.m:1, findPath2.m:1)."""
    N = xy.shape[0]
    dev = xy.device
    path = torch.empty((N,), dtype=torch.int32, device=dev)
    cost = torch.empty((max(N - 1, 1),), dtype=torch.float64, device=dev)
    d2 = torch.empty((N, N), dtype=torch.float64, device=dev) if want_grid else None
    Pc = torch.empty((N,), dtype=torch.float64, device=dev) if model == 1 else None
    cen = torch.zeros((1,), dtype=torch.int32, device=dev) if model == 1 else None
    nv.check(nv.lib().gadmm_find_path(nv.ptr(xy.contiguous()), N, model, nv.ptr(path), nv.ptr(cost),
                                      nv.ptr(d2), nv.ptr(Pc), nv.ptr(cen), nv.stream_ptr()))
    p1 = path.cpu().numpy().astype(np.int64) + 1
    c = cost[:N - 1].cpu().numpy()
    g = d2.cpu().numpy() if want_grid else None
    if model == 0:
        return p1, c, g
    return p1, c, g, Pc.cpu().numpy(), int(cen.item()) + 1


def find_path_batch_device(K: int, N: int, seed: int, stream_id: int = 0, model: int = 1,
                           device=None):
    """Draw K independent topologies with the Philox generator and build their greedy chains on
    the GPU (one CTA per chain).  Returns device tensors ``paths`` (K, N) int32, 0-based, and
    ``cost_sums`` (K,) float64 (sum of hop costs).  D-GADMM uses this to re-chain without the
    host: ``findPath2`` for 128 chains of 1024 nodes is ~1.7 s in NumPy and ~2 ms here."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    area = 250.0 if model == 1 else 50.0
    xy = philox_uniform(K * N * 2, seed=seed, stream_id=stream_id, scale=area, device=device)
    paths = torch.empty((K, N), dtype=torch.int32, device=device)
    cost = torch.zeros((K, max(N - 1, 1)), dtype=torch.float64, device=device)
    nv.check(nv.lib().gadmm_find_path_batch(nv.ptr(xy), K, N, model, nv.ptr(paths), nv.ptr(cost),
                                            nv.stream_ptr()))
    return paths, cost[:, :max(N - 1, 0)].sum(1) if N > 1 else cost.sum(1) * 0.0
