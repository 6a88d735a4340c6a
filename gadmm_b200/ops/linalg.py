"""Setup ops on the GPU: Gram, shifted SPD inverse, packing, packed-symmetric SYMV.

Thin, allocation-owning wrappers over the C ABI (``csrc/setup_fp64.cu``, ``csrc/gadmm_iter.cu``).
Numerics tests compare each against a plain PyTorch/NumPy fp64 reference (tests/test_gpu_ops.py).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import native as nv


def gram(X: torch.Tensor, y: torch.Tensor, dp: int = None):
    """``G_n = X_n' X_n`` (N,dp,dp), ``b_n = X_n' y_n`` (N,dp), ``0.5*y_n'y_n`` (N,) in fp64.

    X: (N, s, d) float64 CUDA, y: (N, s).  (K1, exact path; group_ADMM_closedForm.m:43 ``H'*H``.)
    """
    N, s, d = X.shape
    dp = nv.padded(d) if dp is None else dp
    X = X.contiguous()
    y = y.contiguous()
    G = torch.empty((N, dp, dp), dtype=torch.float64, device=X.device)
    b = torch.empty((N, dp), dtype=torch.float64, device=X.device)
    hyy = torch.empty((N,), dtype=torch.float64, device=X.device)
    nv.check(nv.lib().gadmm_gram_f64(nv.ptr(X), nv.ptr(y), N, s, d, dp, nv.ptr(G), nv.ptr(b),
                                     nv.ptr(hyy), nv.stream_ptr()))
    return G, b, hyy


def gram_tc(X: torch.Tensor, dp: int = None, w: torch.Tensor = None, out: torch.Tensor = None,
            accumulate: bool = False, kseg: int = 0):
    """``G_n (+)= X_n' diag(w_n) X_n`` on the tcgen05 tensor cores (split-tf32, fp32-grade accuracy,
    fp64 result).  X: (N, s, d) float64 or float32 CUDA; w: optional (N, s) float64 row weights."""
    N, s, d = X.shape
    dp = nv.padded(d) if dp is None else dp
    X = X.contiguous()
    L = nv.lib()
    L.gadmm_gram_tc_workspace_floats.restype = C.c_size_t
    nwork = int(L.gadmm_gram_tc_workspace_floats(N, s, d))
    work = torch.empty((nwork,), dtype=torch.float32, device=X.device)
    G = torch.empty((N, dp, dp), dtype=torch.float64, device=X.device) if out is None else out
    nv.check(L.gadmm_gram_tcgen05(nv.ptr(X), 1 if X.dtype == torch.float32 else 0,
                                  nv.ptr(w.contiguous() if w is not None else None), N, s, d, dp,
                                  nv.ptr(work), nv.ptr(G), 1 if accumulate else 0, int(kseg),
                                  nv.stream_ptr()))
    return G


def xty(X: torch.Tensor, y: torch.Tensor, dp: int):
    """``b_n = X_n' y_n`` and ``0.5 y'y`` in fp64 (companion of :func:`gram_tc`)."""
    N, s, d = X.shape
    Xd = X if X.dtype == torch.float64 else X.double()
    b = torch.zeros((N, dp), dtype=torch.float64, device=X.device)
    b[:, :d] = torch.einsum("nsd,ns->nd", Xd, y)
    return b, 0.5 * (y * y).sum(1)


def shifted_inverse(G: torch.Tensor, d: int, shift: float, out: torch.Tensor = None):
    """``(G_n + shift*I)^-1`` for a batch of padded Gram matrices; lower block-triangle is valid.

    Blocked Cholesky + triangular inverse + ``W'W`` in fp64 SIMT (K2).  Returns (N,dp,dp).
    """
    N, dp, _ = G.shape
    A = torch.empty_like(G) if out is None else out
    W = torch.empty_like(G)
    Dinv = torch.empty((N, dp // nv.TILE, nv.TILE, nv.TILE), dtype=torch.float64, device=G.device)
    L = nv.lib()
    nv.check(L.gadmm_shift_copy(nv.ptr(G), nv.ptr(A), N, d, dp, C.c_double(shift), nv.stream_ptr()))
    nv.check(L.gadmm_spd_inverse_f64(nv.ptr(A), nv.ptr(W), nv.ptr(Dinv), N, dp, nv.stream_ptr()))
    return A


def pack_sym(M: torch.Tensor, d: int, store: int = nv.STORE_F64, out: torch.Tensor = None):
    """Lower block-triangle of (N,dp,dp) -> (N, packed_elems) tiles in fp64 or fp32."""
    N, dp, _ = M.shape
    pe = nv.packed_elems(dp)
    dt = torch.float64 if store == nv.STORE_F64 else torch.float32
    P = torch.empty((N, pe), dtype=dt, device=M.device) if out is None else out
    nv.check(nv.lib().gadmm_pack_sym(nv.ptr(M), nv.ptr(P), N, d, dp, store, nv.stream_ptr()))
    return P


def ptr_table(P: torch.Tensor) -> torch.Tensor:
    """Device array of per-matrix base pointers for a (N, packed_elems) tensor."""
    N = P.shape[0]
    base, stride = P.data_ptr(), P.stride(0) * P.element_size()
    return torch.tensor([base + i * stride for i in range(N)], dtype=torch.int64, device=P.device)


class BatchedSymv:
    """y_n = S_n x_n for packed symmetric matrices (reusable scratch; no allocation per call)."""

    def __init__(self, P: torch.Tensor, dp: int):
        self.P, self.dp, self.N = P, dp, P.shape[0]
        self.store = nv.STORE_F64 if P.dtype == torch.float64 else nv.STORE_F32
        self.ptrs = ptr_table(P)
        self.scratch = torch.empty((self.N, nv.symv_scratch_elems(dp)), dtype=torch.float64,
                                   device=P.device)
        self.counters = torch.zeros((self.N,), dtype=torch.int32, device=P.device)

    def __call__(self, x: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
        y = torch.empty_like(x) if out is None else out
        nv.check(nv.lib().gadmm_symv_batched(nv.ptr(self.ptrs), nv.ptr(x), nv.ptr(y), self.N,
                                             self.dp, self.store, nv.ptr(self.scratch),
                                             nv.ptr(self.counters), nv.stream_ptr()))
        return y


def unpack_sym(P: torch.Tensor, dp: int) -> torch.Tensor:
    """Inverse of :func:`pack_sym` (host-side helper for tests): (N,packed) -> full (N,dp,dp)."""
    N = P.shape[0]
    nT = dp // nv.TILE
    out = torch.zeros((N, dp, dp), dtype=torch.float64, device=P.device)
    tiles = P.view(N, -1, nv.TILE, nv.TILE).to(torch.float64)
    t = 0
    for i in range(nT):
        for j in range(i + 1):
            blk = tiles[:, t]
            out[:, i * 64:(i + 1) * 64, j * 64:(j + 1) * 64] = blk
            if i != j:
                out[:, j * 64:(j + 1) * 64, i * 64:(i + 1) * 64] = blk.transpose(1, 2)
            t += 1
    return out
