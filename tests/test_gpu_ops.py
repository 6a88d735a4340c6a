"""Numerics of the native setup ops against plain fp64 PyTorch references (run on the B200)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


@pytest.mark.parametrize("N,s,d", [(3, 50, 50), (2, 37, 70), (1, 200, 130), (4, 16, 64)])
def test_gram_matches_torch(dev, N, s, d):
    from gadmm_b200.ops import linalg as LA
    g = torch.Generator(device="cpu").manual_seed(1)
    X = torch.randn(N, s, d, dtype=torch.float64, generator=g).to(dev)
    y = torch.randn(N, s, dtype=torch.float64, generator=g).to(dev)
    G, b, hyy = LA.gram(X, y)
    dp = G.shape[1]
    assert dp % 64 == 0 and dp >= d
    Gr = X.transpose(1, 2) @ X
    torch.testing.assert_close(G[:, :d, :d], Gr, rtol=1e-12, atol=1e-11)
    assert torch.count_nonzero(G[:, d:, :]) == 0 and torch.count_nonzero(G[:, :, d:]) == 0
    torch.testing.assert_close(b[:, :d], torch.einsum("nsd,ns->nd", X, y), rtol=1e-12, atol=1e-11)
    torch.testing.assert_close(hyy, 0.5 * (y * y).sum(1), rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("N,d,shift", [(3, 50, 2.0), (2, 70, 0.5), (2, 200, 6.0), (1, 320, 1.0)])
def test_shifted_inverse_and_pack_and_symv(dev, N, d, shift):
    from gadmm_b200.ops import linalg as LA
    from gadmm_b200.ops import native as nv
    g = torch.Generator(device="cpu").manual_seed(2)
    X = torch.randn(N, d + 5, d, dtype=torch.float64, generator=g).to(dev)
    y = torch.randn(N, d + 5, dtype=torch.float64, generator=g).to(dev)
    G, b, _ = LA.gram(X, y)
    dp = G.shape[1]
    A = LA.shifted_inverse(G, d, shift)
    ref = torch.linalg.inv(G[:, :d, :d] + shift * torch.eye(d, dtype=torch.float64, device=dev))
    for store, tol in ((nv.STORE_F64, 1e-10), (nv.STORE_F32, 2e-6)):
        P = LA.pack_sym(A, d, store)
        full = LA.unpack_sym(P, dp)
        scale = ref.abs().max()
        assert (full[:, :d, :d] - ref).abs().max() <= tol * scale
        assert torch.count_nonzero(full[:, d:, :]) == 0
        x = torch.zeros(N, dp, dtype=torch.float64, device=dev)
        x[:, :d] = torch.randn(N, d, dtype=torch.float64, generator=g).to(dev)
        symv = LA.BatchedSymv(P, dp)
        yv = symv(x)
        yr = torch.einsum("nij,nj->ni", full, x)
        torch.testing.assert_close(yv, yr, rtol=1e-12, atol=1e-12 * float(scale) * d)
        yv2 = symv(x)     # second call reuses scratch and re-armed counters: bit-identical
        assert torch.equal(yv, yv2)


def test_symv_larger_tiles(dev):
    """nT = 9 (odd, middle tile-row handled alone) with the streaming pipeline wrapping."""
    from gadmm_b200.ops import linalg as LA
    from gadmm_b200.ops import native as nv
    d = 9 * 64 - 3
    dp = nv.padded(d)
    g = torch.Generator(device="cpu").manual_seed(3)
    S = torch.randn(2, dp, dp, dtype=torch.float64, generator=g).to(dev)
    S = S + S.transpose(1, 2)
    for store in (nv.STORE_F64, nv.STORE_F32):
        P = LA.pack_sym(S, d, store)
        full = LA.unpack_sym(P, dp)
        x = torch.randn(2, dp, dtype=torch.float64, generator=g).to(dev)
        yv = LA.BatchedSymv(P, dp)(x)
        torch.testing.assert_close(yv, torch.einsum("nij,nj->ni", full, x), rtol=1e-11, atol=1e-9)
