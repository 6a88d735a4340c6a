"""One-pass X'w(Xx) kernel and the Newton-CG logistic GADMM vs fp64 references."""
import ctypes as C

import numpy as np
import pytest
import torch

from gadmm_b200 import oracle as O

pytestmark = pytest.mark.gpu


def _xtwx(X, y, x, Dw, N, s, d, G, hv, stride, shared):
    from gadmm_b200.ops import native as nv
    v = torch.empty((N, d), dtype=torch.float64, device=X.device)
    loss = torch.zeros((N,), dtype=torch.float64, device=X.device)
    scratch = torch.empty((int(nv.lib().gadmm_logistic_xtwx_scratch_elems(N, d, G)),),
                          dtype=torch.float64, device=X.device)
    nv.check(nv.lib().gadmm_logistic_xtwx(
        nv.ptr(X), nv.ptr(y), nv.ptr(x), C.c_longlong(0 if shared else d), nv.ptr(Dw), nv.ptr(v), d,
        nv.ptr(loss), N, s, d, G, hv, stride, nv.ptr(scratch), nv.stream_ptr()))
    return v, loss


@pytest.mark.parametrize("d,s,G,stride,shared", [(8192, 37, 3, 1, False), (4096, 64, 5, 2, False),
                                                 (260, 50, 2, 1, True), (64, 9, 4, 2, False),
                                                 (1024, 128, 16, 1, False)])
def test_xtwx_kernel_matches_fp64(d, s, G, stride, shared):
    dev = torch.device("cuda")
    gen = torch.Generator(device=dev).manual_seed(d + s)
    Ntot = 6
    N = (Ntot + stride - 1) // stride
    X = torch.randn((Ntot, s, d), dtype=torch.float32, device=dev, generator=gen) / np.sqrt(d)
    y = torch.where(torch.randn((Ntot, s), device=dev, generator=gen) > 0, 1.0, -1.0).double()
    x = 3.0 * torch.randn((1 if shared else N, d), dtype=torch.float64, device=dev, generator=gen)
    Dw = torch.zeros((Ntot, s), dtype=torch.float64, device=dev)
    Xa, ya = X[::stride].double(), y[::stride]
    xa = x.expand(N, d)
    z = torch.einsum("nsd,nd->ns", Xa, xa) * ya
    sig = torch.sigmoid(-z)
    g_ref = -torch.einsum("nsd,ns->nd", Xa, ya * sig)
    v, loss = _xtwx(X, y, x, Dw, N, s, d, G, 0, stride, shared)
    torch.testing.assert_close(v, g_ref, rtol=1e-11, atol=1e-12)
    torch.testing.assert_close(loss, torch.nn.functional.softplus(-z).sum(1), rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(Dw[::stride], sig * (1 - sig), rtol=1e-11, atol=1e-14)
    if stride == 2:
        assert float(Dw[1::2].abs().max()) == 0.0          # the other colour's state is untouched
    p = torch.randn_like(x)
    hv, _ = _xtwx(X, y, p, Dw, N, s, d, G, 1, stride, shared)
    u = torch.einsum("nsd,nd->ns", Xa, p.expand(N, d)) * (sig * (1 - sig))
    torch.testing.assert_close(hv, torch.einsum("nsd,ns->nd", Xa, u), rtol=1e-11, atol=1e-13)
    hv_fast, _ = _xtwx(X, y, p, Dw, N, s, d, G, 2, stride, shared)      # fp32 multiplies
    err = float((hv_fast - hv).norm() / hv.norm())
    assert err < 5e-6, err
    # deterministic: same bits on a second call
    hv2, _ = _xtwx(X, y, p, Dw, N, s, d, G, 1, stride, shared)
    assert torch.equal(hv, hv2)


def test_scaled_logistic_gadmm_matches_oracle():
    from gadmm_b200.engine.logistic_scaled import ScaledLogisticGadmm
    rng = np.random.default_rng(4)
    N, s, d, lam, rho = 7, 24, 16, 1e-3, 5.0
    X = rng.standard_normal((N, s, d)).astype(np.float32)
    y = np.where(rng.standard_normal((N, s)) > 0, 1.0, -1.0)
    XX, YY = X.reshape(N * s, d).astype(np.float64), y.reshape(-1)
    obj0 = O.opt_sol_logistic(XX, YY, d, lam, N)
    eng = ScaledLogisticGadmm(N, d, s, lam, rho, data=(X, y))
    o2, th = eng.solve_optimum()
    assert abs(o2 - obj0) <= 1e-9 * abs(obj0)
    a = O.group_ADMM_logistic(XX, YY, rho, N, d, s, 300, obj0, lam, 1e-3, naive=False, return_state=True)
    b = eng.run(300, obj0, 1e-3)
    assert a[2] == b[2] and a[2] < 300
    np.testing.assert_allclose(a[0], b[0], rtol=1e-8)
    np.testing.assert_allclose(a[4]["theta"], eng.state()["theta"], rtol=1e-5, atol=1e-7)
    assert b[3].shape == (b[2],) and np.all(np.diff(b[3]) > 0)


def test_scaled_logistic_synthetic_converges():
    from gadmm_b200.engine.logistic_scaled import gadmm_logistic_scaled
    obj, loss, it, t, st = gadmm_logistic_scaled(8, 512, 64, rho=0.05, lam_log=1e-3, num_iter=200,
                                                 acc=1e-4, return_state=True)
    assert it < 200 and loss[-1] < 1e-4
    assert st["hessvecs"] > 0 and st["launches"] > 0


def test_api_routes_to_cg_solver():
    import gadmm_b200 as g
    rng = np.random.default_rng(9)
    N, s, d, lam, rho = 5, 20, 10, 1e-3, 5.0            # d % 4 != 0: zero-padded internally
    XX = rng.standard_normal((N * s, d)).astype(np.float32).astype(np.float64)
    YY = np.where(rng.standard_normal(N * s) > 0, 1.0, -1.0)
    obj0 = O.opt_sol_logistic(XX, YY, d, lam, N)
    args = (XX, YY, rho, N, d, s, 300, obj0, lam, 1e-3)
    a = O.group_ADMM_logistic(*args, naive=False, return_state=True)
    b = g.group_ADMM_logistic(*args, backend="b200", solver="cg", return_state=True)
    c = g.group_ADMM_logistic(*args, backend="b200", solver="dense")
    assert a[2] == b[2] == c[2] and a[2] < 300
    np.testing.assert_allclose(a[0], b[0], rtol=1e-8)
    assert b[4]["theta"].shape == (d, N)
    np.testing.assert_allclose(a[4]["theta"], b[4]["theta"], rtol=1e-5, atol=1e-7)
