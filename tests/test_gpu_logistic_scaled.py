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


def test_graph_loops_probe():
    """Conditional (WHILE) graph nodes work on this driver: the engine's device-side loops need them."""
    from gadmm_b200.engine.logistic_scaled import graph_loops_available
    from gadmm_b200.ops import native as nv
    assert nv.lib().gadmm_graph_loops_selftest(7, nv.stream_ptr()) == 0, nv.lib().gadmm_last_error()
    assert graph_loops_available()


@pytest.mark.parametrize("d,s,G,stride", [(8192, 21, 3, 1), (4096, 40, 5, 2), (68, 9, 2, 1), (1024, 64, 8, 1)])
def test_xtwx_kernel_fp64_shards(d, s, G, stride):
    """fp64 shards: user data is never rounded to fp32 (group_ADMM_logistic.m computes in double)."""
    from gadmm_b200.ops import native as nv
    dev = torch.device("cuda")
    gen = torch.Generator(device=dev).manual_seed(7 * d + s)
    Ntot = 4
    N = (Ntot + stride - 1) // stride
    X = torch.randn((Ntot, s, d), dtype=torch.float64, device=dev, generator=gen) / np.sqrt(d)
    y = torch.where(torch.randn((Ntot, s), device=dev, generator=gen) > 0, 1.0, -1.0).double()
    x = 3.0 * torch.randn((N, d), dtype=torch.float64, device=dev, generator=gen)
    Dw = torch.zeros((Ntot, s), dtype=torch.float64, device=dev)
    scratch = torch.empty((int(nv.lib().gadmm_logistic_xtwx_scratch_elems(N, d, G)),),
                          dtype=torch.float64, device=dev)

    def run(vec, hv):
        v = torch.empty((N, d), dtype=torch.float64, device=dev)
        loss = torch.zeros((N,), dtype=torch.float64, device=dev)
        nv.check(nv.lib().gadmm_logistic_xtwx_ex(
            nv.ptr(X), nv.STORE_F64, nv.ptr(y), nv.ptr(vec), C.c_longlong(d), nv.ptr(Dw), nv.ptr(v), d,
            nv.ptr(loss), N, s, d, G, hv, stride, nv.ptr(scratch), nv.stream_ptr()))
        return v, loss

    Xa, ya = X[::stride], y[::stride]
    z = torch.einsum("nsd,nd->ns", Xa, x) * ya
    sig = torch.sigmoid(-z)
    v, loss = run(x, 0)
    torch.testing.assert_close(v, -torch.einsum("nsd,ns->nd", Xa, ya * sig), rtol=1e-12, atol=1e-13)
    torch.testing.assert_close(loss, torch.nn.functional.softplus(-z).sum(1), rtol=1e-12, atol=1e-12)
    p = torch.randn_like(x)
    hv, _ = run(p, 1)
    u = torch.einsum("nsd,nd->ns", Xa, p) * (sig * (1 - sig))
    torch.testing.assert_close(hv, torch.einsum("nsd,ns->nd", Xa, u), rtol=1e-12, atol=1e-13)
    hv2, _ = run(p, 2)                       # hv=2 falls back to the exact product on fp64 shards
    assert torch.equal(hv, hv2)


@pytest.mark.parametrize("mode", ["graph", "host"])
def test_scaled_logistic_gadmm_matches_oracle(mode):
    """Unrounded double data (fp64 shards are chosen automatically); the device-loop graph and its
    host-driven twin run the same kernels and give the same curve."""
    from gadmm_b200.engine.logistic_scaled import ScaledLogisticGadmm
    from gadmm_b200.ops import native as nv
    rng = np.random.default_rng(4)
    N, s, d, lam, rho = 7, 24, 16, 1e-3, 5.0
    X = rng.standard_normal((N, s, d))
    y = np.where(rng.standard_normal((N, s)) > 0, 1.0, -1.0)
    XX, YY = X.reshape(N * s, d), y.reshape(-1)
    obj0 = O.opt_sol_logistic(XX, YY, d, lam, N)
    eng = ScaledLogisticGadmm(N, d, s, lam, rho, data=(X, y))
    assert eng.store == nv.STORE_F64
    o2, th = eng.solve_optimum()
    assert abs(o2 - obj0) <= 1e-9 * abs(obj0)
    a = O.group_ADMM_logistic(XX, YY, rho, N, d, s, 300, obj0, lam, 1e-3, naive=False, return_state=True)
    b = eng.run(300, obj0, 1e-3, mode=mode)
    assert a[2] == b[2] and a[2] < 300
    np.testing.assert_allclose(a[0], b[0], rtol=1e-8)
    np.testing.assert_allclose(a[4]["theta"], eng.state()["theta"], rtol=1e-5, atol=1e-7)
    assert b[3].shape == (b[2],) and np.all(np.diff(b[3]) > 0)
    assert eng.stats["hessvecs"] > 0 and eng.stats["newton_steps"] > 0


def test_scaled_logistic_checkpoint_resume():
    """state() after 15 iterations -> load_state(): the resumed curve continues the uninterrupted one."""
    from gadmm_b200.engine.logistic_scaled import ScaledLogisticGadmm
    rng = np.random.default_rng(14)
    N, s, d, lam, rho = 6, 20, 12, 1e-3, 3.0
    X = rng.standard_normal((N, s, d))
    y = np.where(rng.standard_normal((N, s)) > 0, 1.0, -1.0)
    full = ScaledLogisticGadmm(N, d, s, lam, rho, data=(X, y)).run(40, 0.0, 0.0)
    e1 = ScaledLogisticGadmm(N, d, s, lam, rho, data=(X, y))
    e1.run(15, 0.0, 0.0)
    st = e1.state()
    assert st["iteration"] == 15
    e2 = ScaledLogisticGadmm(N, d, s, lam, rho, data=(X, y))
    e2.load_state(st)
    rest = e2.run(40, 0.0, 0.0)
    assert len(rest[0]) == 40
    np.testing.assert_allclose(rest[0], full[0], rtol=1e-10)


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
    XX = rng.standard_normal((N * s, d))              # double data: nothing is rounded to fp32
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
