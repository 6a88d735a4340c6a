"""GPU paths of the small set-up entry points: opt_sol, opt_sol_logistic, logReg_GD, and the
static-engine chain restriction."""
import numpy as np
import pytest

import gadmm_b200 as g
from gadmm_b200 import oracle as O

pytestmark = pytest.mark.gpu


def test_opt_sol_variants_on_gpu():
    rng = np.random.default_rng(2)
    XX = rng.standard_normal((90, 14))
    YY = rng.standard_normal(90)
    ref = O.opt_sol(XX, YY, 10)
    assert g.opt_sol(XX, YY, 10, backend="b200") == pytest.approx(ref, rel=1e-9)
    assert g.opt_sol(XX, YY, backend="b200") == pytest.approx(O.opt_sol(XX, YY), rel=1e-9)


@pytest.mark.parametrize("d", [10, 32])
def test_opt_sol_logistic_on_gpu(d):
    rng = np.random.default_rng(d)
    N, s, lam = 5, 30, 1e-3
    XX = rng.standard_normal((N * s, d))
    YY = np.where(rng.standard_normal(N * s) > 0, 1.0, -1.0)
    ref, th_ref = O.opt_sol_logistic(XX, YY, d, lam, N, return_theta=True)
    out, th = g.opt_sol_logistic(XX, YY, d, lam, N, backend="b200", return_theta=True)
    assert out == pytest.approx(ref, rel=1e-10)
    np.testing.assert_allclose(th, th_ref, rtol=1e-5, atol=1e-8)


def test_logreg_gd_on_gpu_matches_oracle():
    """logReg_GD.m:1-25 as one launch of the on-chip inner-loop kernel, caller-supplied terms."""
    rng = np.random.default_rng(5)
    s, d = 40, 23
    X = rng.standard_normal((s, d))
    Y = np.where(rng.standard_normal(s) > 0, 1.0, -1.0)
    x0 = 0.1 * rng.standard_normal(d)
    C1, C2, t1, t2 = (0.05 * rng.standard_normal(d) for _ in range(4))
    for step in (0.02, 1e-6):            # hits the 100-step cap / exits early on the tolerance
        xa, na = O.logReg_GD(X, Y, x0, C1, C2, t1, t2, 1e-3, step, return_steps=True)
        xb, nb = g.logReg_GD(X, Y, x0, C1, C2, t1, t2, 1e-3, step, backend="b200", return_steps=True)
        assert na == nb, (step, na, nb)
        np.testing.assert_allclose(xa, xb, rtol=1e-10, atol=1e-13)


def test_static_engine_rejects_foreign_chain_ends():
    import torch
    from gadmm_b200.engine.chain import ChainEngine
    rng = np.random.default_rng(0)
    N, s, d = 5, 10, 6
    eng = ChainEngine(N, d, 1.0, 0.0, 1e-4, 10, dynamic=False)
    eng.ingest_host(rng.standard_normal((N * s, d)), rng.standard_normal(N * s), s)
    eng.set_path(np.arange(N))                                   # natural chain: fine
    with pytest.raises(ValueError, match="dynamic=True"):
        eng.set_path(np.array([2, 0, 1, 3, 4]))                  # worker 2 would be a chain end
    torch.cuda.synchronize()
