"""Star ADMM and logistic GADMM on the GPU vs the CPU oracle."""
import numpy as np
import pytest

import gadmm_b200 as g
from gadmm_b200 import oracle as O

pytestmark = pytest.mark.gpu


def test_star_admm_matches_oracle(small_linear, ref_linear):
    for p, rho, cap in ((small_linear, 1.0, 3000), (ref_linear, 1.0, 3000)):
        obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
        args = (p["XX"], p["YY"], rho, p["N"], p["d"], p["s"], cap, obj0, 1e-4)
        a = O.standared_ADMM(*args)
        b = g.standared_ADMM(*args, backend="b200")
        assert a[2] == b[2] and a[2] < cap
        np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-9)
        assert b[3].shape == (b[2],) and b[3][0] == 0.0


def test_star_admm_two_workers():
    rng = np.random.default_rng(1)
    N, s, d = 2, 12, 70
    XX, YY = rng.standard_normal((N * s, d)), rng.standard_normal(N * s)
    obj0 = O.opt_sol(XX, YY, d)
    args = (XX, YY, 0.7, N, d, s, 80, obj0, 1e-12)
    a, b = O.standared_ADMM(*args), g.standared_ADMM(*args, backend="b200")
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-9)


def test_logistic_gd_matches_oracle(small_logistic):
    p = small_logistic
    star = O.opt_sol_logistic(p["XX"], p["YY"], p["d"], p["lam"], p["N"])
    args = (p["XX"], p["YY"], 0.01, p["N"], p["d"], p["s"], 300, star, p["lam"], 1e-4, 0.5)
    a = O.group_ADMM_logistic_GD(*args, return_state=True)
    b = g.group_ADMM_logistic_GD(*args, backend="b200", return_state=True)
    assert a[2] == b[2] and a[2] < 300
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)
    assert a[4]["inner_steps"] == b[4]["inner_steps"]


def test_logistic_gd_reference_shape():
    """N=24, s=d=50, lam=1e-5, gdStep=2.2 (LogisticRegression_Synthetic.m:40-41,88-99)."""
    from gadmm_b200.data import logreg_synthetic
    X, y, Hmax, XX, YY = logreg_synthetic(N=24, s=50, lam=1e-5, rng=np.random.default_rng(9))
    star = O.opt_sol_logistic(XX, YY, 50, 1e-5, 24)
    args = (XX, YY, 3e-4, 24, 50, 50, 400, star, 1e-5, 1e-4, 2.2)
    a = O.group_ADMM_logistic_GD(*args)
    b = g.group_ADMM_logistic_GD(*args, backend="b200")
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-8)


def test_logistic_newton_matches_oracle(small_logistic):
    p = small_logistic
    star = O.opt_sol_logistic(p["XX"], p["YY"], p["d"], p["lam"], p["N"])
    args = (p["XX"], p["YY"], 0.01, p["N"], p["d"], p["s"], 100, star, p["lam"], 1e-4)
    a = O.group_ADMM_logistic(*args)
    b = g.group_ADMM_logistic(*args, backend="b200")
    assert a[2] == b[2] and a[2] < 100
    np.testing.assert_allclose(a[0], b[0], rtol=1e-7)


def test_logistic_newton_tensor_core_hessian():
    """d=160 -> the Hessian X'DX goes through the tcgen05 Gram; iterates still match the oracle."""
    rng = np.random.default_rng(12)
    N, s, d, lam = 4, 200, 160, 1e-2
    XX = rng.standard_normal((N * s, d)) / np.sqrt(d)
    YY = np.where(XX @ rng.standard_normal(d) + 0.3 * rng.standard_normal(N * s) > 0, 1.0, -1.0)
    star = O.opt_sol_logistic(XX, YY, d, lam, N)
    args = (XX, YY, 0.5, N, d, s, 60, star, lam, 1e-5)
    a = O.group_ADMM_logistic(*args)
    b = g.group_ADMM_logistic(*args, backend="b200")
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-7)
