"""GPU first-order baselines vs the CPU oracle (GD, DGD, LAG-PS, LAG-WK, IAG, dual averaging)."""
import numpy as np
import pytest

import gadmm_b200 as g
from gadmm_b200 import oracle as O

pytestmark = pytest.mark.gpu

NAMES = ["obj_GD", "loss_GD", "time_GD", "obj_DGD", "loss_DGD", "time_DGD", "obj_LAG_PS",
         "loss_LAG_PS", "time_lagPS", "comm_PS", "obj_LAG_WK", "loss_LAG_WK", "time_lagWK",
         "comm_WK", "obj_cIAG", "loss_cIAG", "obj_RIAG", "loss_RIAG", "stepsize"]


def _compare(a, b, n_out):
    assert len(a) == len(b) == n_out
    assert a[18] == pytest.approx(b[18], rel=1e-10)                   # step size (lambda_max)
    for k in (0, 3, 6, 10, 14, 16):                                   # objective curves
        assert a[k].shape == b[k].shape, NAMES[k]
        np.testing.assert_allclose(a[k], b[k], rtol=1e-7, atol=1e-9, err_msg=NAMES[k])
    for k in (9, 13):                                                 # upload counters
        # once the iterate has converged to rounding level the trigger compares noise with noise
        # (both sides ~1e-30), so the counters are only comparable before that point
        live = np.nonzero(a[k - 2] < 1e-7)[0]
        n = int(live[0]) if live.size else len(a[k])
        assert n > 20
        np.testing.assert_array_equal(a[k][:n], b[k][:n], err_msg=NAMES[k])
    for k in (2, 5, 8, 12):                                           # clocks: shape + monotone
        assert b[k].shape == a[k].shape and np.all(np.diff(b[k]) >= 0)


def test_linear_baselines_match_oracle(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    args = (p["X"], p["y"], p["N"], p["d"], p["Hmax"], 1500, obj0, p["XX"], p["YY"], p["s"])
    a = O.GD_DGD_LAG(*args, rng=4)
    b = g.GD_DGD_LAG(*args, rng=4, backend="b200")
    _compare(a, b, 19)
    assert a[9][-1] < 1500 * p["N"]           # LAG really skipped uploads


def test_logistic_baselines_match_oracle(small_logistic):
    p = small_logistic
    args = (p["X"], p["y"], p["N"], p["d"], p["Hmax"], 1200, p["XX"], p["YY"], p["s"], p["lam"], 1e-4)
    a = O.GD_DGD_LAG_logistic(*args, rng=4)
    b = g.GD_DGD_LAG_logistic(*args, rng=4, backend="b200")
    _compare(a, b, 20)
    assert a[19] == pytest.approx(b[19], rel=1e-9)


@pytest.mark.parametrize("mode", ["faithful", "jacobi"])
def test_dual_averaging_matches_oracle(small_linear, small_logistic, mode):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    alpha = 1.0 / np.linalg.eigvalsh(p["XX"].T @ p["XX"])[-1]
    args = (p["XX"], p["YY"], p["N"], p["d"], p["s"], 400, obj0, 1e-4, alpha)
    a = O.dual_averaging(*args, mode=mode)
    b = g.dual_averaging(*args, mode=mode, backend="b200")
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)
    q = small_logistic
    star = O.opt_sol_logistic(q["XX"], q["YY"], q["d"], q["lam"], q["N"])
    al = 1.0 / (0.25 * np.linalg.eigvalsh(q["XX"].T @ q["XX"])[-1] + q["lam"])
    args = (q["XX"], q["YY"], q["N"], q["d"], q["s"], 300, star, 1e-4, al, q["lam"])
    a = O.dual_averaging_logisticReg(*args, mode=mode)
    b = g.dual_averaging_logisticReg(*args, mode=mode, backend="b200")
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)


def test_dual_averaging_stops_early():
    """A generous tolerance makes the stop test fire mid-window; Iter and lengths must agree."""
    rng = np.random.default_rng(2)
    N, s, d = 5, 30, 8
    XX = rng.standard_normal((N * s, d))
    YY = XX @ rng.standard_normal(d)
    obj0 = O.opt_sol_closedForm(XX, YY)
    alpha = 0.5 / np.linalg.eigvalsh(XX.T @ XX)[-1]
    args = (XX, YY, N, d, s, 3000, obj0, 1.0, alpha)
    a = O.dual_averaging(*args)
    b = g.dual_averaging(*args, backend="b200")
    assert a[2] == b[2] and b[0].shape == (b[2],)
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)


def test_checkpoint_resume_star_and_first_order():
    """state after a partial run -> resumed run continues the uninterrupted curve bit for bit
    (star ADMM, LAG-WK, DGD, dual averaging): SURVEY.md section 5.4."""
    from gadmm_b200.engine.first_order import FirstOrder
    from gadmm_b200.engine.star import star_admm
    from gadmm_b200.ops import native as nv
    rng = np.random.default_rng(21)
    N, s, d = 7, 20, 14
    XX = rng.standard_normal((N * s, d))
    YY = XX @ rng.standard_normal(d) + 0.1 * rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    full = star_admm(XX, YY, 1.0, N, d, s, 90, obj0, 0.0)
    part = star_admm(XX, YY, 1.0, N, d, s, 35, obj0, 0.0, return_state=True)
    st = part[4]
    assert st["iteration"] == 35 and st["theta"].shape == (d, N - 1)
    rest = star_admm(XX, YY, 1.0, N, d, s, 90, obj0, 0.0, state=st)
    assert rest[2] == 90
    np.testing.assert_array_equal(rest[0], full[0])
    o = O.standared_ADMM(XX, YY, 1.0, N, d, s, 35, obj0, 0.0)
    np.testing.assert_allclose(o[0], part[0], rtol=1e-9)

    fo = FirstOrder(XX, YY, N, d, s)
    alpha = 1.0 / fo.lambda_max()
    Hmax = fo.hmax_per_worker()
    thr = 1.0 / (alpha ** 2 * N ** 2) / 10
    a = fo.run_shared(nv.FO_LAG_WK, alpha, thr, Hmax, 120, obj0, None)
    b = fo.run_shared(nv.FO_LAG_WK, alpha, thr, Hmax, 50, obj0, None, return_state=True)
    c = fo.run_shared(nv.FO_LAG_WK, alpha, thr, Hmax, 120, obj0, None, state=b[5])
    np.testing.assert_array_equal(a[0], c[0])
    np.testing.assert_array_equal(a[3], c[3])                   # upload counters continue
    a = fo.run_dgd(alpha / 100, 80, obj0)
    b = fo.run_dgd(alpha / 100, 30, obj0, return_state=True)
    c = fo.run_dgd(alpha / 100, 80, obj0, state=b[3])
    np.testing.assert_array_equal(a[0], c[0])
    a = fo.run_dualavg(alpha, 80, obj0, 0.0)
    b = fo.run_dualavg(alpha, 30, obj0, 0.0, return_state=True)
    c = fo.run_dualavg(alpha, 80, obj0, 0.0, state=b[5])
    np.testing.assert_array_equal(a[0], c[0])
