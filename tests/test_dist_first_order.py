"""torch.distributed first-order baselines (backend="gloo"): a world of one process on the CPU
must reproduce the oracle exactly; world_size=2 is covered by tests/test_gloo.py."""
import numpy as np

import gadmm_b200 as g
from gadmm_b200 import oracle as O


def _linear(N=7, s=12, d=9, seed=3):
    rng = np.random.default_rng(seed)
    XX = rng.standard_normal((N * s, d))
    YY = XX @ rng.standard_normal(d) + 0.1 * rng.standard_normal(N * s)
    Xl = [XX[i * s:(i + 1) * s] for i in range(N)]
    yl = [YY[i * s:(i + 1) * s] for i in range(N)]
    Hmax = np.array([np.linalg.eigvalsh(x.T @ x)[-1] for x in Xl])
    return XX, YY, Xl, yl, Hmax, N, s, d


def test_gloo_single_process_linear_matches_oracle():
    XX, YY, Xl, yl, Hmax, N, s, d = _linear()
    obj0 = O.opt_sol_closedForm(XX, YY)
    args = (Xl, yl, N, d, Hmax, 120, obj0, XX, YY, s)
    a = O.GD_DGD_LAG(*args, rng=5)
    b = g.GD_DGD_LAG(*args, rng=5, backend="gloo")
    assert len(a) == len(b) == 19
    for k in (0, 1, 3, 4, 6, 7, 10, 11, 14, 15, 16, 17):
        np.testing.assert_allclose(a[k], b[k], rtol=1e-9, atol=1e-12, err_msg=f"output {k}")
    np.testing.assert_array_equal(a[9], b[9])       # LAG-PS upload counter
    np.testing.assert_array_equal(a[13], b[13])     # LAG-WK upload counter
    assert abs(a[18] - b[18]) <= 1e-12 * abs(a[18])
    for mode in ("faithful", "jacobi"):
        da = O.dual_averaging(XX, YY, N, d, s, 80, obj0, 1e-6, a[18], mode=mode, return_state=True)
        db = g.dual_averaging(XX, YY, N, d, s, 80, obj0, 1e-6, a[18], mode=mode, backend="gloo",
                              return_state=True)
        assert da[2] == db[2]
        np.testing.assert_allclose(da[0], db[0], rtol=1e-10)
        np.testing.assert_allclose(da[4]["theta"], db[4]["theta"], rtol=1e-9, atol=1e-13)


def test_gloo_single_process_logistic_matches_oracle():
    rng = np.random.default_rng(8)
    N, s, d, lam = 5, 16, 6, 1e-2
    XX = rng.standard_normal((N * s, d))
    YY = np.where(rng.standard_normal(N * s) > 0, 1.0, -1.0)
    Xl = [XX[i * s:(i + 1) * s] for i in range(N)]
    yl = [YY[i * s:(i + 1) * s] for i in range(N)]
    Hmax = np.array([0.25 * np.linalg.eigvalsh(x.T @ x)[-1] + lam for x in Xl])
    args = (Xl, yl, N, d, Hmax, 100, XX, YY, s, lam, 1e-9)
    a = O.GD_DGD_LAG_logistic(*args, rng=2)
    b = g.GD_DGD_LAG_logistic(*args, rng=2, backend="gloo")
    assert len(a) == len(b) == 20
    for k in (0, 1, 3, 4, 6, 7, 10, 11, 14, 15, 16, 17):
        np.testing.assert_allclose(a[k], b[k], rtol=1e-9, atol=1e-12, err_msg=f"output {k}")
    assert abs(a[19] - b[19]) <= 1e-12 * abs(a[19])
    star = O.opt_sol_logistic(XX, YY, d, lam, N)
    da = O.dual_averaging_logisticReg(XX, YY, N, d, s, 60, star, 1e-9, a[18], lam)
    db = g.dual_averaging_logisticReg(XX, YY, N, d, s, 60, star, 1e-9, a[18], lam, backend="gloo")
    assert da[2] == db[2]
    np.testing.assert_allclose(da[0], db[0], rtol=1e-10)


def test_gloo_single_process_logistic_gadmm_matches_oracle():
    rng = np.random.default_rng(4)
    N, s, d, lam = 5, 14, 6, 1e-3
    XX = rng.standard_normal((N * s, d))
    YY = np.where(rng.standard_normal(N * s) > 0, 1.0, -1.0)
    star = O.opt_sol_logistic(XX, YY, d, lam, N)
    a = O.group_ADMM_logistic(XX, YY, 5.0, N, d, s, 40, star, lam, 1e-3, return_state=True)
    b = g.group_ADMM_logistic(XX, YY, 5.0, N, d, s, 40, star, lam, 1e-3, backend="gloo", return_state=True)
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-12)
    np.testing.assert_allclose(a[4]["theta"], b[4]["theta"], rtol=1e-10, atol=1e-14)
    a = O.group_ADMM_logistic_GD(XX, YY, 0.01, N, d, s, 30, star, lam, 0.0, 0.05, return_state=True)
    b = g.group_ADMM_logistic_GD(XX, YY, 0.01, N, d, s, 30, star, lam, 0.0, 0.05, backend="gloo",
                                 return_state=True)
    assert a[2] == b[2] and a[4]["inner_steps"] == b[4]["inner_steps"]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-12)
