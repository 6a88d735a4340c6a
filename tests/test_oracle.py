"""Oracle invariants derived from the reference's own redundancy."""
import os

import numpy as np
import pytest

import gadmm_b200 as g
from gadmm_b200 import oracle as O
from gadmm_b200.data import hmax_logistic
from gadmm_b200.topology import chain_tables

REF = "/root/reference"


def test_edge_dual_equals_node_dual(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    a = O.group_ADMM_closedForm(p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 400, obj0, 1e-6)
    costs = np.ones((1000, p["N"] - 1))
    b = O.static_group_ADMM_closedForm(p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 400, obj0,
                                       1e-6, 10, costs)
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-11, atol=1e-11)
    assert b[4].shape == (b[2],)
    np.testing.assert_allclose(b[4], (p["N"] - 1) * np.arange(1, b[2] + 1))


def test_dynamic_with_identity_path_equals_static(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    path = np.arange(1, p["N"] + 1)
    a = O.static_group_ADMM_closedForm(p["XX"], p["YY"], 3.0, p["N"], p["d"], p["s"], 300, obj0,
                                       1e-6, 10 ** 9, np.ones((1, p["N"] - 1)))
    b = O.dynamic_group_ADMM_closedForm(p["XX"], p["YY"], 3.0, p["N"], p["d"], p["s"], 300, obj0,
                                        1e-6, path, np.ones(p["N"] - 1), 1e9)
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-12)


def test_cvx_semantics_equal_closed_form(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    a = O.group_ADMM(p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 200, obj0, 1e-5)
    b = O.group_ADMM_closedForm(p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 200, obj0, 1e-5)
    assert len(a) == 3 and len(b) == 4
    np.testing.assert_allclose(a[0], b[0], rtol=1e-8)


def test_return_contract(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    obj, loss, it, t = O.group_ADMM_closedForm(p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 5000,
                                               obj0, 1e-4)
    assert it < 5000 and obj.shape == loss.shape == t.shape == (it,)
    assert loss[-1] < 1e-4 and np.all(loss[:-1] >= 1e-4)
    assert t[0] == 0.0 and np.all(np.diff(t) >= 0)
    # cap reached -> Iter = num_iter
    obj, loss, it, t = O.group_ADMM_closedForm(p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 7,
                                               obj0, 1e-12)
    assert it == 7 and obj.shape == (7,)


def test_star_admm_converges_to_optimum(small_linear):
    p = small_linear
    obj0, th = O.opt_sol_closedForm(p["XX"], p["YY"], return_theta=True)
    *_, st = O.standared_ADMM(p["XX"], p["YY"], 1.0, p["N"], p["d"], p["s"], 20000, obj0, 1e-10,
                              return_state=True)
    np.testing.assert_allclose(st["theta"], np.tile(th[:, None], (1, p["N"])), atol=1e-4)


def test_opt_sol_variants_agree(small_linear):
    p = small_linear
    assert abs(O.opt_sol(p["XX"], p["YY"], p["d"]) - O.opt_sol_closedForm(p["XX"], p["YY"])) < 1e-9


def test_dgadmm_rechaining_reduces_iterations(ref_linear):
    p = ref_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    path, cost, _ = g.findPath(p["N"], rng=1)
    its = []
    for tau in (1e9, 10):
        r = O.dynamic_group_ADMM_closedForm(p["XX"], p["YY"], 1.0, p["N"], p["d"], p["s"], 6000,
                                            obj0, 1e-4, path, cost, tau, rng=5)
        its.append(r[2])
    assert its[1] < its[0]


def test_v0_uses_pregenerated_chains_and_column_quirk(small_linear):
    p = small_linear
    N = p["N"]
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    rng = np.random.default_rng(0)
    paths, costs = [], []
    for _ in range(60):
        pa, co, _ = g.findPath(N, rng=rng)
        paths.append(pa)
        costs.append(co)
    paths, costs = np.array(paths), np.array(costs)
    a = O.dynamic_group_ADMM_closedForm_v0(p["XX"], p["YY"], 2.0, N, p["d"], p["s"], 50, obj0,
                                           1e-9, paths, costs, 5, faithful=True)
    b = O.dynamic_group_ADMM_closedForm_v0(p["XX"], p["YY"], 2.0, N, p["d"], p["s"], 50, obj0,
                                           1e-9, paths, costs, 5, faithful=False)
    np.testing.assert_allclose(a[0], b[0])
    assert a[4][0] == pytest.approx(costs[:, 0].sum())      # quirk 4: matrix column
    assert b[4][0] == pytest.approx(costs[0].sum())
    # replay through the live variant with an injected re-chain hook gives the same curve
    k = [0]

    def rechain(i):
        k[0] += 1
        return paths[k[0]], costs[k[0]]

    c = O.dynamic_group_ADMM_closedForm(p["XX"], p["YY"], 2.0, N, p["d"], p["s"], 50, obj0, 1e-9,
                                        paths[0], costs[0], 5, rechain=rechain)
    np.testing.assert_allclose(a[0], c[0], rtol=1e-12)


def test_chain_tables():
    left, right, color, cmul = chain_tables(np.array([0, 3, 1, 2]))
    assert list(left) == [-1, 3, 1, 0] and list(right) == [3, 2, -1, 1]
    assert list(color) == [0, 0, 1, 1] and list(cmul) == [1, 2, 1, 2]


def test_baselines_shapes_and_quirks(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    T = 1200
    out = O.GD_DGD_LAG(p["X"], p["y"], p["N"], p["d"], p["Hmax"], T, obj0, p["XX"], p["YY"], p["s"],
                       rng=0)
    assert len(out) == 19
    for k in (0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17):
        assert out[k].shape == (T,), k
    # GD at iteration 1 steps along the ones vector (quirk 3)
    alpha = out[18]
    XX, YY = p["XX"], p["YY"]
    th1 = -alpha * np.ones(p["d"])
    assert out[0][1] == pytest.approx(0.5 * np.sum((XX @ th1 - YY) ** 2))
    # GD converges, LAG uploads fewer gradients than GD
    assert out[1][-1] < out[1][0]
    assert out[9][-1] < T * p["N"] and out[13][-1] < T * p["N"]
    assert np.all(np.diff(out[9]) >= 0)
    # dual averaging runs with GD's step size
    r = O.dual_averaging(XX, YY, p["N"], p["d"], p["s"], 50, obj0, 1e-4, alpha)
    assert r[0].shape == (r[2],)


def test_logistic_family(small_logistic):
    p = small_logistic
    out = O.GD_DGD_LAG_logistic(p["X"], p["y"], p["N"], p["d"], p["Hmax"], 3000, p["XX"], p["YY"],
                                p["s"], p["lam"], 1e-4, rng=0)
    assert len(out) == 20
    obj1 = out[19]
    assert out[1][-1] == 0.0                       # loss_GD(end) = |obj_GD(end) - obj1| = 0
    star = O.opt_sol_logistic(p["XX"], p["YY"], p["d"], p["lam"], p["N"])
    assert obj1 >= star - 1e-9
    r = O.group_ADMM_logistic_GD(p["XX"], p["YY"], 0.01, p["N"], p["d"], p["s"], 300, star,
                                 p["lam"], 1e-4, 0.5)
    rn = O.group_ADMM_logistic(p["XX"], p["YY"], 0.05, p["N"], p["d"], p["s"], 300, star,
                               p["lam"], 1e-4)
    assert r[2] < 300 and rn[2] < 300
    d = O.dual_averaging_logisticReg(p["XX"], p["YY"], p["N"], p["d"], p["s"], 20, star, 1e-4,
                                     out[18], p["lam"])
    assert d[0].shape == (d[2],)


def test_group_admm_logistic_default_iter_quirk(small_logistic):
    p = small_logistic
    r = O.group_ADMM_logistic(p["XX"], p["YY"], 0.05, p["N"], p["d"], p["s"], 3, 0.0, p["lam"],
                              1e-12, faithful=True)
    assert r[2] == 100 and r[0].shape == (3,)     # quirk 2: Iter=100 default


@pytest.mark.skipif(not os.path.exists(f"{REF}/inputData.mat"), reason="reference not mounted")
def test_survey_replay_on_input_data():
    """GADMM-GD logistic on inputData.mat, rho=2e-4 -> 53 iterations."""
    X, y = g.io.load_input_data(f"{REF}/inputData.mat")
    N, s, d, lam = 24, 50, 50, 1e-5
    obj0 = O.opt_sol_logistic(X, y, d, lam, N)
    assert obj0 == pytest.approx(0.7177269845, abs=2e-9)
    r = O.group_ADMM_logistic_GD(X, y, 2e-4, N, d, s, 400, obj0, lam, 1e-4, 2.2)
    assert r[2] == 53
