import os

import numpy as np
import pytest
import scipy.io as sio

import gadmm_b200 as g
from gadmm_b200 import io as gio
from gadmm_b200.results import (cumulative_com, energy_cumulative, iterations_to, star_cost,
                                with_rechain_overhead)

REF = "/root/reference"


def test_input_data_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    X = rng.standard_normal((60, 5))
    y = np.where(rng.random(60) < 0.3, 1.0, -1.0)
    p = str(tmp_path / "inputData.mat")
    gio.save_input_data(p, X, y)
    raw = sio.loadmat(p)
    assert raw["X_fede"].shape == (60, 5) and raw["X_fede"].dtype == np.float64
    assert raw["y_fede"].shape == (60, 1) and raw["y_fede"].dtype == np.int16   # MATLAB compaction
    X2, y2 = gio.load_input_data(p)
    assert y2.dtype == np.float64
    np.testing.assert_array_equal(X, X2)
    np.testing.assert_array_equal(y, y2)


def test_optimal_sol_roundtrip(tmp_path):
    p = str(tmp_path / "optimalSol.mat")
    gio.save_optimal_sol(p, 0.7248784914, 400)
    raw = sio.loadmat(p)
    assert raw["obj0"].shape == (1, 1) and raw["opt_obj"].shape == (1, 400)
    obj0, opt = gio.load_optimal_sol(p)
    assert obj0 == 0.7248784914 and np.all(opt == obj0)


@pytest.mark.skipif(not os.path.exists(f"{REF}/inputData.mat"), reason="reference not mounted")
def test_reference_fixtures_load():
    X, y = gio.load_input_data(f"{REF}/inputData.mat")
    assert X.shape == (1200, 50) and y.shape == (1200,) and set(np.unique(y)) == {-1.0, 1.0}
    obj0, opt = gio.load_optimal_sol(f"{REF}/optimalSol.mat")
    assert opt.shape == (40000,) and obj0 == pytest.approx(0.7248784914)


def test_curve_writers(tmp_path):
    c = {"loss_GADMM": np.array([1.0, 0.1, 0.01]), "Iter": np.array([3.0])}
    for ext in ("mat", "csv", "json"):
        gio.save_curves(str(tmp_path / f"c.{ext}"), c)
        assert os.path.getsize(tmp_path / f"c.{ext}") > 0
    with pytest.raises(ValueError):
        gio.save_curves(str(tmp_path / "c.png"), c)


def test_text_dataset(tmp_path):
    d = tmp_path / "data29"
    d.mkdir()
    np.savetxt(d / "data.txt", np.arange(12.0).reshape(4, 3))
    np.savetxt(d / "y.txt", np.arange(4.0))
    X, y = gio.load_text_dataset(str(d))
    assert X.shape == (4, 3) and y.shape == (4,)


def test_find_path_contract():
    for fn, area in ((g.findPath, 50.0), (g.findPath2, 250.0)):
        out = fn(17, rng=3)
        path, cost, d2 = out[0], out[1], out[2]
        assert path[0] == 1 and sorted(path) == list(range(1, 18))
        assert cost.shape == (16,) and d2.shape == (17, 17)
        assert np.allclose(d2, d2.T) and d2.max() <= 2 * area ** 2
    path, cost, d2, P_central, center = g.findPath2(17, rng=3)
    scale = 1e-6 * 2e6 * 2.0 ** 5
    np.testing.assert_allclose(cost, d2[path[:-1] - 1, path[1:] - 1] * scale)
    assert P_central[center - 1] == 0.0 and 1 <= center <= 17
    # greedy: every hop goes to the nearest unvisited node
    seen = {path[0]}
    for a, b in zip(path[:-1], path[1:]):
        rest = [m for m in range(1, 18) if m not in seen]
        assert d2[a - 1, b - 1] == min(d2[a - 1, m - 1] for m in rest)
        seen.add(b)


def test_calc_cost():
    grid = np.arange(16.0).reshape(4, 4)
    np.testing.assert_array_equal(g.calc_cost(grid, [1, 3, 2, 4]), [grid[0, 2], grid[2, 1], grid[1, 3]])


def test_comm_accounting():
    np.testing.assert_array_equal(cumulative_com("GD", 3, 24), [25, 50, 75])
    np.testing.assert_array_equal(cumulative_com("GADMM", 2, 24), [24, 48])
    np.testing.assert_array_equal(cumulative_com("LAG_PS", 3, 24, [1, 5, 9]), [2, 7, 12])
    assert star_cost([1.0, 2.0, 3.0]) == 9.0
    np.testing.assert_allclose(with_rechain_overhead([3.0]), [4.0])
    np.testing.assert_array_equal(energy_cumulative(2.0, 3), [2, 4, 6])
    assert iterations_to([1, 0.5, 1e-5, 1e-6], 1e-4) == 3


def test_checkpoint_files_resume_exactly(tmp_path):
    """save_state / load_state (.mat, .npz, .pt): a run resumed from the file continues the
    uninterrupted run bit for bit."""
    import numpy as np
    from gadmm_b200 import io, oracle as O
    rng = np.random.default_rng(2)
    N, s, d = 6, 10, 8
    XX, YY = rng.standard_normal((N * s, d)), rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    path0 = np.arange(1, N + 1)
    args = (XX, YY, 2.0, N, d, s)
    full = O.dynamic_group_ADMM_closedForm(*args, 60, obj0, 0.0, path0, np.ones(N - 1), 1e9)
    part = O.dynamic_group_ADMM_closedForm(*args, 25, obj0, 0.0, path0, np.ones(N - 1), 1e9,
                                           return_state=True)
    for ext in (".mat", ".npz", ".pt"):
        f = str(tmp_path / ("ckpt" + ext))
        io.save_state(f, part[-1])
        st = io.load_state(f)
        assert st["iteration"] == 25 and st["theta"].shape == (d, N) and st["path"].tolist() == path0.tolist()
        rest = O.dynamic_group_ADMM_closedForm(*args, 35, obj0, 0.0, path0, np.ones(N - 1), 1e9, state=st)
        np.testing.assert_array_equal(np.concatenate([part[0], rest[0]]), full[0])
