"""Curve parity of the GPU chain engine against the CPU oracle."""
import numpy as np
import pytest

import gadmm_b200 as g
from gadmm_b200 import oracle as O

pytestmark = pytest.mark.gpu


def _cmp(a, b, rtol=1e-9):
    assert a[2] == b[2], (a[2], b[2])
    np.testing.assert_allclose(a[0], b[0], rtol=rtol, atol=1e-9)
    np.testing.assert_allclose(a[1], b[1], rtol=1e-5, atol=1e-8)


def test_gadmm_closed_form_matches_oracle_small(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    args = (p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 3000, obj0, 1e-4)
    a = O.group_ADMM_closedForm(*args)
    b = g.group_ADMM_closedForm(*args, backend="b200")
    _cmp(a, b)
    assert b[3].shape == (b[2],) and b[3][0] == 0.0 and np.all(np.diff(b[3]) >= 0) and b[3][-1] > 0


@pytest.mark.parametrize("rho", [3.0, 7.0])
def test_gadmm_reference_shape(ref_linear, rho):
    p = ref_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    args = (p["XX"], p["YY"], rho, p["N"], p["d"], p["s"], 4000, obj0, 1e-4)
    a = O.group_ADMM_closedForm(*args)
    b = g.group_ADMM_closedForm(*args, backend="b200")
    _cmp(a, b)
    b3 = g.group_ADMM(*args, backend="b200")
    assert len(b3) == 3 and b3[2] == a[2]


def test_gadmm_odd_and_tiny_chains():
    rng = np.random.default_rng(5)
    # (the last shape has 97 tile-rows: the finisher's four-group column reduction)
    for N, s, d in ((1, 9, 5), (2, 9, 5), (3, 9, 5), (7, 20, 66), (4, 30, 530), (2, 12, 6200)):
        XX = rng.standard_normal((N * s, d))
        YY = rng.standard_normal(N * s)
        obj0 = O.opt_sol_closedForm(XX, YY)
        args = (XX, YY, 1.5, N, d, s, 60, obj0, 1e-9)
        _cmp(O.group_ADMM_closedForm(*args), g.group_ADMM_closedForm(*args, backend="b200"))


def test_fp32_storage_tracks_oracle(small_linear):
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    args = (p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 3000, obj0, 1e-4)
    a = O.group_ADMM_closedForm(*args)
    b = g.group_ADMM_closedForm(*args, backend="b200", store="f32")
    assert abs(a[2] - b[2]) <= 2
    n = min(a[2], b[2])
    np.testing.assert_allclose(a[0][:n], b[0][:n], rtol=1e-5)


def test_static_and_dynamic_match_oracle(small_linear):
    p = small_linear
    N = p["N"]
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    rng = np.random.default_rng(0)
    paths, costs = [], []
    for _ in range(400):
        pa, co, _ = g.findPath(N, rng=rng)
        paths.append(pa)
        costs.append(co)
    paths, costs = np.array(paths), np.array(costs)
    base = (p["XX"], p["YY"], 2.0, N, p["d"], p["s"], 300, obj0, 1e-7)
    a = O.static_group_ADMM_closedForm(*base, 10, costs)
    b = g.static_group_ADMM_closedForm(*base, 10, costs, backend="b200")
    _cmp(a, b)
    np.testing.assert_allclose(a[4], b[4], rtol=1e-12)
    for tau in (1, 7):
        a = O.dynamic_group_ADMM_closedForm_v0(*base, paths, costs, tau)
        b = g.dynamic_group_ADMM_closedForm_v0(*base, paths, costs, tau, backend="b200")
        _cmp(a, b)
        np.testing.assert_allclose(a[4], b[4], rtol=1e-12)
    k = [0, 0]

    def mk(idx):
        def f(i):
            k[idx] += 1
            return paths[k[idx]], costs[k[idx]]
        return f

    a = O.dynamic_group_ADMM_closedForm(*base, paths[0], costs[0], 5, rechain=mk(0))
    b = g.dynamic_group_ADMM_closedForm(*base, paths[0], costs[0], 5, rechain=mk(1), backend="b200")
    _cmp(a, b)


def test_opt_sol_on_gpu(small_linear):
    p = small_linear
    assert g.opt_sol_closedForm(p["XX"], p["YY"], backend="b200") == pytest.approx(
        O.opt_sol_closedForm(p["XX"], p["YY"]), rel=1e-9, abs=1e-9)


def test_checkpoint_resume_matches_uninterrupted_run(small_linear):
    """state_out after 40 iterations -> state_in: the resumed curve continues bit-for-bit."""
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    args = (p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"])
    full = g.group_ADMM_closedForm(*args, 120, obj0, 0.0, backend="b200")
    part = g.group_ADMM_closedForm(*args, 40, obj0, 0.0, backend="b200", return_state=True)
    st = part[4]
    assert st["iteration"] == 40 and st["theta"].shape == (p["d"], p["N"])
    rest = g.group_ADMM_closedForm(*args, 120, obj0, 0.0, backend="b200", state=st)
    assert rest[2] == 120
    np.testing.assert_array_equal(rest[0], full[0])
    # the node duals of the oracle's node-dual run are interchangeable with the engine's
    o = O.static_group_ADMM_closedForm(*args, 40, obj0, 0.0, 10 ** 9, np.ones((1, p["N"] - 1)),
                                       return_state=True)
    np.testing.assert_allclose(o[5]["dual"], st["dual"], rtol=1e-9, atol=1e-12)


def test_dgadmm_with_device_generated_chains(small_linear):
    """Chains drawn and built on the GPU: replaying the very same chains through the oracle gives
    the same curve, and re-chaining converges faster than the static chain."""
    import torch
    from gadmm_b200.ops.topology import find_path_batch_device
    p = small_linear
    N = p["N"]
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    base = (p["XX"], p["YY"], 2.0, N, p["d"], p["s"], 300, obj0, 1e-7)
    path0 = np.arange(1, N + 1)
    b = g.dynamic_group_ADMM_closedForm(*base, path0, np.ones(N - 1), 4, backend="b200",
                                        chain_gen="device", rng=77)
    paths, sums = find_path_batch_device(256, N, seed=77, stream_id=1, model=1,
                                         device=torch.device("cuda", 0))
    paths = paths.cpu().numpy() + 1
    k = [-1]

    def replay(i):
        k[0] += 1
        return paths[k[0]], np.array([float(sums[k[0]])])

    a = O.dynamic_group_ADMM_closedForm(*base, path0, np.ones(N - 1), 4, rechain=replay)
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(a[4], b[4], rtol=1e-12)


def test_simulated_clock_mode(small_linear):
    """clock="simulated": the reference's 2 x (one worker's solve) model, iteration 1 untimed."""
    p = small_linear
    obj0 = O.opt_sol_closedForm(p["XX"], p["YY"])
    args = (p["XX"], p["YY"], 2.0, p["N"], p["d"], p["s"], 3000, obj0, 1e-4)
    a = g.group_ADMM_closedForm(*args, backend="b200")
    b = g.group_ADMM_closedForm(*args, backend="b200", clock="simulated")
    assert a[2] == b[2]
    np.testing.assert_array_equal(a[0], b[0])
    t = b[3]
    assert t.shape == (b[2],) and t[0] == 0.0
    step = np.diff(t)
    assert step[0] > 0 and np.allclose(step, step[0], rtol=1e-12)
    assert 1e-6 < step[0] < 1e-2            # two SYMV launches of a 64x64 problem: microseconds


@pytest.mark.parametrize("d", [12, 150])
def test_state_after_early_stop_is_the_state_of_iter(d):
    """The stop test fires inside a launch window (persistent kernel: between two grid barriers;
    streaming path: with the next head phase already launched).  The returned state must be the
    one of iteration ``Iter`` exactly, as if the run had broken there like the reference does
    (group_ADMM_closedForm.m:105-108): it equals the oracle's state and resumes bit-for-bit."""
    rng = np.random.default_rng(11)
    N, s = 7, 20
    XX = rng.standard_normal((N * s, d))
    YY = rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    args = (XX, YY, 2.0, N, d, s)
    acc = 1e-3
    o = O.static_group_ADMM_closedForm(*args, 3000, obj0, acc, 10 ** 9, np.ones((1, N - 1)),
                                       return_state=True)
    b = g.group_ADMM_closedForm(*args, 3000, obj0, acc, backend="b200", return_state=True)
    assert b[2] == o[2] and b[2] % 8 != 0 and b[2] % 32 != 0, b[2]
    st = b[4]
    assert st["iteration"] == b[2]
    np.testing.assert_allclose(st["theta"], o[5]["theta"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(st["dual"], o[5]["dual"], rtol=1e-9, atol=1e-12)
    # continuing from that state reproduces the uninterrupted run
    more = b[2] + 37
    full = g.group_ADMM_closedForm(*args, more, obj0, 0.0, backend="b200")
    rest = g.group_ADMM_closedForm(*args, more, obj0, 0.0, backend="b200", state=st)
    np.testing.assert_allclose(rest[0][b[2]:], full[0][b[2]:], rtol=1e-12)


def test_mixed_storage_switches_to_fp64_and_stops_exactly():
    """store="mixed": fp32-stored inverses while the gap is large (half the bytes per iteration,
    objective only estimated), fp64 inverses -- exact objective identity and stop test -- once the
    estimated gap is below the switch tolerance.  The run must stop where the true gap is below
    acc, within a couple of iterations of the all-fp64 run."""
    import warnings
    rng = np.random.default_rng(3)
    N, s, d = 6, 40, 150
    XX = rng.standard_normal((N * s, d))
    YY = XX @ rng.standard_normal(d) + 0.1 * rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    args = (XX, YY, 2.0, N, d, s, 3000, obj0, 1e-6)
    a = O.group_ADMM_closedForm(*args)
    b = g.group_ADMM_closedForm(*args, backend="b200", store="mixed", return_engine=True)
    eng = b[-1]
    assert 0 < eng.switched_at < b[2]
    assert abs(a[2] - b[2]) <= 3, (a[2], b[2])
    assert b[1][-1] < 1e-6
    # the final objective is exact (fp64 identity), not an fp32 estimate
    th = eng.state()["theta"]
    true = sum(0.5 * np.sum((XX[n * s:(n + 1) * s] @ th[:, n] - YY[n * s:(n + 1) * s]) ** 2) for n in range(N))
    assert abs(true - b[0][-1]) <= 1e-9 * max(1.0, abs(true))
    with warnings.catch_warnings(record=True) as w:          # plain fp32 storage + stop test: warned
        warnings.simplefilter("always")
        g.group_ADMM_closedForm(*args[:6], 20, obj0, 1e-6, backend="b200", store="f32")
        assert any("first-order" in str(x.message) for x in w)
