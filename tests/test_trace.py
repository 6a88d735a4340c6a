"""Chrome-trace export of the per-iteration clocks."""
import json

import numpy as np
import pytest

from gadmm_b200.utils.trace import chrome_trace, iteration_stats


def test_trace_of_an_oracle_run(tmp_path):
    import gadmm_b200.api as G
    rng = np.random.default_rng(0)
    N, d, s = 4, 6, 6
    XX, YY = rng.standard_normal((N * s, d)), rng.standard_normal(N * s)
    obj0 = G.opt_sol_closedForm(XX, YY, backend="oracle")
    obj, loss, it, tv = G.group_ADMM_closedForm(XX, YY, 3.0, N, d, s, 50, obj0, 1e-4, backend="oracle")
    p = tmp_path / "t.json"
    tr = chrome_trace({"GADMM": tv}, str(p), loss={"GADMM": loss})
    ev = json.load(open(p))["traceEvents"]
    assert ev == tr["traceEvents"]
    slices = [e for e in ev if e["ph"] == "X"]
    assert 0 < len(slices) <= it
    assert abs(sum(e["dur"] for e in slices) - tv[-1] * 1e6) < 1e-3 * max(tv[-1] * 1e6, 1.0)
    assert sum(e["ph"] == "C" for e in ev) == it
    st = iteration_stats(tv)
    assert st["n"] == it - 1 and st["max"] >= st["median"] >= 0.0


def test_trace_rejects_a_non_cumulative_clock():
    with pytest.raises(ValueError):
        chrome_trace({"x": [0.0, 2.0, 1.0]})
    assert iteration_stats([0.0]) == {"n": 0}
