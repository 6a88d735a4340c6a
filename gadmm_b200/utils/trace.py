"""Per-iteration traces.

Every solver returns a cumulative clock vector (``gadmm_time``, ``time_GD``, ... -- on the GPU
backend the device time since iteration 1, max over ranks).  ``chrome_trace`` turns a set of such
vectors into a Chrome / Perfetto trace (one track per algorithm, one slice per iteration, the
objective gap as a counter), so a run can be inspected on a timeline without a profiler attached.
The reference only keeps the vectors (group_ADMM_closedForm.m:40-42,53-55).
"""
from __future__ import annotations

import json
from typing import Mapping, Optional, Sequence

import numpy as np


def chrome_trace(times: Mapping[str, Sequence[float]], path: Optional[str] = None,
                 loss: Optional[Mapping[str, Sequence[float]]] = None, unit_s: float = 1.0) -> dict:
    """``times[name][i]`` = cumulative seconds (x ``unit_s``) at the end of iteration ``i+1``
    (``times[name][0]`` may be 0: the reference's ``time(1)=0`` convention).  Returns the trace
    object and writes it to ``path`` when given."""
    events = []
    for tid, (name, tv) in enumerate(times.items(), start=1):
        tv = np.asarray(tv, dtype=np.float64).ravel() * unit_s * 1e6      # microseconds
        if tv.size and np.any(np.diff(tv) < 0):
            raise ValueError(f"clock vector of {name!r} is not cumulative")
        events.append({"ph": "M", "pid": 1, "tid": tid, "name": "thread_name", "args": {"name": name}})
        prev = 0.0
        for i, t in enumerate(tv):
            if t > prev:
                events.append({"ph": "X", "pid": 1, "tid": tid, "name": f"iter {i + 1}",
                               "ts": prev, "dur": t - prev})
            prev = max(prev, float(t))
        if loss is not None and name in loss:
            lv = np.asarray(loss[name], dtype=np.float64).ravel()
            for i in range(min(lv.size, tv.size)):
                events.append({"ph": "C", "pid": 1, "tid": tid, "name": f"gap {name}",
                               "ts": float(tv[i]), "args": {"gap": float(lv[i])}})
    trace = {"traceEvents": events, "displayTimeUnit": "ms"}
    if path:
        with open(path, "w") as f:
            json.dump(trace, f)
    return trace


def iteration_stats(tv: Sequence[float]) -> dict:
    """Median / p99 / max per-iteration time of a cumulative clock vector."""
    dt = np.diff(np.asarray(tv, dtype=np.float64).ravel())
    if dt.size == 0:
        return {"n": 0}
    return {"n": int(dt.size), "median": float(np.median(dt)), "p99": float(np.quantile(dt, 0.99)),
            "max": float(dt.max()), "total": float(dt.sum())}
