"""CPU/gloo world_size=2: same curves as the single-process oracle."""
import os
import sys

import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def test_gloo_world2_matches_oracle():
    from mp_gloo_check import run
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=run, args=(r, 2, 29731, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=280) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
    assert [r[1] for r in res] == ["ok", "ok"] and all(p.exitcode == 0 for p in procs)
