"""Cost of one iteration of a CUDA-graph WHILE node whose body is a single 1-thread kernel
(gadmm_graph_loops_selftest): wall time of n iterations minus wall time of 1, per iteration."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gadmm_b200.ops import native as nv

L = nv.lib()
torch.cuda.init()
st = nv.stream_ptr()
L.gadmm_graph_loops_selftest(3, st)
res = {}
for n in (1, 20001):
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = L.gadmm_graph_loops_selftest(n, st)
        best = min(best, time.perf_counter() - t0)
        assert rc == 0
    res[n] = best
print({"us_per_while_iteration": round((res[20001] - res[1]) / 20000 * 1e6, 3), "raw_s": res})
