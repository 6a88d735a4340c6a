"""``backend="nccl"`` / ``backend="gloo"`` implementations of the public API (library ops +
torch.distributed collectives; see :mod:`gadmm_b200.parallel.nccl_baseline`)."""
from __future__ import annotations

import numpy as np

from ..topology import findPath2
from .nccl_baseline import NcclChainGadmm


class _Backend:
    def __init__(self, name):
        self.name = name

    def _eng(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, dynamic):
        e = NcclChainGadmm(N, d, rho, obj0, acc, num_iter, backend=self.name, dynamic=dynamic)
        e.ingest_host(XX, YY, s)
        return e

    def gadmm_closed_form(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, **kw):
        e = self._eng(XX, YY, rho, N, d, s, num_iter, obj0, acc, False)
        e.set_path(np.arange(N))
        return e.results(e.run(num_iter))

    def static_gadmm(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, tau, costs, **kw):
        costs = np.atleast_2d(np.asarray(costs, dtype=np.float64))
        obj, loss, it, tv = self.gadmm_closed_form(XX, YY, rho, N, d, s, num_iter, obj0, acc)
        per = np.array([costs[min(int(i // tau) - (1 if tau == 1 else 0), len(costs) - 1)].sum()
                        for i in range(1, it + 1)])
        return obj, loss, it, tv, np.cumsum(per)

    def _dynamic(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, p0, c0, tau, nxt):
        e = self._eng(XX, YY, rho, N, d, s, num_iter, obj0, acc, True)
        e.set_path(np.asarray(p0, dtype=np.int64) - 1)
        cost_at = {1: float(c0)}

        def rechain(i):
            p1, pc = nxt(i)
            cost_at[i] = float(np.sum(pc))
            return np.asarray(p1, dtype=np.int64) - 1

        it = e.run(num_iter, rechain=rechain, coherence=tau)
        obj, loss, it, tv = e.results(it)
        per, c = np.empty(it), cost_at[1]
        for i in range(1, it + 1):
            c = cost_at.get(i, c)
            per[i - 1] = c
        return obj, loss, it, tv, np.cumsum(per)

    def dynamic_gadmm(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, path, pathCost, tau,
                      rng=None, rechain=None, **kw):
        g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng

        def nxt(i):
            if rechain is not None:
                return rechain(i)
            out = findPath2(N, g)     # same seed on every rank -> same chain everywhere
            return out[0], out[1]

        return self._dynamic(XX, YY, rho, N, d, s, num_iter, obj0, acc, path, np.sum(pathCost),
                             tau, nxt)

    def dynamic_gadmm_v0(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, path_matrix,
                         pathCost_matrix, tau, faithful=True, **kw):
        paths = np.atleast_2d(np.asarray(path_matrix, dtype=np.int64))
        costs = np.atleast_2d(np.asarray(pathCost_matrix, dtype=np.float64))
        kk = [0]

        def nxt(i):
            kk[0] += 1
            return paths[kk[0]], costs[kk[0]]

        c0 = costs[:, 0].sum() if faithful else costs[0].sum()
        return self._dynamic(XX, YY, rho, N, d, s, num_iter, obj0, acc, paths[0], c0, tau, nxt)

    def star_admm(self, XX, YY, rho, N, d, s, num_iter, obj0, acc, **kw):
        from .nccl_star import star_admm
        return star_admm(XX, YY, rho, N, d, s, num_iter, obj0, acc, backend=self.name)


    # Families that live in their own modules (first-order baselines: workers split over the ranks,
    # all_gather / all_reduce only; logistic GADMM: per-rank local solves, all_gather of the model
    # matrix per colour): resolved by name on first use, called with this backend.
    _LAZY = {
        "gd_dgd_lag": ("dist_first_order", "gd_dgd_lag"),
        "gd_dgd_lag_logistic": ("dist_first_order", "gd_dgd_lag_logistic"),
        "dual_averaging": ("dist_first_order", "dual_averaging"),
        "dual_averaging_logistic": ("dist_first_order", "dual_averaging_logistic"),
        "gadmm_logistic": ("dist_logistic", "group_admm_logistic"),
        "gadmm_logistic_gd": ("dist_logistic", "group_admm_logistic_gd"),
    }

    def __getattr__(self, name):
        if name in _Backend._LAZY:
            import functools
            import importlib
            mod, fn = _Backend._LAZY[name]
            f = getattr(importlib.import_module(f"{__package__}.{mod}"), fn)
            return functools.partial(f, backend=self.name)
        raise AttributeError(name)


def get(name: str) -> _Backend:
    return _Backend(name)
