"""Result records and the reference's communication-cost accounting.

Unit = "one worker transmission" (LinearRegression_Synthetic.m:100-142):
  GD ``iter*N + iter``; DGD ``iter*N``; LAG ``comm_iter_final(iter) + iter``; dual averaging
  ``iter*N``; GADMM ``iter*N``.  Energy-weighted variants: chain ``sum(pathCost)`` per iteration
  (LinearRegression_gadmm_vs_admm.m:143-151), star ``sum(P_central)+max(P_central)`` per iteration
  (``:84-86``), D-GADMM ``x(1+5/15)`` re-chaining overhead on cost and time
  (Dynamic_LinearRegression_Synthetic.m:123,125).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np

from .io.matio import save_curves

RECHAIN_OVERHEAD = 5.0 / 15.0


@dataclass
class Result:
    """One algorithm's curves with the reference's variable names."""

    name: str
    obj: np.ndarray
    loss: np.ndarray
    Iter: int
    time: Optional[np.ndarray] = None          # simulated or device clock, cumulative seconds
    com_cost: Optional[np.ndarray] = None      # cumulative energy cost (D-GADMM family)
    comm_iter: Optional[np.ndarray] = None     # LAG upload counter
    extra: Dict[str, object] = field(default_factory=dict)

    def curves(self) -> dict:
        out = {f"obj_{self.name}": self.obj, f"loss_{self.name}": self.loss,
               f"Iter_{self.name}": np.array([self.Iter], dtype=np.float64)}
        if self.time is not None:
            out[f"time_{self.name}"] = self.time
        if self.com_cost is not None:
            out[f"com_cost_{self.name}"] = self.com_cost
        if self.comm_iter is not None:
            out[f"comm_iter_final_{self.name}"] = self.comm_iter
        return out

    def save(self, path: str):
        save_curves(path, self.curves())


def cumulative_com(kind: str, n_iter: int, num_workers: int, comm_iter_final=None) -> np.ndarray:
    it = np.arange(1, int(n_iter) + 1, dtype=np.float64)
    if kind == "GD":
        return it * num_workers + it
    if kind in ("DGD", "dualAvg", "GADMM", "ADMM_rounds"):
        return it * num_workers
    if kind in ("LAG_PS", "LAG_WK"):
        return np.asarray(comm_iter_final, dtype=np.float64)[: int(n_iter)] + it
    if kind in ("cyclic_IAG", "R_IAG"):
        return 2.0 * it   # one upload + one broadcast per iteration (not plotted by the reference)
    raise ValueError(kind)


def error_percent(loss, opt_obj) -> np.ndarray:
    """``errorPer = |loss/opt_obj|*100`` (LinearRegression_Synthetic.m:102)."""
    return np.abs(np.asarray(loss) / np.asarray(opt_obj)) * 100.0


def energy_cumulative(per_iter_cost: float, n_iter: int) -> np.ndarray:
    return per_iter_cost * np.arange(1, int(n_iter) + 1, dtype=np.float64)


def star_cost(P_central) -> float:
    P = np.asarray(P_central, dtype=np.float64)
    return float(P.sum() + P.max())


def with_rechain_overhead(v) -> np.ndarray:
    v = np.asarray(v, dtype=np.float64)
    return v + RECHAIN_OVERHEAD * v


def iterations_to(loss, acc: float) -> int:
    """First 1-based iteration with ``loss < acc`` or ``len(loss)`` if never."""
    loss = np.asarray(loss)
    hit = np.nonzero(loss < acc)[0]
    return int(hit[0]) + 1 if hit.size else int(loss.size)
