"""Chain topology builders and the link-energy model (host side, float64).

Behavioural parity targets (reference, read-only):
  * ``findPath.m:1-33``   -- N nodes uniform on [0,50]^2, greedy nearest-neighbour chain
    starting at node 1, ``pathCost`` = squared hop lengths.
  * ``findPath2.m:1-62``  -- same on [0,250]^2 with the link-energy model
    ``P = d^2 * eta * B * 2^(R/B)`` and the star model ``P_central = d^2/2 * eta*B*2^(2R/B)``
    around the node closest to the area centre.
  * ``calc_cost.m:1-12``  -- cost of walking a *given* path over a *given* cost grid.

Worker ids are 1-based at this API boundary (``.mat`` parity) and the returned ``path`` is a
permutation of ``1..N`` that begins with 1.  The device-side equivalent (one CTA, warp-shuffle
argmin) lives in ``csrc/topology.cu`` and is exposed as ``gadmm_b200.ops.find_path_device``; both
share :func:`greedy_chain` semantics (first minimum wins, strict ``<`` comparison, the 1e12
sentinel of the reference).
"""
from __future__ import annotations

import numpy as np

# findPath2.m:2-6
AREA2 = 250.0
ETA = 1e-6
RATE = 10e6
BANDWIDTH = 2e6
AREA1 = 50.0  # findPath.m:4-5


def _rng(rng):
    if rng is None:
        return np.random.default_rng()
    if isinstance(rng, (int, np.integer)):
        return np.random.default_rng(int(rng))
    return rng


def pairwise_sq_dist(x: np.ndarray, y: np.ndarray) -> np.ndarray:
    """d_square(n,m) = (x_n-x_m)^2 + (y_n-y_m)^2 with a zero diagonal (findPath.m:8-14)."""
    dx = x[:, None] - x[None, :]
    dy = y[:, None] - y[None, :]
    d2 = dx * dx + dy * dy
    np.fill_diagonal(d2, 0.0)
    return d2


def greedy_chain(d_square: np.ndarray, start: int = 0):
    """Greedy nearest-neighbour chain (0-based) -- findPath.m:16-30.

    Returns the path as an int64 array.  The scan keeps the reference's strict ``temp < min`` test
    with ``min = 1e12`` so ties resolve to the lowest index.
    """
    n_nodes = d_square.shape[0]
    visited = np.zeros(n_nodes, dtype=bool)
    path = [start]
    visited[start] = True
    while len(path) < n_nodes:
        n = path[-1]
        row = np.where(visited, np.inf, d_square[n])
        row[n] = np.inf
        m = int(np.argmin(row))  # first minimum == lowest index, as the reference's strict '<'
        if not row[m] < 1e12:
            raise ValueError("greedy_chain: no reachable unvisited node below the 1e12 sentinel")
        path.append(m)
        visited[m] = True
    return np.asarray(path, dtype=np.int64)


def findPath(N: int, rng=None, points=None):
    """``[path, pathCost, d_square] = findPath(N)`` (findPath.m:1).

    ``rng`` seeds the node positions (the reference is unseeded); ``points`` = (x, y) overrides
    the random draw for testing.  ``path`` is 1-based.
    """
    g = _rng(rng)
    if points is None:
        # the reference draws x(n), y(n) interleaved per node (findPath.m:3-6)
        xy = g.random((N, 2)) * AREA1
        x, y = xy[:, 0].copy(), xy[:, 1].copy()
    else:
        x, y = (np.asarray(p, dtype=np.float64) for p in points)
    d2 = pairwise_sq_dist(x, y)
    p0 = greedy_chain(d2, 0)
    cost = d2[p0[:-1], p0[1:]]
    return p0 + 1, cost, d2


def findPath2(N: int, rng=None, points=None):
    """``[path, pathCost, d_square, P_central, center] = findPath2(N)`` (findPath2.m:1)."""
    g = _rng(rng)
    if points is None:
        xy = g.random((N, 2)) * AREA2
        x, y = xy[:, 0].copy(), xy[:, 1].copy()
    else:
        x, y = (np.asarray(p, dtype=np.float64) for p in points)
    # centre node: first node strictly closer than all previous (findPath2.m:16-23)
    dist_c = np.sqrt((x - AREA2 / 2) ** 2 + (y - AREA2 / 2) ** 2)
    center0 = int(np.argmin(dist_c))
    d2c = (x - x[center0]) ** 2 + (y - y[center0]) ** 2
    P_central = 0.5 * d2c * ETA * BANDWIDTH * 2.0 ** (2 * RATE / BANDWIDTH)
    d2 = pairwise_sq_dist(x, y)
    P = d2 * ETA * BANDWIDTH * 2.0 ** (RATE / BANDWIDTH)
    p0 = greedy_chain(d2, 0)
    cost = P[p0[:-1], p0[1:]]
    return p0 + 1, cost, d2, P_central, center0 + 1


def calc_cost(grid: np.ndarray, path0) -> np.ndarray:
    """``pathCost_static = calc_cost(grid, path0)`` (calc_cost.m:1-12); ``path0`` is 1-based."""
    p = np.asarray(path0, dtype=np.int64).ravel() - 1
    grid = np.asarray(grid)
    return grid[p[:-1], p[1:]].astype(np.float64)


def chain_tables(path0: np.ndarray):
    """Neighbour tables of a 0-based chain ``path0`` (a permutation of 0..N-1).

    Returns ``left, right, color, cmul`` indexed by *worker id*:
      left/right -- worker id of the predecessor / successor on the chain, -1 at the ends;
      color      -- 0 for heads (odd positions of ``path`` in the reference's 1-based counting,
                    dynamic_group_ADMM_closedForm.m:23), 1 for tails (``:103``);
      cmul       -- 1 at the chain ends, 2 inside (the ``c_n`` of ``(G_n + c_n rho I)``,
                    group_ADMM_closedForm.m:39-46).
    """
    path0 = np.asarray(path0, dtype=np.int64)
    n = path0.shape[0]
    left = np.full(n, -1, dtype=np.int32)
    right = np.full(n, -1, dtype=np.int32)
    color = np.zeros(n, dtype=np.int32)
    cmul = np.full(n, 2, dtype=np.int32)
    left[path0[1:]] = path0[:-1]
    right[path0[:-1]] = path0[1:]
    color[path0] = np.arange(n) % 2
    cmul[path0[0]] = 1
    cmul[path0[-1]] = 1
    return left, right, color, cmul
