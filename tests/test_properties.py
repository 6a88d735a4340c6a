"""Property-based checks (hypothesis) of the host-side building blocks every backend relies on:
chain tables, worker partition, the greedy chain, communication accounting and the algebraic
identities between the GADMM variants on random small problems."""
import numpy as np
from hypothesis import given, settings, strategies as st

from gadmm_b200 import oracle as O
from gadmm_b200.parallel.runtime import partition
from gadmm_b200.results import cumulative_com, iterations_to
from gadmm_b200.topology import chain_tables, findPath, findPath2, greedy_chain, pairwise_sq_dist

perm = st.integers(min_value=1, max_value=40).flatmap(
    lambda n: st.permutations(list(range(n))))


@given(perm)
@settings(max_examples=80, deadline=None)
def test_chain_tables_describe_the_path(path):
    path = np.asarray(path)
    left, right, color, cmul = chain_tables(path)
    N = path.size
    pos = np.empty(N, dtype=int)
    pos[path] = np.arange(N)
    for n in range(N):
        p = pos[n]
        assert left[n] == (path[p - 1] if p > 0 else -1)
        assert right[n] == (path[p + 1] if p < N - 1 else -1)
        assert color[n] == p % 2                      # heads = odd positions (1-based) of the path
        assert cmul[n] == (1 if p in (0, N - 1) else 2)   # c = 1 at the chain ends (reference :39-45)
    # neighbours always have the other colour: the two-colour schedule is valid on any chain
    for n in range(N):
        for v in (left[n], right[n]):
            if v >= 0:
                assert color[v] != color[n]


@given(st.integers(1, 300), st.integers(1, 16))
@settings(max_examples=100, deadline=None)
def test_partition_is_contiguous_and_balanced(n, world):
    owner = partition(n, world)
    assert owner.shape == (n,) and np.all(np.diff(owner) >= 0)
    counts = np.bincount(owner, minlength=world)
    assert counts.sum() == n and counts.max() - counts.min() <= 1
    assert np.all(np.diff(counts) <= 0)               # the first n % world ranks hold the extra worker


@given(st.integers(2, 30), st.integers(0, 2 ** 31 - 1))
@settings(max_examples=60, deadline=None)
def test_greedy_chain_is_a_permutation_from_node_one(n, seed):
    rng = np.random.default_rng(seed)
    pts = rng.random((n, 2)) * 50
    d2 = pairwise_sq_dist(pts[:, 0], pts[:, 1])
    path = greedy_chain(d2)
    cost = np.array([d2[path[k], path[k + 1]] for k in range(n - 1)])
    assert path[0] == 0 and sorted(path.tolist()) == list(range(n))
    for k in range(n - 1):                            # each hop goes to the nearest unvisited node
        rest = [m for m in range(n) if m not in path[:k + 1].tolist()]
        assert abs(cost[k] - min(d2[path[k], m] for m in rest)) <= 1e-12 * max(1.0, cost[k])
    p1, c1, _ = findPath(n, rng=seed)
    p2, c2, *_ = findPath2(n, rng=seed)
    assert p1[0] == 1 and p2[0] == 1 and len(c1) == len(c2) == n - 1


@given(st.integers(1, 500), st.integers(1, 64))
@settings(max_examples=50, deadline=None)
def test_communication_accounting(n_iter, workers):
    it = np.arange(1, n_iter + 1)
    np.testing.assert_array_equal(cumulative_com("GD", n_iter, workers), it * workers + it)
    np.testing.assert_array_equal(cumulative_com("GADMM", n_iter, workers), it * workers)
    up = np.cumsum(np.ones(n_iter))
    np.testing.assert_array_equal(cumulative_com("LAG_WK", n_iter, workers, up), up + it)
    assert iterations_to(1.0 / it, 0.5) == min(3, n_iter) if n_iter >= 3 else True


@given(st.integers(2, 7), st.integers(2, 6), st.integers(0, 10 ** 6), st.floats(0.2, 8.0))
@settings(max_examples=25, deadline=None)
def test_edge_dual_node_dual_and_identity_chain_agree(N, d, seed, rho):
    rng = np.random.default_rng(seed)
    s = d + 2
    XX, YY = rng.standard_normal((N * s, d)), rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    a = O.group_ADMM_closedForm(XX, YY, rho, N, d, s, 25, obj0, 0.0)
    b = O.static_group_ADMM_closedForm(XX, YY, rho, N, d, s, 25, obj0, 0.0, 5, np.ones((6, N - 1)))
    c = O.dynamic_group_ADMM_closedForm(XX, YY, rho, N, d, s, 25, obj0, 0.0, np.arange(1, N + 1),
                                        np.ones(N - 1), 10 ** 9)
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(a[0], c[0], rtol=1e-9, atol=1e-10)
    assert np.all(np.isfinite(a[0])) and a[0].shape == (25,)
