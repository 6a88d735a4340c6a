"""Exhaustive interleaving check of the cross-GPU hand-off protocol of the GADMM hot loop
(compute-sanitizer is closed on this pool, so the protocol is checked as a
model instead).

Model of csrc/gadmm_iter.cu: every worker is its own asynchronous process (the worst case: every
worker on a different GPU, no kernel boundary ordering anything).  Per iteration ``it`` a worker
  1. waits until both neighbours' version flags are ``>= need`` (heads: ``it-1``, tails: ``it``),
  2. reads the neighbours' models from replica slot ``need & 1``,
  3. writes its own model into slot ``it & 1`` (two steps: a torn write is observable),
  4. releases its flag ``= it``,
  5. dual update: waits for the neighbours' flags ``>= it`` and reads THEIR models of iteration
     ``it`` (slot ``it & 1``) -- this is the reader that a neighbour's next head phase can overlap.
The checker explores ALL interleavings (depth-first over the product state space) and fails if a
read ever observes anything but the complete model of iteration ``need``.  With the double-buffered
replica the protocol is safe WITHOUT any global barrier; with a single slot the same checker finds
the overwrite race, which shows it has teeth.
"""
import sys

import pytest


def _explore(n_workers: int, n_iter: int, slots: int):
    """Returns (states_visited, violation or None)."""
    colour = [w % 2 for w in range(n_workers)]
    nbrs = [[v for v in (w - 1, w + 1) if 0 <= v < n_workers] for w in range(n_workers)]
    # per worker: (it, pc); pc 0 = wait+read (one neighbour at a time: pc 0..len(nbrs)-1),
    # then write-begin, write-end, release.  Slot content: version (int) or -1 while being written.
    init_slots = tuple(tuple(0 for _ in range(slots)) for _ in range(n_workers))   # version 0 = theta0
    init = (tuple((1, 0) for _ in range(n_workers)), init_slots, tuple(0 for _ in range(n_workers)))
    seen = {init}
    stack = [init]
    sys.setrecursionlimit(10000)
    while stack:
        procs, mem, flags = stack.pop()
        enabled = 0
        for w in range(n_workers):
            it, pc = procs[w]
            if it > n_iter:
                continue
            need = it - 1 if colour[w] == 0 else it
            k = len(nbrs[w])
            nmem, nflags = mem, flags
            if pc < k:                                   # wait for + read neighbour pc
                v = nbrs[w][pc]
                if flags[v] < need:
                    continue                             # blocked on the flag
                got = mem[v][need % slots]
                if got != need:
                    return len(seen), (w, it, v, need, got)
                nxt = (it, pc + 1)
            elif pc == k:                                # begin writing own slot
                row = list(mem[w])
                row[it % slots] = -1
                nmem = mem[:w] + (tuple(row),) + mem[w + 1:]
                nxt = (it, pc + 1)
            elif pc == k + 1:                            # write complete
                row = list(mem[w])
                row[it % slots] = it
                nmem = mem[:w] + (tuple(row),) + mem[w + 1:]
                nxt = (it, pc + 1)
            elif pc == k + 2:                            # release the version flag
                nflags = flags[:w] + (it,) + flags[w + 1:]
                nxt = (it, pc + 1)
            else:                                        # dual update: neighbours' models of THIS iteration
                v = nbrs[w][pc - (k + 3)]
                if flags[v] < it:
                    continue
                got = mem[v][it % slots]
                if got != it:
                    return len(seen), (w, it, v, it, got)
                nxt = (it, pc + 1) if pc - (k + 3) + 1 < k else (it + 1, 0)
            enabled += 1
            state = (procs[:w] + (nxt,) + procs[w + 1:], nmem, nflags)
            if state not in seen:
                seen.add(state)
                stack.append(state)
        if enabled == 0 and any(it <= n_iter for it, _ in procs):
            return len(seen), ("deadlock", procs, flags, None, None)
    return len(seen), None


@pytest.mark.parametrize("n_workers,n_iter", [(2, 4), (3, 4), (4, 3), (5, 2)])
def test_double_buffered_replica_is_safe_under_every_interleaving(n_workers, n_iter):
    states, bad = _explore(n_workers, n_iter, slots=2)
    assert bad is None, f"worker {bad[0]} it {bad[1]} read worker {bad[2]}: wanted v{bad[3]}, saw {bad[4]}"
    assert states >= 30


def test_single_slot_replica_is_caught_by_the_checker():
    """Negative control: without the iteration-parity double buffer a head's next write lands on
    the model a neighbour's dual update has yet to read -- the checker must find that interleaving."""
    _, bad = _explore(3, 3, slots=1)
    assert bad is not None


def test_no_deadlock_all_workers_finish():
    """Every maximal execution ends with all workers past the last iteration (the DFS above only
    stops at states without enabled steps; check that the only such state is the final one)."""
    n_workers, n_iter = 4, 3
    colour = [w % 2 for w in range(n_workers)]
    # replay one greedy schedule: always run the lowest-index enabled worker
    flags = [0] * n_workers
    it = [1] * n_workers
    progress = True
    while progress:
        progress = False
        for w in range(n_workers):
            if it[w] > n_iter:
                continue
            need = it[w] - 1 if colour[w] == 0 else it[w]
            if all(flags[v] >= need for v in (w - 1, w + 1) if 0 <= v < n_workers):
                flags[w] = it[w]
                it[w] += 1
                progress = True
    assert all(i == n_iter + 1 for i in it)


# ------------------------------------------------------------------------------------------------
# The per-iteration objective exchange (gadmm_dual_objective_kernel): every rank stores its partial
# objective into slot ``it % RING`` of EVERY rank (entry = writer's rank), releases a flag per
# destination, then waits for all ranks' flags ``>= it`` and reads their entries.  It doubles as the
# bound on rank skew (nobody starts iteration it+1 before everybody has published iteration it).

def _explore_ring(n_ranks: int, n_iter: int, ring: int, invariant=None):
    """Returns (states_visited, violation or None); ``invariant(procs)`` is asserted in every
    reachable state."""
    # memory[dst][slot][src] = version (or -1 while torn), flag[dst][slot][src] = iteration
    z = tuple(tuple(tuple(0 for _ in range(n_ranks)) for _ in range(ring)) for _ in range(n_ranks))
    init = (tuple((1, 0) for _ in range(n_ranks)), z, z)
    seen, stack = {init}, [init]

    def put(a, dst, slot, src, val):
        row = list(a[dst][slot])
        row[src] = val
        sl = list(a[dst])
        sl[slot] = tuple(row)
        return a[:dst] + (tuple(sl),) + a[dst + 1:]

    while stack:
        procs, mem, flg = stack.pop()
        if invariant is not None:
            invariant(procs)
        enabled = 0
        for r in range(n_ranks):
            it, pc = procs[r]
            if it > n_iter:
                continue
            slot = it % ring
            nmem, nflg = mem, flg
            if pc < 2 * n_ranks:                       # store to destination pc//2 in two steps
                dst = pc // 2
                nmem = put(mem, dst, slot, r, -1 if pc % 2 == 0 else it)
                nxt = (it, pc + 1)
            elif pc < 3 * n_ranks:                     # release the flag at each destination
                dst = pc - 2 * n_ranks
                nflg = put(flg, dst, slot, r, it)
                nxt = (it, pc + 1)
            else:                                      # wait for + read source q's entry here
                q = pc - 3 * n_ranks
                if flg[r][slot][q] < it:
                    continue
                got = mem[r][slot][q]
                if got != it:
                    return len(seen), (r, it, q, got)
                nxt = (it, pc + 1) if q + 1 < n_ranks else (it + 1, 0)
            enabled += 1
            state = (procs[:r] + (nxt,) + procs[r + 1:], nmem, nflg)
            if state not in seen:
                seen.add(state)
                stack.append(state)
        if enabled == 0 and any(it <= n_iter for it, _ in procs):
            return len(seen), ("deadlock", procs, None, None)
    return len(seen), None


@pytest.mark.parametrize("n_ranks,n_iter,ring", [(2, 6, 4), (3, 3, 4), (2, 6, 2)])
def test_objective_exchange_ring_is_safe(n_ranks, n_iter, ring):
    states, bad = _explore_ring(n_ranks, n_iter, ring)
    assert bad is None, bad
    assert states > 50


def test_objective_exchange_needs_more_than_one_slot():
    """Negative control: with a single slot a fast rank's next partial objective overwrites the one
    a slower rank has not read yet (the flag only says 'written', not 'consumed')."""
    _, bad = _explore_ring(2, 3, 1)
    assert bad is not None and bad[0] != "deadlock"


def test_objective_exchange_bounds_rank_skew():
    """A rank finishes iteration it only after reading every rank's iteration-it entry, so in every
    reachable state the ranks are at most one iteration apart."""
    def skew(procs):
        its = [it for it, _ in procs]
        assert max(its) - min(its) <= 1, procs

    _, bad = _explore_ring(3, 3, 4, invariant=skew)
    assert bad is None


# ------------------------------------------------------------------------------------------------
# D-GADMM: the whole iteration with a NEW chain every iteration (tau = 1), one worker per rank.
# Per iteration a rank (1) pulls the latest models (version it-1) of its new neighbours straight out
# of the owners' replicas (csrc/rechain.cu), (2) runs its half-iteration in the colour the new
# chain gives it, (3) does the dual update, (4) takes part in the objective exchange.  The exchange
# is the only thing that keeps a fast rank from overwriting slot (it-1)&1 while a slow rank still
# pulls from it, so the composite model is checked with and without it.

def _explore_dgadmm(chains, n_iter: int, with_exchange: bool):
    n = len(chains[0])

    def program(rank, it):
        chain = chains[(it - 1) % len(chains)]
        pos = chain.index(rank)
        nb = [chain[p] for p in (pos - 1, pos + 1) if 0 <= p < n]
        ops = [("read", q, it - 1, it - 1) for q in nb]                 # re-chain pull (flag >= it-1)
        if pos % 2 == 1:                                                # tail: heads of THIS iteration
            ops += [("read", q, it, it) for q in nb]
        ops += [("wbegin",), ("wend",), ("release",)]
        ops += [("read", q, it, it) for q in nb]                        # dual update
        if with_exchange:
            ops += [("xput", q) for q in range(n)] + [("xget", q) for q in range(n)]
        return ops

    theta0 = tuple((0, 0) for _ in range(n))                            # per owner: (slot0, slot1)
    zero = tuple(0 for _ in range(n))
    xz = tuple(tuple(0 for _ in range(n)) for _ in range(n))            # xflag[dst][src]
    init = (tuple((1, 0) for _ in range(n)), theta0, zero, xz)
    seen, stack = {init}, [init]
    while stack:
        procs, th, fl, xf = stack.pop()
        enabled = 0
        for r in range(n):
            it, pc = procs[r]
            if it > n_iter:
                continue
            ops = program(r, it)
            op = ops[pc]
            nth, nfl, nxf = th, fl, xf
            if op[0] == "read":
                _, q, need, expect = op
                if fl[q] < need:
                    continue
                got = th[q][expect % 2]
                if got != expect:
                    return len(seen), (r, it, q, expect, got)
            elif op[0] in ("wbegin", "wend"):
                row = list(th[r])
                row[it % 2] = -1 if op[0] == "wbegin" else it
                nth = th[:r] + (tuple(row),) + th[r + 1:]
            elif op[0] == "release":
                nfl = fl[:r] + (it,) + fl[r + 1:]
            elif op[0] == "xput":
                row = list(xf[op[1]])
                row[r] = it
                nxf = xf[:op[1]] + (tuple(row),) + xf[op[1] + 1:]
            elif op[0] == "xget":
                if xf[r][op[1]] < it:
                    continue
            nxt = (it, pc + 1) if pc + 1 < len(ops) else (it + 1, 0)
            enabled += 1
            state = (procs[:r] + (nxt,) + procs[r + 1:], nth, nfl, nxf)
            if state not in seen:
                seen.add(state)
                stack.append(state)
        if enabled == 0 and any(it <= n_iter for it, _ in procs):
            return len(seen), ("deadlock", procs, fl, None, None)
    return len(seen), None


_CHAINS3 = [[0, 1, 2], [1, 0, 2], [2, 1, 0], [0, 2, 1]]
_CHAINS4 = [[0, 1, 2, 3], [0, 2, 1, 3], [3, 0, 2, 1], [1, 3, 0, 2]]


@pytest.mark.parametrize("chains,n_iter", [(_CHAINS3, 4), (_CHAINS4, 3)])
def test_dgadmm_rechain_every_iteration_is_safe_with_the_exchange(chains, n_iter):
    states, bad = _explore_dgadmm(chains, n_iter, with_exchange=True)
    assert bad is None, bad
    assert states > 100


def test_dgadmm_rechain_without_the_exchange_is_caught():
    """Negative control: drop the per-iteration exchange and a rank that is two iterations ahead
    overwrites the replica slot a slower rank is still reading (the checker finds a torn read of
    rank 3's iteration-1 model while rank 3 writes iteration 3) -- the reason the stop-test
    exchange stays on every iteration of a re-chained run."""
    _, bad = _explore_dgadmm(_CHAINS4, 4, with_exchange=False)
    assert bad is not None and bad[0] != "deadlock" and bad[4] == -1


# ------------------------------------------------------------------------------------------------
# Round 2: whole windows run under programmatic dependent launch, so a worker's SOLVE (streaming
# kernel + finisher) and its DUAL UPDATE (gadmm_dual_objective_kernel) are different CTAs of
# different, overlapping kernels.  They are ordered only by data: the solve of iteration it waits
# for nuflag[w] >= it-1 (csrc/gadmm_iter.cu: wait_flag_ge_gpu(c.nuflag + k, it - 1)), the dual update
# of iteration it waits for the version flags of w and of its neighbours >= it.  Model: two
# asynchronous processes per worker, every interleaving explored.
def _explore_split(n_workers: int, n_iter: int, use_nuflag: bool = True):
    colour = [w % 2 for w in range(n_workers)]
    nbrs = [[v for v in (w - 1, w + 1) if 0 <= v < n_workers] for w in range(n_workers)]
    init = (tuple((1, 0) for _ in range(n_workers)),            # solver  (it, pc)
            tuple((1, 0) for _ in range(n_workers)),            # dual    (it, pc)
            tuple((0, 0) for _ in range(n_workers)),            # theta slots (versions)
            tuple(0 for _ in range(n_workers)),                 # version flags
            tuple(0 for _ in range(n_workers)),                 # nu version (-1 = torn)
            tuple(0 for _ in range(n_workers)))                 # nuflag
    seen = {init}
    stack = [init]
    while stack:
        S, D, mem, flags, nu, nuflag = stack.pop()
        moved = False
        for w in range(n_workers):
            k = len(nbrs[w])
            # ---------------- solver process
            it, pc = S[w]
            if it <= n_iter:
                need = it - 1 if colour[w] == 0 else it
                nS, nmem, nflags = None, mem, flags
                if pc == 0:                                   # own dual of iteration it-1
                    if not use_nuflag or nuflag[w] >= it - 1:
                        if nu[w] != it - 1:
                            return len(seen), ("nu", w, it, nu[w])
                        nS = (it, 1)
                elif pc <= k:
                    v = nbrs[w][pc - 1]
                    if flags[v] >= need:
                        if mem[v][need % 2] != need:
                            return len(seen), ("theta-solve", w, it, v, mem[v][need % 2])
                        nS = (it, pc + 1)
                elif pc == k + 1:
                    row = list(mem[w]); row[it % 2] = -1
                    nmem = mem[:w] + (tuple(row),) + mem[w + 1:]
                    nS = (it, pc + 1)
                elif pc == k + 2:
                    row = list(mem[w]); row[it % 2] = it
                    nmem = mem[:w] + (tuple(row),) + mem[w + 1:]
                    nS = (it, pc + 1)
                else:
                    nflags = flags[:w] + (it,) + flags[w + 1:]
                    nS = (it + 1, 0)
                if nS is not None:
                    moved = True
                    st = (S[:w] + (nS,) + S[w + 1:], D, nmem, nflags, nu, nuflag)
                    if st not in seen:
                        seen.add(st); stack.append(st)
            # ---------------- dual process
            it, pc = D[w]
            if it <= n_iter:
                readers = [w] + nbrs[w]
                nD, nnu, nnf = None, nu, nuflag
                if pc < len(readers):
                    v = readers[pc]
                    if flags[v] >= it:
                        if mem[v][it % 2] != it:
                            return len(seen), ("theta-dual", w, it, v, mem[v][it % 2])
                        nD = (it, pc + 1)
                elif pc == len(readers):
                    nnu = nu[:w] + (-1,) + nu[w + 1:]
                    nD = (it, pc + 1)
                elif pc == len(readers) + 1:
                    nnu = nu[:w] + (it,) + nu[w + 1:]
                    nD = (it, pc + 1)
                else:
                    nnf = nuflag[:w] + (it,) + nuflag[w + 1:]
                    nD = (it + 1, 0)
                if nD is not None:
                    moved = True
                    st = (S, D[:w] + (nD,) + D[w + 1:], mem, flags, nnu, nnf)
                    if st not in seen:
                        seen.add(st); stack.append(st)
        if not moved and (any(i <= n_iter for i, _ in S) or any(i <= n_iter for i, _ in D)):
            return len(seen), ("deadlock", S, D, flags, nuflag)
    return len(seen), None


@pytest.mark.parametrize("n_workers,n_iter", [(2, 4), (3, 3)])
def test_window_protocol_solver_and_dual_as_separate_processes(n_workers, n_iter):
    states, bad = _explore_split(n_workers, n_iter, use_nuflag=True)
    assert bad is None, bad
    assert states > 100


def test_window_protocol_without_the_dual_flag_is_caught():
    """Negative control: if the next solve did not wait for nuflag, it could assemble its right-hand
    side from a node dual that is stale or half written."""
    _, bad = _explore_split(2, 3, use_nuflag=False)
    assert bad is not None and bad[0] == "nu"


# ------------------------------------------------------------------------------------------------
# DGD across GPUs (csrc/first_order.cu: own_exchange_kernel / dgd_update_kernel): the neighbour's
# boundary gradient row lands in a SINGLE-buffered halo row.  Per iteration a rank (one stream, in
# order): posts its objective partial fx(it); pushes its boundary row (version it) into the
# neighbour; update(it) waits for the neighbour's row flag >= it and reads the halo.  The push of
# iteration it waits for the neighbour's fx post of iteration it (back-pressure): that post comes
# after the neighbour's update(it-1) in its stream, so the halo row of it-1 has been consumed.
def _explore_dgd(n_iter: int, backpressure: bool):
    # two ranks; state per rank: (it, pc); halo[r] = version of the row that rank r RECEIVED
    init = (((1, 0), (1, 0)), (0, 0), (0, 0), (0, 0))       # procs, halo, rowflag, fxpost
    seen = {init}
    stack = [init]
    while stack:
        procs, halo, rowflag, fx = stack.pop()
        moved = False
        for r in (0, 1):
            o = 1 - r
            it, pc = procs[r]
            if it > n_iter:
                continue
            nh, nrf, nfx, nxt = halo, rowflag, fx, None
            if pc == 0:                                    # post fx(it)
                nfx = tuple(it if q == r else fx[q] for q in (0, 1))
                nxt = (it, 1)
            elif pc == 1:                                  # push own row into the peer's halo
                if not backpressure or fx[o] >= it:
                    nh = tuple(it if q == o else halo[q] for q in (0, 1))
                    nxt = (it, 2)
            elif pc == 2:                                  # release the row flag at the peer
                nrf = tuple(it if q == o else rowflag[q] for q in (0, 1))
                nxt = (it, 3)
            else:                                          # update(it): wait for the row, read it
                if rowflag[r] >= it:
                    if halo[r] != it:
                        return len(seen), (r, it, halo[r])
                    nxt = (it + 1, 0)
            if nxt is not None:
                moved = True
                st = (tuple(nxt if q == r else procs[q] for q in (0, 1)), nh, nrf, nfx)
                if st not in seen:
                    seen.add(st); stack.append(st)
        if not moved and any(i <= n_iter for i, _ in procs):
            return len(seen), ("deadlock", procs)
    return len(seen), None


def test_dgd_halo_with_backpressure_is_safe_and_without_is_caught():
    _, bad = _explore_dgd(5, backpressure=True)
    assert bad is None, bad
    _, bad = _explore_dgd(5, backpressure=False)
    assert bad is not None and bad[0] != "deadlock"       # reads the row of iteration it+1


# ------------------------------------------------------------------------------------------------
# Tile ring of the streaming SYMV (csrc/symv_core.cuh, producer loop): the producer warp issues the
# bulk copy of a tile one GROUP ahead of the tile it hands to the consumers.  Model: the producer
# runs the op list of the loop, the copy engine completes issued copies in any order, the
# consumers take the tiles in order.  Checked: no deadlock, no stage (tile buffer or RHS slice)
# overwritten before its previous tile was consumed, the consumers see tile t in stage t % STAGES.
def _producer_ops(total: int, G: int):
    ops = [("issue", g) for g in range(min(G, total))]
    for t0 in range(0, total, G):
        for g in range(G):
            if t0 + G + g < total:
                ops.append(("issue", t0 + G + g))
            if t0 + g < total:
                ops.append(("hand", t0 + g))
    return ops


def _explore_tile_ring(total: int, stages: int, G: int):
    ops = _producer_ops(total, G)
    # state: producer pc, frozenset of copies in flight, per-stage (tile whose data landed or -1),
    # per-stage (tile whose slice was written or -1), consumed count
    init = (0, frozenset(), (-1,) * stages, (-1,) * stages, 0)
    seen = {init}
    stack = [init]
    while stack:
        pc, flight, data, xs, done = stack.pop()
        succ = []
        if pc < len(ops):
            kind, t = ops[pc]
            s = t % stages
            if kind == "issue":
                if t < stages or done > t - stages:          # empty[s] of the previous user
                    succ.append((pc + 1, flight | {t}, data, xs, done))
            else:
                if xs[s] >= 0 and done <= xs[s]:             # slice of an unconsumed tile
                    return len(seen), ("slice overwritten", t, xs[s])
                succ.append((pc + 1, flight, data, tuple(t if q == s else xs[q] for q in range(stages)), done))
        for t in flight:                                     # a copy lands
            s = t % stages
            if data[s] >= 0 and done <= data[s]:
                return len(seen), ("tile overwritten", t, data[s])
            succ.append((pc, flight - {t}, tuple(t if q == s else data[q] for q in range(stages)), xs, done))
        if done < total:                                     # consumers: full[s] = copy landed + hand-off
            s = done % stages
            if data[s] == done and xs[s] == done:
                succ.append((pc, flight, data, xs, done + 1))
            elif data[s] > done or xs[s] > done:
                return len(seen), ("wrong tile in stage", done, data[s], xs[s])
        if not succ and (done < total or pc < len(ops)):
            return len(seen), ("deadlock", pc, done)
        for st in succ:
            if st not in seen:
                seen.add(st)
                stack.append(st)
    return len(seen), None


@pytest.mark.parametrize("stages,G", [(3, 1), (3, 2), (6, 3), (6, 5), (12, 3)])
@pytest.mark.parametrize("total", [0, 1, 2, 7, 13])
def test_tile_ring_with_lookahead_is_safe(total, stages, G):
    _, bad = _explore_tile_ring(total, stages, G)
    assert bad is None, bad


def test_tile_ring_group_as_deep_as_the_ring_deadlocks():
    # issuing tile t + G needs the stage of tile t + G - STAGES back, i.e. G <= STAGES - 1
    # (the kernel clamps G accordingly)
    _, bad = _explore_tile_ring(9, 3, 3)
    assert bad is not None and bad[0] == "deadlock"
