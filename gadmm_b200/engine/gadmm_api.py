"""Reference-signature wrappers over :class:`ChainEngine` (linear closed-form GADMM family)."""
from __future__ import annotations

import numpy as np
import torch

from ..parallel.runtime import Dist
from ..topology import findPath2
from .chain import ChainEngine


def _engine(XX, YY, rho, N, d, s, num_iter, obj0, acc, dynamic, store="f64", dist=None,
            theta0=None):
    if dist is None and not dynamic and _small_enough(N, d):
        from ..parallel.runtime import SoloDist
        dist = SoloDist()      # tiny problem: every rank runs the persistent single-GPU kernel
    eng = ChainEngine(N, d, rho, obj0, acc, num_iter, dist=dist or Dist(), store=store,
                      dynamic=dynamic)
    eng.ingest_host(XX, YY, s)
    if theta0 is not None:
        eng.set_theta(theta0)
    return eng


def _small_enough(N, d) -> bool:
    """Problems the persistent kernel can hold (one CTA per worker, dp <= 128)."""
    import torch
    try:
        sms = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
    except Exception:
        return False
    return N <= sms and (d + 63) // 64 * 64 <= 128


def _energy(cost_of_iter, Iter):
    c = np.array([cost_of_iter(i) for i in range(1, Iter + 1)], dtype=np.float64)
    return np.cumsum(c)


def gadmm_closed_form(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                      store="f64", window=32, dist=None, theta0=None, return_engine=False,
                      state=None, return_state=False, clock="device"):
    """``group_ADMM_closedForm`` on the GPU.  Node-dual form internally; the iterates equal the
    edge-dual ones (nu_n = lambda_n - lambda_{n-1})."""
    eng = _engine(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc, False,
                  store, dist, theta0)
    eng.set_path(np.arange(no_workers))
    if state is not None:
        eng.load_state(state)
    it = eng.run(num_iter, window=window)
    res = eng.results(it, clock=clock)
    if return_state:
        res = res + (eng.state(),)
    return res + (eng,) if return_engine else res


def static_gadmm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                 coherence_Time, pathCost_static_matrix, store="f64", window=32, dist=None,
                 return_engine=False):
    costs = np.atleast_2d(np.asarray(pathCost_static_matrix, dtype=np.float64))
    eng = _engine(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc, False,
                  store, dist)
    eng.set_path(np.arange(no_workers))
    it = eng.run(num_iter, window=window)
    obj, loss, it, tv = eng.results(it)
    tau = coherence_Time

    def cost(i):
        # row index = number of refreshes so far: j in [2, i] with j % tau == 0
        kk = int(i // tau) - (1 if tau == 1 else 0)
        return float(costs[min(kk, costs.shape[0] - 1)].sum())

    res = (obj, loss, it, tv, _energy(cost, it))
    return res + (eng,) if return_engine else res


def _dynamic(XX, YY, rho, N, d, s, num_iter, obj0, acc, first_path, first_cost, tau, next_chain,
             store, window, dist, return_engine, rechain_mode="device", device_chain_seed=None):
    """``rechain_mode="device"``: chains are uploaded in chunks and switched by the re-chain kernels
    inside the CUDA graph (no host round trip per window); ``"host"``: tables are rebuilt on the host
    at every re-chain (kept for comparison / debugging)."""
    eng = _engine(XX, YY, rho, N, d, s, num_iter, obj0, acc, True, store, dist)
    eng.set_path(np.asarray(first_path, dtype=np.int64) - 1)
    cost_at = {1: float(first_cost)}
    tau_i = None if tau >= 1e8 else int(tau)

    def rechain(i):
        p1, pc = next_chain(i)
        cost_at[i] = float(np.sum(pc))
        return np.asarray(p1, dtype=np.int64) - 1

    if rechain_mode == "device" and device_chain_seed is not None and tau_i is not None:
        # chains are drawn and built on the GPU too (Philox positions + one CTA per chain)
        from ..ops.topology import find_path_batch_device
        chunks = []

        def chunk_source(base, K):
            paths, sums = find_path_batch_device(K, N, seed=int(device_chain_seed), stream_id=base,
                                                 model=1, device=eng.dev)
            chunks.append((base, sums))
            return paths

        it = eng.run_device_rechain(num_iter, None, tau, window=window, chunk_source=chunk_source)
        for base, sums in chunks:
            for q, v in enumerate(sums.cpu().numpy()):
                kk = base + q
                cost_at[kk * tau_i if tau_i > 1 else kk + 1] = float(v)
    elif rechain_mode == "device":
        def chain_source(kk):
            return rechain(kk * tau_i if tau_i > 1 else kk + 1)
        it = eng.run_device_rechain(num_iter, chain_source, tau, window=window)
    else:
        it = eng.run(num_iter, window=window, rechain=rechain, coherence=tau)
    obj, loss, it, tv = eng.results(it)
    per = np.empty(it)
    c = cost_at[1]
    for i in range(1, it + 1):
        c = cost_at.get(i, c)
        per[i - 1] = c
    res = (obj, loss, it, tv, np.cumsum(per))
    return res + (eng,) if return_engine else res


def dynamic_gadmm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc, path,
                  pathCost, coherence_Time, rng=None, rechain=None, store="f64", window=32,
                  dist=None, return_engine=False, rechain_mode="device", chain_gen="auto"):
    """D-GADMM with live ``findPath2`` re-chaining (dynamic_group_ADMM_closedForm.m:18-21).
    ``chain_gen="device"`` (the default unless a ``rechain`` callback supplies the chains):
    positions come from the Philox generator (seed = ``rng``) and the greedy chains are built on
    the GPU, one CTA per chain, so a fresh chain every tau iterations costs no host round trip;
    ``chain_gen="host"``: chains are drawn with NumPy on rank 0 and broadcast (bit-identical to
    the oracle for the same ``rng``; 13 ms per chain at N = 1024)."""
    dist = dist or Dist()
    if chain_gen == "auto":
        chain_gen = "host" if rechain is not None else "device"
    if chain_gen == "device" and isinstance(rng, np.random.Generator):
        rng = int(rng.integers(2 ** 31 - 1))          # one draw seeds the device generator
    g = np.random.default_rng(rng) if not isinstance(rng, np.random.Generator) else rng

    def next_chain(i):
        if rechain is not None:
            return rechain(i)
        out = findPath2(no_workers, g) if dist.rank == 0 else None
        out = dist.broadcast_object(out)
        return out[0], out[1]

    seed = None
    if chain_gen == "device" and rechain is None:
        seed = int(rng) if isinstance(rng, (int, np.integer)) else 0
    return _dynamic(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc, path,
                    np.sum(pathCost), coherence_Time, next_chain, store, window, dist,
                    return_engine, rechain_mode, seed)


def dynamic_gadmm_v0(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                     path_matrix, pathCost_matrix, coherence_Time, faithful=True, store="f64",
                     window=32, dist=None, return_engine=False, rechain_mode="device"):
    paths = np.atleast_2d(np.asarray(path_matrix, dtype=np.int64))
    costs = np.atleast_2d(np.asarray(pathCost_matrix, dtype=np.float64))
    kk = [0]
    first_cost = costs[:, 0].sum() if faithful else costs[0].sum()   # quirk 4 (v0.m:18)

    def next_chain(i):
        kk[0] += 1
        return paths[kk[0]], costs[kk[0]]

    def next_chain_safe(i):
        if kk[0] + 1 >= paths.shape[0]:          # chunk pre-fetch may look past the supplied table
            return paths[-1], costs[-1]
        return next_chain(i)

    return _dynamic(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                    paths[0], first_cost, coherence_Time, next_chain_safe, store, window, dist,
                    return_engine, rechain_mode)


def opt_sol_closed_form(XX, YY, return_theta=False):
    """Global optimum on the GPU in fp64: normal equations on the summed Gram (K12)."""
    from ..ops import linalg as LA
    dev = torch.device("cuda", torch.cuda.current_device())
    X = torch.as_tensor(np.asarray(XX, dtype=np.float64), device=dev)
    y = torch.as_tensor(np.asarray(YY, dtype=np.float64).reshape(-1), device=dev)
    d = X.shape[1]
    G, b, hyy = LA.gram(X.unsqueeze(0), y.unsqueeze(0))
    A = LA.shifted_inverse(G, d, 0.0)
    P = LA.pack_sym(A, d)
    th = LA.BatchedSymv(P, G.shape[1])(b)[0, :d]
    r = X @ th - y
    out = 0.5 * float((r @ r).item())
    return (out, th.cpu().numpy()) if return_theta else out
