"""Chain GADMM / D-GADMM engine: device state, worker->GPU packing, launch loop.

One instance per process (= per GPU).  Logical workers are packed as contiguous id ranges onto
the ranks (:func:`gadmm_b200.parallel.runtime.partition`); every rank holds

  * the packed inverses ``(G_n + c rho I)^-1`` of its own workers (the d^2 state never moves),
  * a full replica of the theta table ``[2][N][dp]`` in symmetric memory -- own rows are written
    by the local finisher CTAs, rows of remote chain neighbours arrive as NVLink P2P stores from
    the owning GPU's kernel, guarded by per-worker version flags,
  * node duals, chain tables, per-iteration objective / loss / clock rings.

An iteration is three launches (head half-iteration, tail half-iteration, dual+objective), all
iteration-agnostic (the iteration counter lives on the device), so a window of iterations is
captured once in a CUDA graph and replayed.  Reference semantics: group_ADMM_closedForm.m:11-108,
static_group_ADMM_closedForm.m:19-160, dynamic_group_ADMM_closedForm.m:14-190.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np
import torch

from ..ops import linalg as LA
from ..ops import native as nv
from ..parallel.runtime import Dist, partition
from ..topology import chain_tables


def _carve(buf: torch.Tensor, off: int, nbytes: int, dtype) -> torch.Tensor:
    return buf[off:off + nbytes].view(dtype)


class ChainEngine:
    def __init__(self, N: int, d: int, rho: float, obj0: float, acc: float, max_iter: int,
                 dist: Optional[Dist] = None, store: str = "f64", dynamic: bool = False,
                 stop: bool = True, pdl: bool = True):
        self.dist = dist or Dist()
        self.dev = self.dist.device
        self.N, self.d, self.rho = int(N), int(d), float(rho)
        self.dp = nv.padded(self.d)
        self.obj0, self.acc, self.max_iter = float(obj0), float(acc), int(max_iter)
        # store="mixed": iterate on fp32-stored inverses (half the bytes per iteration) until the
        # estimated gap is small, then switch to the fp64 inverses, whose objective / stop test are
        # exact (see switch_to_f64); both sets are kept.
        self.mixed = store == "mixed"
        self.store = nv.STORE_F64 if store in ("f64", "fp64", nv.STORE_F64) else nv.STORE_F32
        if store == "f32" and stop:
            import warnings
            warnings.warn("store='f32': the reported objective uses G theta = r - c rho theta, which "
                          "is only first-order accurate with fp32 inverses (error ~ cond(A) x 6e-8 "
                          "relative); the stop test is unreliable -- use store='mixed' or 'f64'",
                          stacklevel=3)
        self.dynamic = dynamic
        self.stop = stop
        self.pdl = bool(pdl) and os.environ.get("GADMM_B200_NO_PDL") != "1"
        self.owner_np = partition(self.N, self.dist.world)
        self.gids = np.nonzero(self.owner_np == self.dist.rank)[0].astype(np.int32)
        self.Nloc = int(self.gids.shape[0])
        if self.Nloc == 0:
            raise ValueError("every rank must own at least one worker (N >= world size)")
        self._alloc()
        self._graphs = {}
        self.launches = 0
        self.path0 = None

    # ------------------------------------------------------------------------------ memory
    def _alloc(self):
        N, dp, Nloc, W = self.N, self.dp, self.Nloc, self.dist.world
        dev = self.dev
        f64, i32, i64 = torch.float64, torch.int32, torch.int64
        # symmetric region: theta[2][N][dp] | flags[N] | objpart[4][W] | objflag[4][W]
        sz_theta = 2 * N * dp * 8
        sz_flags = (N * 8 + 255) // 256 * 256
        sz_op = (4 * W * 8 + 255) // 256 * 256
        self._sym = self.dist.symm_bytes(sz_theta + sz_flags + 2 * sz_op)
        buf = self._sym.local
        self.theta = _carve(buf, 0, sz_theta, f64).view(2, N, dp)
        self.flags = _carve(buf, sz_theta, N * 8, i64)
        self.objpart = _carve(buf, sz_theta + sz_flags, 4 * W * 8, f64)
        self.objflag = _carve(buf, sz_theta + sz_flags + sz_op, 4 * W * 8, i64)
        peers = self._sym.peer_ptrs

        def table(off):
            return torch.tensor([p + off for p in peers], dtype=i64, device=dev)

        self.theta_peer = table(0)
        self.flags_peer = table(sz_theta)
        self.objpart_peer = table(sz_theta + sz_flags)
        self.objflag_peer = table(sz_theta + sz_flags + sz_op)

        self.b = torch.zeros((Nloc, dp), dtype=f64, device=dev)
        self.hyy = torch.zeros((Nloc,), dtype=f64, device=dev)
        self.nu = torch.zeros((Nloc, dp), dtype=f64, device=dev)
        cpe = max(1, (dp // 64) * (dp // 64 - 1) // 2 * 64)
        # few workers per GPU: divide every tile-row pair over several CTAs so that ~2 CTAs per SM
        # are streaming (a single CTA sustains ~25 GB/s)
        active = max(1, (Nloc * ((dp // 64 + 1) // 2) + 1) // 2)
        want = max(1, -(-296 // active))
        self.nsplit = int(min(8, 1 << (want - 1).bit_length())) if dp // 64 >= 8 else 1
        if os.environ.get("GADMM_B200_NSPLIT"):
            self.nsplit = int(os.environ["GADMM_B200_NSPLIT"])
        self.rowpart = torch.empty((Nloc, self.nsplit, dp), dtype=f64, device=dev)
        self.colpart = torch.empty((Nloc, cpe), dtype=f64, device=dev)
        self.counters = torch.zeros((Nloc + 2,), dtype=i32, device=dev)
        self.objw = torch.zeros((2, Nloc), dtype=f64, device=dev)
        self.sync = torch.zeros((2,), dtype=i64, device=dev)       # window base, decided-through
        self.nuflag = torch.zeros((Nloc,), dtype=i64, device=dev)
        # parallel finisher (csrc/gadmm_iter.cu: gadmm_finish_kernel) for d >= 512
        self.split_finish = (dp // 64 >= 8) and os.environ.get("GADMM_B200_NO_SPLIT") != "1"
        self.wready = torch.zeros((Nloc,), dtype=i64, device=dev)
        self.sblk = torch.zeros((Nloc, dp // 64, 2), dtype=f64, device=dev)
        self.fin_counters = torch.zeros((Nloc,), dtype=i32, device=dev)
        self.ctrl = torch.zeros((4,), dtype=i64, device=dev)
        self.obj_hist = torch.zeros((self.max_iter + 2,), dtype=f64, device=dev)
        self.loss_hist = torch.zeros((self.max_iter + 2,), dtype=f64, device=dev)
        self.t_hist = torch.zeros((self.max_iter + 2,), dtype=i64, device=dev)
        # chain tables: one int32 block [7][Nloc] so a re-chain is a single H2D copy
        self.tab = torch.zeros((7, Nloc), dtype=i32, device=dev)
        self.tab_host = torch.zeros((7, Nloc), dtype=i32).pin_memory()
        self.tab_host[0] = torch.from_numpy(self.gids)
        self.owner = torch.from_numpy(self.owner_np).to(dev)
        self.order = torch.arange(Nloc, dtype=i32, device=dev)
        self.Mptr = torch.zeros((Nloc,), dtype=i64, device=dev)
        self.Mptr_host = torch.zeros((Nloc,), dtype=i64).pin_memory()
        pe = nv.packed_elems(dp)
        dt = f64 if self.store == nv.STORE_F64 else torch.float32
        # static chains need one inverse per worker; dynamic chains keep c=1 and c=2 (a worker
        # becomes / ceases to be a chain end when the chain is re-drawn)
        self.P = {2: torch.empty((Nloc, pe), dtype=dt, device=dev)}
        if self.dynamic:
            self.P[1] = torch.empty((Nloc, pe), dtype=dt, device=dev)
        else:
            self.P_end = {}   # local index -> (1, pe) tensor for chain-end workers (c=1)
        if self.mixed:        # the fp64 twins, used after switch_to_f64()
            self.P64 = {c: torch.empty((Nloc, pe), dtype=f64, device=dev) for c in self.P}
            self.P64_end = {}
            self.P32, self.P32_end = self.P, getattr(self, "P_end", None)
        self.ctrl[0] = 1
        self.ctrl[2] = 1 if (self.stop and not self.mixed) else 0   # mixed: stop test after the switch
        self.ctrl[3] = -1          # first-crossing register of the persistent kernel (= ~0)
        self.objring = torch.zeros((32 * Nloc,), dtype=f64, device=dev)
        self.ringcnt = torch.zeros((32,), dtype=i32, device=dev)

    # ------------------------------------------------------------------------------- setup
    def ingest_host(self, XX, YY, s: int, batch: int = 0):
        """Upload the local workers' shards, build Gram / RHS / inverses on the GPU."""
        XX = np.asarray(XX, dtype=np.float64)
        YY = np.asarray(YY, dtype=np.float64).reshape(-1)
        d, Nloc = self.d, self.Nloc
        if batch <= 0:
            per = self.dp * self.dp * 8 * 4 + s * d * 8
            batch = max(1, min(Nloc, int(8e9 // per)))
        for k0 in range(0, Nloc, batch):
            k1 = min(Nloc, k0 + batch)
            g0, g1 = int(self.gids[k0]), int(self.gids[k1 - 1]) + 1
            Xb = torch.from_numpy(XX[g0 * s:g1 * s, :d].reshape(g1 - g0, s, d)).to(self.dev)
            yb = torch.from_numpy(YY[g0 * s:g1 * s].reshape(g1 - g0, s)).to(self.dev)
            self.ingest_device(Xb, yb, k0)

    def ingest_device(self, Xb: torch.Tensor, yb: torch.Tensor, k0: int):
        """Xb: (nb, s, d) fp64 on this device for local workers k0..k0+nb."""
        nb = Xb.shape[0]
        G, b, hyy = LA.gram(Xb, yb, self.dp)
        self.ingest_gram(G, b, hyy, k0)

    def ingest_gram(self, G: torch.Tensor, b: torch.Tensor, hyy: torch.Tensor, k0: int):
        nb = G.shape[0]
        self.b[k0:k0 + nb] = b
        self.hyy[k0:k0 + nb] = hyy
        A = LA.shifted_inverse(G, self.d, 2.0 * self.rho)
        LA.pack_sym(A, self.d, self.store, out=self.P[2][k0:k0 + nb])
        if self.mixed:
            LA.pack_sym(A, self.d, nv.STORE_F64, out=self.P64[2][k0:k0 + nb])
        if self.dynamic:
            A = LA.shifted_inverse(G, self.d, 1.0 * self.rho, out=A)
            LA.pack_sym(A, self.d, self.store, out=self.P[1][k0:k0 + nb])
            if self.mixed:
                LA.pack_sym(A, self.d, nv.STORE_F64, out=self.P64[1][k0:k0 + nb])
        else:
            self._G_keep = getattr(self, "_G_keep", {})
            # keep the Gram of potential chain ends (first / last worker id) for the c=1 inverse
            for g_end in (0, self.N - 1):
                k = int(np.searchsorted(self.gids, g_end))
                if k < self.Nloc and self.gids[k] == g_end and k0 <= k < k0 + nb:
                    self._G_keep[k] = G[k - k0:k - k0 + 1].clone()

    def set_theta(self, theta0):
        """Optional non-zero start (BASELINE.json: 'random-init theta'); default zeros = reference."""
        t = torch.as_tensor(np.asarray(theta0, dtype=np.float64).T).to(self.dev)   # (N, d)
        self.theta[:, :, :self.d] = t.unsqueeze(0)

    # ------------------------------------------------------------------------------ topology
    def set_path(self, path0, sync_tables: bool = True):
        """Install a chain (0-based permutation).  Stream-ordered: safe between iterations."""
        path0 = np.asarray(path0, dtype=np.int64)
        left, right, color, cmul = chain_tables(path0)
        g = self.gids
        own = self.owner_np
        rank = self.dist.rank
        push = np.full((2, self.Nloc), -1, dtype=np.int32)
        for k, gid in enumerate(g):
            tg = []
            for nb in (left[gid], right[gid]):
                if nb >= 0 and own[nb] != rank and own[nb] not in tg:
                    tg.append(int(own[nb]))
            for i, r in enumerate(tg):
                push[i, k] = r
        th = self.tab_host
        th[1] = torch.from_numpy(left[g].astype(np.int32))
        th[2] = torch.from_numpy(right[g].astype(np.int32))
        th[3] = torch.from_numpy(color[g].astype(np.int32))
        th[4] = torch.from_numpy(cmul[g].astype(np.int32))
        th[5] = torch.from_numpy(push[0])
        th[6] = torch.from_numpy(push[1])
        pe_bytes = self.P[2].stride(0) * self.P[2].element_size()
        base2 = self.P[2].data_ptr()
        mp = self.Mptr_host
        for k, gid in enumerate(g):
            c = int(cmul[gid])
            if c == 2:
                mp[k] = base2 + k * pe_bytes
            elif self.dynamic:
                mp[k] = self.P[1].data_ptr() + k * pe_bytes
            else:
                if k not in self.P_end:
                    if k not in getattr(self, "_G_keep", {}):
                        raise ValueError(
                            "a static engine (dynamic=False) only supports chains that start and "
                            f"end at workers 0 and N-1; worker {int(gid)} is a chain end of this "
                            "chain -- construct the engine with dynamic=True")
                    A = LA.shifted_inverse(self._G_keep[k], self.d, 1.0 * self.rho)
                    self.P_end[k] = LA.pack_sym(A, self.d, self.store)
                    if self.mixed:
                        self.P64_end[k] = LA.pack_sym(A, self.d, nv.STORE_F64)
                mp[k] = self.P_end[k].data_ptr()
        self.tab.copy_(th, non_blocking=True)
        self.Mptr.copy_(mp, non_blocking=True)
        # launch order: workers that exchange theta with another GPU go first in every phase
        remote = (push[0] >= 0) | (push[1] >= 0)
        order = np.concatenate([np.nonzero(remote)[0], np.nonzero(~remote)[0]]).astype(np.int32)
        self.order.copy_(torch.from_numpy(order))
        prev = self.path0
        self.path0 = path0
        self.left, self.right = left, right
        if self.dist.world > 1 and prev is not None:
            self._gather_new_neighbours()

    def _gather_new_neighbours(self):
        """After a re-chain: pull theta (latest version) of remote neighbours that were not
        neighbours before.  All ranks are iteration-aligned here (the per-iteration objective
        exchange bounds the skew), so the owners' rows are final."""
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()
        it = int(self.ctrl[0].item())           # iteration about to run
        ver = it - 1
        slot = ver & 1
        rank = self.dist.rank
        need = set()
        for gid in self.gids:
            for nb in (self.left[gid], self.right[gid]):
                if nb >= 0 and self.owner_np[nb] != rank:
                    need.add(int(nb))
        row_bytes = self.dp * 8
        for nb in sorted(need):
            src = self._sym.peer_ptrs[int(self.owner_np[nb])] + (slot * self.N + nb) * row_bytes
            dst = self.theta[slot, nb]
            nv_cudart_memcpy(dst.data_ptr(), src, row_bytes)
            self.flags[nb] = max(int(self.flags[nb].item()), ver)
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()

    def switch_to_f64(self):
        """``store="mixed"``: from the next iteration on use the fp64 inverses.  Models and duals
        carry over (any (theta, nu) is a valid GADMM state); from here the objective identity is
        exact and the stop test is enabled.  Stream-ordered; captured graphs are dropped."""
        if not self.mixed or self.store == nv.STORE_F64:
            return
        self.store = nv.STORE_F64
        self.P = self.P64
        if not self.dynamic:
            self.P_end = self.P64_end
        self._desc = None
        self._graphs = {}
        self.ctrl[2] = 1 if self.stop else 0
        if self.path0 is not None:
            prev, self.path0 = self.path0, None      # (no neighbour gather: same chain)
            self.set_path(prev)

    def _mixed_ready(self, switch_tol: float) -> bool:
        """True when the (estimated) gap of the last completed iteration is below ``switch_tol``."""
        it = int(self.ctrl[0].item()) - 1
        return it >= 1 and float(self.loss_hist[it].item()) < switch_tol

    # ---------------------------------------------------------------------------- descriptor
    def desc(self) -> nv.GadmmChain:
        if getattr(self, "_desc", None) is not None:
            return self._desc
        c = nv.GadmmChain()
        t = self.tab
        for i, name in enumerate(("gid", "left", "right", "color", "cmul", "push0", "push1")):
            setattr(c, name, t[i].data_ptr())
        c.Mptr = self.Mptr.data_ptr()
        c.owner = self.owner.data_ptr()
        c.b, c.hyy, c.nu = self.b.data_ptr(), self.hyy.data_ptr(), self.nu.data_ptr()
        c.theta, c.flags = self.theta.data_ptr(), self.flags.data_ptr()
        c.theta_peer, c.flags_peer = self.theta_peer.data_ptr(), self.flags_peer.data_ptr()
        c.objpart, c.objflag = self.objpart.data_ptr(), self.objflag.data_ptr()
        c.objpart_peer, c.objflag_peer = self.objpart_peer.data_ptr(), self.objflag_peer.data_ptr()
        c.rowpart, c.colpart = self.rowpart.data_ptr(), self.colpart.data_ptr()
        c.counters, c.objw, c.ctrl = self.counters.data_ptr(), self.objw.data_ptr(), self.ctrl.data_ptr()
        c.obj_hist, c.loss_hist, c.t_hist = (self.obj_hist.data_ptr(), self.loss_hist.data_ptr(),
                                             self.t_hist.data_ptr())
        c.rho, c.obj0, c.acc = self.rho, self.obj0, self.acc
        c.d, c.dp, c.Nglob, c.Nloc = self.d, self.dp, self.N, self.Nloc
        c.rank, c.nranks, c.store, c.max_iter = self.dist.rank, self.dist.world, self.store, self.max_iter
        c.pdl = int(self.pdl)
        c.split_finish = int(self.split_finish)
        c.nsplit = int(self.nsplit)
        c.order = self.order.data_ptr()
        c.sync, c.nuflag = self.sync.data_ptr(), self.nuflag.data_ptr()
        c.wready, c.sblk, c.fin_counters = (self.wready.data_ptr(), self.sblk.data_ptr(),
                                            self.fin_counters.data_ptr())
        self._desc = c
        return c

    # --------------------------------------------------------------------------------- run
    def iterate(self, n_iter: int, use_graph: bool = True):
        """Enqueue ``n_iter`` iterations on the current stream (no host sync)."""
        L = nv.lib()
        c = self.desc()
        if not use_graph or n_iter < 4:
            nv.check(L.gadmm_chain_iterate(C.byref(c), int(n_iter), nv.stream_ptr()))
        else:
            self._capture(n_iter).replay()
        self.launches += (5 if self.split_finish else 3) * int(n_iter) + 1   # + window begin

    def _capture(self, n_iter: int):
        g = self._graphs.get(n_iter)
        if g is None:
            nv.lib().gadmm_chain_configure()
            c = self.desc()
            torch.cuda.synchronize(self.dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                nv.check(nv.lib().gadmm_chain_iterate(C.byref(c), int(n_iter), nv.stream_ptr()))
            self._graphs[n_iter] = g        # capture does not execute anything
        return g

    def _switch_tol(self, switch_tol):
        if switch_tol is not None:
            return float(switch_tol)
        return max(1e3 * self.acc, 1e-5 * abs(self.obj0))

    def run(self, num_iter: Optional[int] = None, window: int = 32, rechain=None,
            coherence: Optional[float] = None, use_graph: bool = True,
            switch_tol: Optional[float] = None) -> int:
        """Run until the stop test fires or ``num_iter`` iterations are done.  ``rechain(i)``
        returns the 0-based chain installed before iteration ``i`` when ``i % coherence == 0``
        (dynamic_group_ADMM_closedForm.m:18-21).  ``store="mixed"``: fp32 inverses until the
        estimated gap is below ``switch_tol`` (default max(1e3 acc, 1e-5 |obj0|), checked once per
        window), fp64 inverses and the exact stop test from there.  Returns ``Iter``."""
        num_iter = self.max_iter if num_iter is None else min(int(num_iter), self.max_iter)
        nv.lib().gadmm_chain_configure()
        done = getattr(self, "_resume_from", 0)
        if rechain is None and self.persistent_ok():
            return self._run_persistent(num_iter, done)
        if use_graph and rechain is None and num_iter - done >= max(4, window):
            self._capture(window)              # capture before the device clock of iteration 1
        while done < num_iter:
            it_next = done + 1
            if rechain is not None and coherence is not None and it_next > 1 \
                    and it_next % coherence == 0:
                self.set_path(rechain(it_next))
            n = min(window, num_iter - done)
            if rechain is not None and coherence is not None and coherence < 1e8:
                tau = int(coherence)
                nxt = (it_next // tau + 1) * tau      # next re-chain iteration
                n = max(1, min(n, nxt - it_next))
            self.iterate(n, use_graph=use_graph)
            done += n
            if self.mixed and self.store == nv.STORE_F32 and \
                    self._mixed_ready(self._switch_tol(switch_tol)):
                self.switch_to_f64()
                self.switched_at = done
            if self.stop:
                hit = int(self.ctrl[1].item())      # host sync once per window
                if hit:
                    return hit
        return num_iter

    # ------------------------------------------------------- D-GADMM with device-side re-chaining
    def run_device_rechain(self, num_iter, chain_source, tau, window: int = 32, chunk: int = 256,
                           chunk_source=None, switch_tol: Optional[float] = None):
        """D-GADMM without the host in the loop: the chains of the coming windows are uploaded in
        chunks of ``chunk`` rows; two tiny kernels in front of every head kernel switch tables and
        pull the new remote neighbours' theta over NVLink when ``it % tau == 0`` (csrc/rechain.cu).
        ``chain_source(kk)`` returns the 0-based chain of window ``kk`` (kk = 1, 2, ...; called in
        increasing order).  ``chunk_source(base, K)`` (optional) returns a device int32 tensor
        (K, N) with the chains of windows base..base+K-1 -- e.g. from
        :func:`gadmm_b200.ops.topology.find_path_batch_device`, which keeps chain generation on
        the GPU as well.  Returns ``Iter``."""
        assert self.dynamic, "device re-chaining needs both inverses (dynamic=True)"
        num_iter = min(int(num_iter), self.max_iter)
        L = nv.lib()
        L.gadmm_chain_configure()
        N, dev = self.N, self.dev
        tau_i = 0 if tau >= 1e8 else int(tau)
        if tau_i > 1:
            # windows that are whole multiples of tau start at the same offset from a re-chain
            # point every time, so ONE captured graph serves the whole run (a new graph per window
            # offset would be captured inside the run, with the GPU idle meanwhile)
            window = max(tau_i, (int(window) // tau_i) * tau_i)
        K = int(chunk)
        paths_dev = torch.zeros((K, N), dtype=torch.int32, device=dev)
        paths_host = torch.zeros((K, N), dtype=torch.int32).pin_memory()
        pos = torch.zeros((N,), dtype=torch.int32, device=dev)
        dyn = torch.zeros((2,), dtype=torch.int64, device=dev)
        dyn_host = torch.zeros((2,), dtype=torch.int64).pin_memory()
        r = nv.GadmmRechain()
        r.paths, r.pos, r.tab, r.Mptr = paths_dev.data_ptr(), pos.data_ptr(), self.tab.data_ptr(), self.Mptr.data_ptr()
        r.owner, r.theta, r.flags = self.owner.data_ptr(), self.theta.data_ptr(), self.flags.data_ptr()
        r.theta_peer, r.ctrl, r.dyn = self.theta_peer.data_ptr(), self.ctrl.data_ptr(), dyn.data_ptr()
        r.base_c1, r.base_c2 = self.P[1].data_ptr(), self.P[2].data_ptr()
        r.stride_bytes = self.P[2].stride(0) * self.P[2].element_size()
        r.tau, r.N, r.Nloc, r.dp = tau_i, N, self.Nloc, self.dp
        r.rank, r.nranks, r.K = self.dist.rank, self.dist.world, K
        r.order = 0 if os.environ.get("GADMM_B200_NO_REORDER") == "1" else self.order.data_ptr()
        c = self.desc()
        graphs = {}

        def launch(it0, n):
            # iterations it0 .. it0+n-1: the re-chain kernels run only in front of an iteration that
            # can re-chain (it % tau == 0); the iterations up to the next such point form one
            # window of gadmm_chain_iterate, i.e. one programmatic-dependent-launch chain
            i, end = it0, it0 + n
            while i < end:
                seg = 1 if tau_i <= 1 else min(end - i, tau_i - (i % tau_i))
                if tau_i and i % tau_i == 0:        # (a no-op on the device for iteration 1)
                    nv.check(L.gadmm_rechain(C.byref(r), nv.stream_ptr()))
                    self.launches += 2
                nv.check(L.gadmm_chain_iterate(C.byref(c), int(seg), nv.stream_ptr()))
                self.launches += (5 if self.split_finish else 3) * seg + 1
                i += seg

        def iterate(it0, n):
            key = (it0 % tau_i if tau_i > 1 else 0, n)
            if n >= 4:
                g = graphs.get(key)
                if g is None:
                    l0 = self.launches
                    torch.cuda.synchronize(dev)
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        launch(it0, n)
                    graphs[key] = (g, self.launches - l0)
                    self.launches = l0
                g, cnt = graphs[key]
                g.replay()
                self.launches += cnt
            else:
                launch(it0, n)

        done, base = 0, 1
        while done < num_iter:
            if tau_i:
                torch.cuda.current_stream().synchronize()       # previous chunk fully consumed
                dyn_host[0], dyn_host[1] = base, K
                if chunk_source is not None:
                    paths_dev.copy_(chunk_source(base, K))
                else:
                    for q in range(K):
                        paths_host[q] = torch.from_numpy(np.asarray(chain_source(base + q),
                                                                    dtype=np.int32))
                    paths_dev.copy_(paths_host, non_blocking=True)
                dyn.copy_(dyn_host, non_blocking=True)
                it_star = (base + K) * tau_i if tau_i > 1 else base + K + 1
                target = min(num_iter, it_star - 1)
            else:
                target = num_iter
            while done < target:
                n = min(window, target - done)
                iterate(done + 1, n)
                done += n
                if self.mixed and self.store == nv.STORE_F32 and \
                        self._mixed_ready(self._switch_tol(switch_tol)):
                    torch.cuda.synchronize(dev)
                    # same chain tables (they live on the device), fp64 inverses from here on
                    self.store, self.P = nv.STORE_F64, self.P64
                    self._desc = None
                    self.ctrl[2] = 1 if self.stop else 0
                    r.base_c1, r.base_c2 = self.P[1].data_ptr(), self.P[2].data_ptr()
                    r.stride_bytes = self.P[2].stride(0) * self.P[2].element_size()
                    # pointer table: same workers, same c, other base
                    off = self.Mptr - self.P32[2].data_ptr()
                    is1 = (off < 0) | (off >= self.P32[2].numel() * 4)
                    k = torch.arange(self.Nloc, device=dev)
                    self.Mptr.copy_(torch.where(is1, self.P[1].data_ptr() + k * r.stride_bytes,
                                                self.P[2].data_ptr() + k * r.stride_bytes))
                    c = self.desc()
                    graphs.clear()
                    self.switched_at = done
                if self.stop:
                    hit = int(self.ctrl[1].item())
                    if hit:
                        self._keep = (paths_dev, pos, dyn)
                        return hit
            base += K
        self._keep = (paths_dev, pos, dyn)
        return num_iter

    # ------------------------------------------------------------------- streaming public API
    def step(self, path0=None):
        """One GADMM iteration driven from the host, end to end: the step's control block (chain
        tables + inverse pointer table; a new chain if ``path0`` is given) is copied host->device
        from pinned memory, the three kernels run, and ``(objective, loss)`` of the iteration is
        read back device->host.  This is the user-facing online mode (e.g. D-GADMM where an
        external radio layer supplies the chain every step).  Returns ``(obj, loss)``."""
        if not hasattr(self, "_step_out"):
            self._step_out = torch.zeros((2,), dtype=torch.float64).pin_memory()
            nv.lib().gadmm_chain_configure()
            self._it_host = int(self.ctrl[0].item())
        if path0 is not None:
            self.set_path(path0)
        else:
            self.tab.copy_(self.tab_host, non_blocking=True)
            self.Mptr.copy_(self.Mptr_host, non_blocking=True)
        self.iterate(1, use_graph=False)
        it = self._it_host
        self._step_out[0:1].copy_(self.obj_hist[it:it + 1], non_blocking=True)
        self._step_out[1:2].copy_(self.loss_hist[it:it + 1], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self._it_host += 1
        return float(self._step_out[0]), float(self._step_out[1])

    @property
    def step_h2d_bytes(self) -> int:
        return self.tab_host.numel() * 4 + self.Mptr_host.numel() * 8

    step_d2h_bytes = 16

    # ------------------------------------------------------ persistent kernel (small problems)
    def persistent_ok(self) -> bool:
        if os.environ.get("GADMM_B200_NO_PERSISTENT") == "1" or self.path0 is None:
            return False
        return nv.lib().gadmm_chain_persistent_supported(C.byref(self.desc())) == 0

    def _run_persistent(self, num_iter: int, done: int, window: int = 4096) -> int:
        """Reference-size problems: the whole solve is one cooperative launch per ``window``
        iterations (csrc/gadmm_persistent.cu); state stays in shared memory."""
        L = nv.lib()
        c = self.desc()
        while done < num_iter:
            n = min(window, num_iter - done)
            nv.check(L.gadmm_chain_persistent(C.byref(c), int(n), nv.ptr(self.objring),
                                              nv.ptr(self.ringcnt), nv.stream_ptr()))
            self.launches += 1
            done += n
            if self.stop:
                hit = int(self.ctrl[1].item())
                if hit:
                    return hit
        torch.cuda.synchronize(self.dev)
        return num_iter

    # ----------------------------------------------------------------------------- results
    def one_worker_solve_seconds(self, reps: int = 21) -> float:
        """Device time of ONE worker's solve ``theta = M r`` (median of ``reps`` launches of the
        streaming SYMV on the first local worker's matrix) -- the quantity the reference puts
        between ``tic`` and ``toc`` (group_ADMM_closedForm.m:40-42,53-55)."""
        ptrs = self.Mptr[:1].clone()
        x = torch.ones((1, self.dp), dtype=torch.float64, device=self.dev)
        y = torch.empty_like(x)
        scratch = torch.empty((nv.symv_scratch_elems(self.dp),), dtype=torch.float64, device=self.dev)
        cnt = torch.zeros((1,), dtype=torch.int32, device=self.dev)
        times = []
        for k in range(reps + 3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            nv.check(nv.lib().gadmm_symv_batched(nv.ptr(ptrs), nv.ptr(x), nv.ptr(y), 1, self.dp,
                                                 self.store, nv.ptr(scratch), nv.ptr(cnt),
                                                 nv.stream_ptr()))
            e1.record()
            e1.synchronize()
            if k >= 3:
                times.append(e0.elapsed_time(e1) * 1e-3)
        return float(np.median(times))

    def results(self, Iter: int, clock: str = "device"):
        """``(obj, loss, Iter, time)``.  ``clock="device"``: device time since iteration 1, max
        over ranks.  ``clock="simulated"``: the reference's model of a parallel deployment,
        ``time(i) = time(i-1) + 2 * (one worker's solve)`` with iteration 1 untimed (SURVEY.md
        section 5.1); the solve is timed on this GPU."""
        torch.cuda.synchronize(self.dev)
        if clock == "simulated":
            obj = self.obj_hist[1:Iter + 1].cpu().numpy().copy()
            loss = self.loss_hist[1:Iter + 1].cpu().numpy().copy()
            dt = self.dist.all_reduce_max(self.one_worker_solve_seconds())
            return obj, loss, Iter, 2.0 * dt * np.arange(Iter, dtype=np.float64)
        if clock != "device":
            raise ValueError("clock must be 'device' or 'simulated'")
        obj = self.obj_hist[1:Iter + 1].cpu().numpy().copy()
        loss = self.loss_hist[1:Iter + 1].cpu().numpy().copy()
        t = self.t_hist[1:Iter + 1].to(torch.float64)
        if Iter > 0:
            live = t > 0                       # resumed runs carry no clock for restored iterations
            t0 = t[live][0] if bool(live.any()) else t[0]
            tv = torch.clamp((t - t0) * 1e-9, min=0.0)
        else:
            tv = t
        if self.dist.world > 1:
            import torch.distributed as dist
            dist.all_reduce(tv, op=dist.ReduceOp.MAX)
        return obj, loss, Iter, tv.cpu().numpy()

    def state(self):
        """Checkpointable solver state: theta and node duals of *all*
        workers (d x N, gathered over the ranks), iteration, chain, curves so far.  Identical on
        every rank; ``np.savez`` / ``scipy.io.savemat`` friendly."""
        torch.cuda.synchronize(self.dev)
        it = int(self.ctrl[0].item()) - 1
        slot = it & 1
        idx = torch.from_numpy(self.gids).long().to(self.dev)
        th = torch.zeros((self.N, self.d), dtype=torch.float64, device=self.dev)
        nu = torch.zeros((self.N, self.d), dtype=torch.float64, device=self.dev)
        th[idx] = self.theta[slot, idx, :self.d]
        nu[idx] = self.nu[:, :self.d]
        self.dist.all_reduce_sum_(th)
        self.dist.all_reduce_sum_(nu)
        return {"theta": th.cpu().numpy().T.copy(), "dual": nu.cpu().numpy().T.copy(),
                "iteration": it, "path": None if self.path0 is None else self.path0 + 1,
                "obj": self.obj_hist[1:it + 1].cpu().numpy().copy(),
                "loss": self.loss_hist[1:it + 1].cpu().numpy().copy(),
                "rho": self.rho}

    def reset(self, stop: Optional[bool] = None, acc: Optional[float] = None,
              obj0: Optional[float] = None):
        """Back to iteration 0 (zero models and duals) on the same data, inverses and chain, e.g.
        to measure time-to-accuracy after a throughput run.  Curves and clocks are cleared; the
        stop test / tolerance / optimum can be changed (captured graphs are dropped with them)."""
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()
        if stop is not None:
            self.stop = bool(stop)
        if acc is not None:
            self.acc = float(acc)
        if obj0 is not None:
            self.obj0 = float(obj0)
        for t in (self.theta, self.flags, self.objpart, self.objflag, self.nu, self.counters,
                  self.fin_counters, self.wready, self.objw, self.obj_hist, self.loss_hist,
                  self.t_hist, self.ringcnt, self.nuflag, self.sync):
            t.zero_()
        self.switched_at = None
        if self.mixed:                      # back to the fp32 phase
            self.store, self.P = nv.STORE_F32, self.P32
            if not self.dynamic:
                self.P_end = self.P32_end
            if self.path0 is not None:
                prev, self.path0 = self.path0, None
                self.set_path(prev)
        self.ctrl.zero_()
        self.ctrl[0] = 1
        self.ctrl[2] = 1 if (self.stop and not self.mixed) else 0
        self.ctrl[3] = -1
        self._resume_from = 0
        self._desc = None
        self._graphs = {}
        for name in ("_step_out", "_it_host"):
            if hasattr(self, name):
                delattr(self, name)
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()

    def load_state(self, st):
        """Resume from :meth:`state` (or from an oracle ``return_state`` dict of a node-dual run):
        the next iteration to run is ``st['iteration'] + 1``."""
        it = int(st["iteration"])
        th = torch.as_tensor(np.asarray(st["theta"], dtype=np.float64).T.copy(), device=self.dev)
        nu = torch.as_tensor(np.asarray(st["dual"], dtype=np.float64).T.copy(), device=self.dev)
        idx = torch.from_numpy(self.gids).long().to(self.dev)
        self.theta.zero_()
        self.theta[:, :, :self.d] = th.unsqueeze(0)
        self.nu.zero_()
        self.nu[:, :self.d] = nu[idx]
        self.flags.fill_(it)
        self.objflag.zero_()
        n = min(it, len(st.get("obj", [])))
        if n:
            self.obj_hist[1:n + 1] = torch.as_tensor(np.asarray(st["obj"][:n]), device=self.dev)
            self.loss_hist[1:n + 1] = torch.as_tensor(np.asarray(st["loss"][:n]), device=self.dev)
        self.ctrl[0] = it + 1
        self.ctrl[1] = 0
        self.ctrl[3] = -1
        self._resume_from = it
        if st.get("path") is not None and self.path0 is None:
            self.set_path(np.asarray(st["path"], dtype=np.int64) - 1)
        self.dist.barrier()


def nv_cudart_memcpy(dst_ptr: int, src_ptr: int, nbytes: int):
    """Peer-to-peer cudaMemcpy through the runtime already loaded by torch (UVA handles peers)."""
    import ctypes
    rt = _cudart()
    err = rt.cudaMemcpy(ctypes.c_void_p(dst_ptr), ctypes.c_void_p(src_ptr), ctypes.c_size_t(nbytes),
                        ctypes.c_int(4))   # cudaMemcpyDefault
    if err != 0:
        raise RuntimeError(f"cudaMemcpy (peer gather) failed with {err}")


_rt = None


def _cudart():
    global _rt
    if _rt is None:
        import ctypes
        import glob
        import os
        cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia",
                                       "cuda_runtime", "lib", "libcudart.so.12"))
        _rt = ctypes.CDLL(cands[0] if cands else "libcudart.so.12")
    return _rt
