"""Star (parameter-server) ADMM engine -- ``standared_ADMM`` on the GPU (K6).

The N-1 workers are split into contiguous id ranges over the ranks; the server shard (worker N)
is replicated on every rank so that the z-solve needs no broadcast.  Per iteration: one workers
kernel (solves + fused dual update + in-kernel reduction pushed to all GPUs over NVLink) and one
server kernel.  See ``csrc/star.cu``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..ops import linalg as LA
from ..ops import native as nv
from ..parallel.runtime import Dist, partition


class StarEngine:
    def __init__(self, XX, YY, rho, N, d, s, max_iter, obj0, acc, dist=None, store="f64", stop=True,
                 nvls="auto"):
        self.dist = dist or Dist()
        D = self.dist
        dev = self.dev = D.device
        self.N, self.d, self.dp, self.rho = int(N), int(d), nv.padded(int(d)), float(rho)
        self.store = nv.STORE_F64 if store in ("f64", nv.STORE_F64) else nv.STORE_F32
        self.max_iter = int(max_iter)
        if N < 2:
            raise ValueError("star ADMM needs at least one worker besides the server")
        owner = partition(N - 1, D.world)
        self.gids = np.nonzero(owner == D.rank)[0]
        Nw = self.Nw = len(self.gids)
        if Nw == 0:
            raise ValueError("every rank must own at least one worker")
        dp, f64 = self.dp, torch.float64
        pe = nv.packed_elems(dp)
        pdt = f64 if self.store == nv.STORE_F64 else torch.float32
        self.P = torch.empty((Nw, pe), dtype=pdt, device=dev)
        self.Ps = torch.empty((1, pe), dtype=pdt, device=dev)
        self.b = torch.zeros((Nw, dp), dtype=f64, device=dev)
        self.hyy = torch.zeros((Nw,), dtype=f64, device=dev)
        self.bs = torch.zeros((dp,), dtype=f64, device=dev)
        self.hyy_s = 0.0
        self.Mptr = LA.ptr_table(self.P)
        self.lam = torch.zeros((Nw, dp), dtype=f64, device=dev)
        self.theta = torch.zeros((Nw, dp), dtype=f64, device=dev)
        self.U = torch.zeros((Nw, dp), dtype=f64, device=dev)
        self.fw = torch.zeros((Nw,), dtype=f64, device=dev)
        W = D.world
        row = dp + 64
        # symmetric region: ubuf[2][W][row] | uflag[2][W] | z[dp] | zobj[4] | zflag[1]
        sz_u = 2 * W * row * 8
        off_flag = sz_u
        off_z = off_flag + (2 * W * 8 + 255) // 256 * 256
        off_zobj = off_z + dp * 8
        off_zflag = off_zobj + 256
        self._sym = D.symm_bytes(off_zflag + 256)
        loc = self._sym.local
        self.ubuf = loc[:sz_u].view(f64)
        self.uflag = loc[off_flag:off_flag + 2 * W * 8].view(torch.int64)
        self.z = loc[off_z:off_z + dp * 8].view(f64)
        self.zobj = loc[off_zobj:off_zobj + 32].view(f64)
        self.zflag = loc[off_zflag:off_zflag + 8].view(torch.int64)
        peers = self._sym.peer_ptrs
        self.ubuf_peer = torch.tensor(peers, dtype=torch.int64, device=dev)
        self.uflag_peer = torch.tensor([q + off_flag for q in peers], dtype=torch.int64, device=dev)
        self.zflag_peer = torch.tensor([q + off_zflag for q in peers], dtype=torch.int64, device=dev)
        mc = self._sym.mc_ptr
        self.nvls = bool(W > 1 and mc and nvls in ("auto", True))
        if nvls is True and W > 1 and not mc:
            raise RuntimeError("NVLS multicast is not available for this process group")
        cpe = max(1, (dp // 64) * (dp // 64 - 1) // 2 * 64)
        # server solve: the tile columns of every tile-row pair are divided over nsplit CTAs so that
        # ~2 CTAs per SM stream the one matrix (a single CTA sustains ~25 GB/s)
        npairs = (dp // 64 + 1) // 2
        want = max(1, -(-296 // npairs))
        self.nsplit = int(min(8, 1 << (want - 1).bit_length())) if dp // 64 >= 8 else 1
        self.rowpart = torch.empty((Nw + self.nsplit, dp), dtype=f64, device=dev)
        self.colpart = torch.empty((Nw + 1, cpe), dtype=f64, device=dev)
        self.counters = torch.zeros((Nw + 4,), dtype=torch.int32, device=dev)
        self.zblk = torch.zeros((dp // 64, 2), dtype=f64, device=dev)
        self.ctrl = torch.tensor([1, 0, 1 if stop else 0, 0], dtype=torch.int64, device=dev)
        self.obj_hist = torch.zeros((self.max_iter + 2,), dtype=f64, device=dev)
        self.loss_hist = torch.zeros((self.max_iter + 2,), dtype=f64, device=dev)
        self.t_hist = torch.zeros((self.max_iter + 2,), dtype=torch.int64, device=dev)
        p = self.p = nv.GadmmStar()
        p.Mptr, p.Ms, p.b, p.hyy, p.bs = (self.Mptr.data_ptr(), self.Ps.data_ptr(), self.b.data_ptr(),
                                          self.hyy.data_ptr(), self.bs.data_ptr())
        p.lam, p.theta, p.z, p.U, p.fw = (self.lam.data_ptr(), self.theta.data_ptr(), self.z.data_ptr(),
                                          self.U.data_ptr(), self.fw.data_ptr())
        p.ubuf, p.uflag = self.ubuf.data_ptr(), self.uflag.data_ptr()
        p.ubuf_peer, p.uflag_peer = self.ubuf_peer.data_ptr(), self.uflag_peer.data_ptr()
        p.rowpart, p.colpart, p.counters = (self.rowpart.data_ptr(), self.colpart.data_ptr(),
                                            self.counters.data_ptr())
        p.ctrl, p.obj_hist, p.loss_hist, p.t_hist = (self.ctrl.data_ptr(), self.obj_hist.data_ptr(),
                                                     self.loss_hist.data_ptr(), self.t_hist.data_ptr())
        p.ubuf_mc, p.z_mc, p.zobj_mc = mc, mc + off_z, mc + off_zobj
        p.zobj, p.zflag, p.zflag_peer = self.zobj.data_ptr(), self.zflag.data_ptr(), self.zflag_peer.data_ptr()
        p.nvls, p.server_rank = int(self.nvls), 0
        p.rho, p.obj0, p.acc, p.hyy_s = self.rho, float(obj0), float(acc), self.hyy_s
        p.d, p.dp, p.N, p.Nw, p.rank, p.nranks = self.d, dp, self.N, Nw, D.rank, D.world
        p.store, p.max_iter = self.store, self.max_iter
        p.nsplit, p.zblk = self.nsplit, self.zblk.data_ptr()
        self.launches = 0
        self._graphs = {}
        self.stop = stop
        if XX is not None:
            XX = np.asarray(XX, dtype=np.float64)
            YY = np.asarray(YY, dtype=np.float64).reshape(-1)
            g0, g1 = int(self.gids[0]), int(self.gids[-1]) + 1
            X = torch.from_numpy(XX[g0 * s:g1 * s, :d].reshape(Nw, s, d)).to(dev)
            y = torch.from_numpy(YY[g0 * s:g1 * s].reshape(Nw, s)).to(dev)
            self.ingest_gram(*LA.gram(X, y, dp), 0)
            Xs = torch.from_numpy(XX[(N - 1) * s:N * s, :d].reshape(1, s, d)).to(dev)
            ys = torch.from_numpy(YY[(N - 1) * s:N * s].reshape(1, s)).to(dev)
            self.ingest_server(*LA.gram(Xs, ys, dp))
        D.barrier()

    # ---- set-up from Gram matrices (scaled configs draw the shards on the device batch by batch)
    def ingest_gram(self, G, b, hyy, k0: int):
        """Workers k0..k0+nb of this rank: ``(G_n + rho I)^-1`` packed, ``b_n``, ``0.5 y'y``."""
        nb = G.shape[0]
        self.b[k0:k0 + nb] = b
        self.hyy[k0:k0 + nb] = hyy
        LA.pack_sym(LA.shifted_inverse(G, self.d, self.rho), self.d, self.store, out=self.P[k0:k0 + nb])

    def ingest_server(self, Gs, bs, hs):
        """The server's own shard (worker N, replicated on every rank): ``(G_N + (N-1) rho I)^-1``."""
        LA.pack_sym(LA.shifted_inverse(Gs, self.d, (self.N - 1) * self.rho), self.d, self.store,
                    out=self.Ps)
        self.bs.copy_(bs.reshape(-1))
        self.hyy_s = float(hs.reshape(-1)[0].item())
        self.p.hyy_s = self.hyy_s

    def iterate(self, n, use_graph=True):
        if use_graph and n >= 8:
            self._graph(n).replay()
        else:
            nv.check(nv.lib().gadmm_star_iterate(C.byref(self.p), int(n), nv.stream_ptr()))
        self.launches += 4 * n

    def _graph(self, n):
        g = self._graphs.get(n)
        if g is None:
            L = nv.lib()
            nv.check(L.gadmm_star_iterate(C.byref(self.p), 0, nv.stream_ptr()))   # attributes only
            torch.cuda.synchronize(self.dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                nv.check(L.gadmm_star_iterate(C.byref(self.p), int(n), nv.stream_ptr()))
            self._graphs[n] = g
        return g

    def run(self, num_iter=None, window=64):
        T = self.max_iter if num_iter is None else min(int(num_iter), self.max_iter)
        if T >= 8:
            self._graph(min(window, T))        # capture before the clock starts
        done = getattr(self, "_done", 0)       # > 0 after load_state()
        while done < T:
            n = min(window, T - done)
            self.iterate(n)
            done += n
            if self.stop and int(self.ctrl[1].item()):
                break
        torch.cuda.synchronize(self.dev)
        hit = int(self.ctrl[1].item())
        return hit if hit else T

    def results(self, Iter):
        obj = self.obj_hist[1:Iter + 1].cpu().numpy()
        loss = self.loss_hist[1:Iter + 1].cpu().numpy()
        t = self.t_hist[1:Iter + 1].to(torch.float64)
        live = t > 0                                # restored iterations carry no clock
        t0 = t[live][0] if bool(live.any()) else t[0]
        tv = torch.clamp((t - t0) * 1e-9, min=0.0)
        if self.dist.world > 1:
            import torch.distributed as dist
            dist.all_reduce(tv, op=dist.ReduceOp.MAX)
        return obj, loss, Iter, tv.cpu().numpy()


    # ------------------------------------------------------------------ checkpoint / resume
    def state(self):
        """Solver state: this rank's worker models and duals, the consensus
        model z, the iteration, the curves.  ``dual`` is the reference's lambda_n after the dual
        update of the last iteration (the kernel itself folds that update into the next solve)."""
        torch.cuda.synchronize(self.dev)
        it = int(self.ctrl[0].item()) - 1
        d = self.d
        lam_ref = self.lam + self.rho * (self.theta - self.z.unsqueeze(0))
        return {"theta": self.theta[:, :d].cpu().numpy().T.copy(), "z": self.z[:d].cpu().numpy().copy(),
                "dual": lam_ref[:, :d].cpu().numpy().T.copy(),
                "lam_lagged": self.lam[:, :d].cpu().numpy().T.copy(),
                "g0": int(self.gids[0]), "iteration": it,
                "obj": self.obj_hist[1:it + 1].cpu().numpy().copy(),
                "loss": self.loss_hist[1:it + 1].cpu().numpy().copy(), "rho": self.rho}

    def load_state(self, st):
        """Resume from :meth:`state` (same partition): the next iteration is ``iteration + 1``."""
        it, d = int(st["iteration"]), self.d

        def dev_t(a):
            return torch.as_tensor(np.asarray(a, dtype=np.float64).T.copy(), device=self.dev)

        self.theta.zero_()
        self.lam.zero_()
        self.z.zero_()
        self.theta[:, :d] = dev_t(st["theta"])
        self.lam[:, :d] = dev_t(st["lam_lagged"])
        self.z[:d] = torch.as_tensor(np.asarray(st["z"], dtype=np.float64), device=self.dev)
        n = min(it, len(st.get("obj", [])))
        if n:
            self.obj_hist[1:n + 1] = torch.as_tensor(np.asarray(st["obj"][:n]), device=self.dev)
            self.loss_hist[1:n + 1] = torch.as_tensor(np.asarray(st["loss"][:n]), device=self.dev)
        self.uflag.zero_()
        self.zflag.zero_()
        self.ctrl[0] = it + 1
        self.ctrl[1] = 0
        self._done = it
        torch.cuda.synchronize(self.dev)
        self.dist.barrier()


def star_admm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc, store="f64",
              dist=None, window=64, return_engine=False, nvls="auto", state=None,
              return_state=False):
    eng = StarEngine(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc, dist=dist,
                     store=store, nvls=nvls)
    if state is not None:
        eng.load_state(state)
    it = eng.run(num_iter, window=window)
    res = eng.results(it)
    if return_state:
        res = res + (eng.state(),)
    return res + (eng,) if return_engine else res
