"""Process-group runtime: one process per GPU, NCCL for plumbing, symmetric memory for the product.

``Dist`` hides whether we run single-GPU (plain tensors, peer table = self) or multi-GPU
(``torch.distributed._symmetric_memory`` buffers mapped into every peer; the kernels then issue
``st.global`` / ``ld.acquire.sys`` on those peer pointers directly -- SURVEY.md section 5.8).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist


def env_rank_world() -> Tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_process_group(backend: Optional[str] = None):
    """Initialise torch.distributed from the torchrun environment (idempotent)."""
    rank, world, local = env_rank_world()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world, local


def partition(n_workers: int, world: int) -> np.ndarray:
    """Contiguous chain segments: worker g is owned by rank ``g*world // n_workers``-style blocks
    (first ``n % world`` ranks get one extra worker)."""
    base, extra = divmod(n_workers, world)
    owner = np.empty(n_workers, dtype=np.int32)
    g = 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        owner[g:g + cnt] = r
        g += cnt
    return owner


@dataclass
class SymmBuffer:
    """A byte buffer that exists at the same size on every rank, with all peers' base pointers."""
    local: torch.Tensor          # uint8, this rank's buffer
    peer_ptrs: List[int]         # device pointer of the buffer on every rank (index = rank)
    handle: object = None
    mc_ptr: int = 0              # NVLS multicast alias of the buffer (0 = no multicast support)


class Dist:
    def __init__(self, device: Optional[torch.device] = None):
        self.rank, self.world, self.local_rank = env_rank_world()
        if self.world > 1 and not dist.is_initialized():
            init_process_group()
        if device is None:
            device = torch.device("cuda", self.local_rank if self.world > 1
                                  else torch.cuda.current_device())
        self.device = device
        if self.world > 1:
            torch.cuda.set_device(device)

    # ------------------------------------------------------------------ symmetric memory
    def symm_bytes(self, nbytes: int) -> SymmBuffer:
        nbytes = (nbytes + 255) // 256 * 256
        if self.world == 1:
            t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            return SymmBuffer(t, [t.data_ptr()])
        import torch.distributed._symmetric_memory as symm_mem
        group = dist.group.WORLD
        t = symm_mem.empty(nbytes, dtype=torch.uint8, device=self.device)
        try:
            hdl = symm_mem.rendezvous(t, group=group)
        except TypeError:
            hdl = symm_mem.rendezvous(t, group.group_name)
        t.zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier()
        mc = 0
        try:
            mc = int(hdl.multicast_ptr or 0)
        except Exception:
            mc = 0
        return SymmBuffer(t, [int(p) for p in hdl.buffer_ptrs], hdl, mc)

    def barrier(self):
        if self.world > 1:
            torch.cuda.synchronize(self.device)
            dist.barrier()

    def all_reduce_max(self, x: float) -> float:
        if self.world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def broadcast_object(self, obj, src: int = 0):
        if self.world == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src)
        return box[0]


class SoloDist(Dist):
    """A world-of-one view on this process's GPU, used under torchrun for problems so small that
    splitting them over GPUs only adds NVLink latency: every rank then solves the whole problem
    with the single-GPU persistent kernel (identical results everywhere, no communication)."""

    def __init__(self, device: Optional[torch.device] = None):
        _, world, local = env_rank_world()
        self.rank, self.world, self.local_rank = 0, 1, local
        self.device = device or torch.device("cuda", local if world > 1 else torch.cuda.current_device())
        if world > 1:
            torch.cuda.set_device(self.device)
