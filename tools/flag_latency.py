"""NVLink flag latency on this box (run under torchrun, >= 2 GPUs): round trip of the
store -> release flag -> acquire hand-off between GPU 0 and every other GPU, and the time of the
all-rank scalar exchange used by the dual/objective kernels.  Writes one JSON line on rank 0."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from gadmm_b200.ops import native as nv
from gadmm_b200.parallel.runtime import Dist, init_process_group


def main():
    rank, world, local = init_process_group("nccl")
    torch.cuda.set_device(local)
    D = Dist()
    dev = D.device
    L = nv.lib()
    iters = 2000
    # symmetric region: flag[world] | word[world] | objpart[4*world] | objflag[4*world]
    W = world
    sym = D.symm_bytes(256 * 4 + 8 * W * 2 + 64 * W * 2 + 4096)
    o_flag, o_word, o_part, o_oflag = 0, 1024, 2048, 2048 + 64 * W
    base = sym.local.data_ptr()
    out = torch.zeros((1,), dtype=torch.int64, device=dev)
    osum = torch.zeros((1,), dtype=torch.float64, device=dev)
    res = {"gpus": world, "iters": iters, "pingpong_round_trip_us": {}}
    for peer in range(1, world):
        sym.local.zero_()
        torch.cuda.synchronize()
        D.barrier()
        if rank in (0, peer):
            other = peer if rank == 0 else 0
            # slot `peer` of both ranks is used for this pair
            nv.check(L.gadmm_probe_pingpong(
                C.c_void_p(base + o_flag + 8 * peer), C.c_void_p(base + o_word + 8 * peer),
                C.c_void_p(sym.peer_ptrs[other] + o_flag + 8 * peer),
                C.c_void_p(sym.peer_ptrs[other] + o_word + 8 * peer),
                iters, 0 if rank == 0 else 1, nv.ptr(out), nv.stream_ptr()))
        torch.cuda.synchronize()
        D.barrier()
        if rank == 0:
            res["pingpong_round_trip_us"][f"0<->{peer}"] = round(int(out.item()) / iters * 1e-3, 3)
    sym.local.zero_()
    torch.cuda.synchronize()
    D.barrier()
    pp = torch.tensor([q + o_part for q in sym.peer_ptrs], dtype=torch.int64, device=dev)
    fp = torch.tensor([q + o_oflag for q in sym.peer_ptrs], dtype=torch.int64, device=dev)
    nv.check(L.gadmm_probe_exchange(C.c_void_p(base + o_part), C.c_void_p(base + o_oflag), nv.ptr(pp),
                                    nv.ptr(fp), rank, world, iters, C.c_ulonglong(0), nv.ptr(out),
                                    nv.ptr(osum), nv.stream_ptr()))
    torch.cuda.synchronize()
    t = torch.tensor([float(out.item())], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        res["all_rank_exchange_us"] = round(float(t.item()) / iters * 1e-3, 3)
        res["exchange_sum_check"] = float(osum.item())
        res["expected_sum"] = world * (world + 1) / 2
        print(json.dumps(res))
    D.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
