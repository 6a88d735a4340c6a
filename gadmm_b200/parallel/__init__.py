"""Process-group runtime (one process per GPU), symmetric-memory buffers, and the
``torch.distributed`` baselines (NCCL on GPUs, gloo on CPU)."""
from .runtime import Dist, SymmBuffer, env_rank_world, init_process_group, partition

__all__ = ["Dist", "SymmBuffer", "env_rank_world", "init_process_group", "partition"]
