from .clocks import ClockSampler
from .timing import cuda_time_ms, flush_l2

__all__ = ["ClockSampler", "cuda_time_ms", "flush_l2"]
