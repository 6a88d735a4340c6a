from .clocks import ClockSampler
from .timing import cuda_time_ms, flush_l2
from .trace import chrome_trace, iteration_stats

__all__ = ["ClockSampler", "cuda_time_ms", "flush_l2", "chrome_trace", "iteration_stats"]
