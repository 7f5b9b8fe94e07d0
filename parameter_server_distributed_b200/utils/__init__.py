"""Timing, clock sampling, tracing helpers."""
from .timing import ClockSampler, cuda_time_ms, flush_l2
from .tracing import Metrics, nvtx_range

__all__ = ["ClockSampler", "cuda_time_ms", "flush_l2", "Metrics", "nvtx_range"]
