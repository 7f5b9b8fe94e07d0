"""Timing, clock sampling, tracing helpers."""
from .timing import ClockSampler, cuda_time_ms, flush_l2

__all__ = ["ClockSampler", "cuda_time_ms", "flush_l2"]
