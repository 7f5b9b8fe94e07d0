"""Device ops (sm_100a kernels, :mod:`.native`) and their fp32 PyTorch oracles (:mod:`.reference`)."""
from . import native, reference
from .native import (OptimConfig, cast_scale, clip_coef, counter_add_, fill_, gather_to_bucket,
                     gemm_bf16, optim_step, sumsq)

__all__ = ["native", "reference", "OptimConfig", "cast_scale", "clip_coef", "counter_add_", "fill_",
           "gather_to_bucket", "gemm_bf16", "optim_step", "sumsq"]
