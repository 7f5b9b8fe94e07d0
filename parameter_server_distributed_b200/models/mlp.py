"""Synthetic MLP — BASELINE.json config 2 ("synthetic MLP sync-SGD with 8 PS shards / 8
workers").  The reference has no model at all: its only tensor is a fabricated
``weight[10,10]`` (src/worker.cpp:346-353) with constant gradients (:316-320).

The first Linear can run through the fused all-gather + tcgen05 GEMM (K2f) when the
parameter server is in pull mode.  On a GPU every layer is a ``[GELU ->] Linear [-> GELU]``
link of :mod:`..ops.linear`: forward, dX and dW run on the repo's CTA-pair tcgen05 kernel with
the activation and its derivative fused into the GEMM epilogues; on the CPU the same links
are plain PyTorch.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import linear, native


@dataclass
class MLPConfig:
    d_in: int = 4096
    hidden: Sequence[int] = (8192, 8192)
    d_out: int = 1024
    bias: bool = True

    @property
    def dims(self):
        return [self.d_in, *self.hidden, self.d_out]

    def num_params(self) -> int:
        d = self.dims
        return sum(d[i] * d[i + 1] + (d[i + 1] if self.bias else 0) for i in range(len(d) - 1))

    def flops_per_sample(self) -> int:
        """Forward + dW of every layer, dX of every layer but the first (the input needs no
        gradient, so that GEMM is never run): 6 P - 2 d0 d1 per sample."""
        d = self.dims
        return 6 * sum(d[i] * d[i + 1] for i in range(len(d) - 1)) - 2 * d[0] * d[1]


class MLP(nn.Module):
    grad_sink_layers = True  # its links write dW / db through the engine's gradient sinks

    def __init__(self, cfg: MLPConfig = MLPConfig()):
        super().__init__()
        self.cfg = cfg
        d = cfg.dims
        self.layers = nn.ModuleList(
            [nn.Linear(d[i], d[i + 1], bias=cfg.bias) for i in range(len(d) - 1)])
        self.first_linear_fn = None  # set by the trainer to route layer 0 through K2f

    def first_weight_name(self) -> str:
        return "layers.0.weight"

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        n = len(self.layers)
        z, act = x, None  # pre-activation of the previous layer and gelu() of it
        for i, layer in enumerate(self.layers):
            if i == 0 and self.first_linear_fn is not None:
                z, act = self.first_linear_fn(x, layer.weight, layer.bias, gelu_out=n > 1)
            else:
                z, act = linear.link(z, act, layer.weight, layer.bias,
                                     gelu_in=i > 0, gelu_out=i + 1 < n)
        return z


class _MseFn(torch.autograd.Function):
    """mean((out - target)^2) in fp32 from bf16 inputs: one kernel forward, one backward."""

    @staticmethod
    def forward(ctx, out, target):
        ctx.save_for_backward(out, target)
        return native.mse_fwd(out, target)

    @staticmethod
    def backward(ctx, g):
        out, target = ctx.saved_tensors
        return native.mse_bwd(out, target, g.float().contiguous()), None


def mse_loss(out: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    if (linear.ENABLED and out.is_cuda and out.dtype == torch.bfloat16
            and target.dtype == torch.bfloat16 and out.shape == target.shape
            and out.is_contiguous() and target.is_contiguous()):
        return _MseFn.apply(out, target)
    return F.mse_loss(out.float(), target.float())
