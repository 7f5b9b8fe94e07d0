"""Plain-PyTorch fp32 restatements of the device ops.

Two uses: (1) the oracle every CUDA-kernel test compares against, (2) the host backend of
the engine (CPU tensors, ``gloo``), which keeps the control plane and the bucket / shard
logic testable without a GPU.  The math generalises the reference's server update
(src/parameter_server.cpp:49-60 sum + mean, :83-90 ``p -= g``) with lr, momentum, Adam(W).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch

from .native import OptimConfig


def optim_step(cfg: OptimConfig, master: torch.Tensor, grad: torch.Tensor,
               m: Optional[torch.Tensor] = None, v: Optional[torch.Tensor] = None,
               out_bf16: Optional[torch.Tensor] = None, clip_coef: float = 1.0,
               step: int = 0) -> None:
    """In-place update of fp32 ``master`` (and ``m``/``v``); ``step`` = completed steps."""
    g = grad.to(torch.float32) * (cfg.grad_scale * float(clip_coef))
    if cfg.kind == "sgd":
        g = g + cfg.weight_decay * master
        master.add_(g, alpha=-cfg.lr)
    elif cfg.kind == "momentum":
        g = g + cfg.weight_decay * master
        m.mul_(cfg.momentum).add_(g)
        master.add_(m, alpha=-cfg.lr)
    elif cfg.kind in ("adam", "adamw"):
        t = step + 1
        if cfg.kind == "adam":
            g = g + cfg.weight_decay * master
        else:
            master.mul_(1.0 - cfg.lr * cfg.weight_decay)
        m.mul_(cfg.beta1).add_(g, alpha=1.0 - cfg.beta1)
        v.mul_(cfg.beta2).addcmul_(g, g, value=1.0 - cfg.beta2)
        bias1 = 1.0 - cfg.beta1 ** t
        bias2 = 1.0 - cfg.beta2 ** t
        denom = v.sqrt() / math.sqrt(bias2) + cfg.eps
        master.addcdiv_(m, denom, value=-cfg.lr / bias1)
    else:
        raise ValueError(f"unknown optimizer kind {cfg.kind!r}")
    if out_bf16 is not None:
        out_bf16.copy_(master.to(torch.bfloat16))


def reduce_mean(grads: Sequence[torch.Tensor], scale: Optional[float] = None) -> torch.Tensor:
    """Fixed-order (w = 0..N-1) fp32 sum times ``scale`` (default 1/N)."""
    acc = torch.zeros_like(grads[0], dtype=torch.float32)
    for g in grads:
        acc += g.to(torch.float32)
    return acc * (1.0 / len(grads) if scale is None else scale)


def sumsq(x: torch.Tensor) -> torch.Tensor:
    return (x.to(torch.float32) ** 2).sum().reshape(1)


def clip_coef(total_sumsq: float, max_norm: float, eps: float = 1e-6) -> float:
    norm = math.sqrt(total_sumsq)
    return min(1.0, max_norm / (norm + eps))


def gemm_bf16(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    y = x.to(torch.float32) @ w.to(torch.float32).t()
    if bias is not None:
        y = y + bias.to(torch.float32)
    return y.to(out_dtype)
