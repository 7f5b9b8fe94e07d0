"""First-layer Linear through K2f: parameter pull (all-gather of the row-sharded weight)
fused with the tcgen05 GEMM that consumes it.

In pull mode the owner of each row-slice of ``W`` publishes only its own slice (K1 without
broadcast).  The forward of the first layer then launches ONE kernel that pulls the remote
slices over NVLink with bulk copies into the local replica while the tensor cores already
work on the locally owned tiles.  Backward: dX and dW on the same tcgen05 kernel (MN-major
operands, ``ops.native.gemm``), written through the engine's gradient sinks.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _lib
from .._lib import PsbGemmAG, check
from ..ops import linear, native


class FusedFirstLinear:
    """Callable ``(x, weight, bias) -> y`` bound to one dedicated bucket of a parameter server."""

    @staticmethod
    def supported(shape, world: int) -> bool:
        """The fused kernel needs every 256-row N-tile to have exactly one owner and K to be a
        multiple of 64; otherwise the layer is pulled by K2 like any other bucket."""
        if len(shape) != 2:
            return False
        n, k = shape
        return n % (256 * world) == 0 and k % 64 == 0 and (n * k) % (world * 512) == 0

    def __init__(self, ps, weight_name: str):
        slot = ps.plan.slots[weight_name]
        bk = ps.plan.buckets[slot.bucket]
        if not bk.dedicated or slot.offset != 0:
            raise ValueError(f"{weight_name} must live in a dedicated bucket")
        if len(slot.shape) != 2:
            raise ValueError("fused first layer needs a 2-D weight")
        N, K = slot.shape
        if bk.numel != N * K:
            raise ValueError(f"{weight_name}: {N}x{K} is not divisible into {ps.world} row slices "
                             "of 512-element granularity")
        be = ps.backend
        a = PsbGemmAG()
        a.fabric = ps.fabric.c_fabric()
        for r in range(ps.world):
            a.w[r] = be.P.peer_ptr(r) + bk.base * 2
        a.N, a.K = N, K
        a.bucket = slot.bucket
        a.epoch = be.epoch[slot.bucket:].data_ptr()
        a.skip_wait = 0
        self.args = a
        self.ps = ps
        self.device = ps.device
        self.lib = _lib.load()
        ps._fused_first_layer = True

    def launch(self, x2d: torch.Tensor, bias: Optional[torch.Tensor], gelu: bool = False,
               out_dtype: torch.dtype = torch.bfloat16):
        """Returns ``(z, act)``: ``z = x·Wᵀ + b`` and, with ``gelu``, ``act = gelu(z)`` from the
        same epilogue (``act`` is None otherwise)."""
        a = self.args
        M = x2d.shape[0]
        z = torch.empty((M, a.N), dtype=out_dtype, device=x2d.device)
        act = torch.empty_like(z) if gelu else None
        a.x = x2d.data_ptr()
        a.bias = bias.data_ptr() if bias is not None else None
        a.bias_dtype = _lib.BF16 if (bias is not None and bias.dtype == torch.bfloat16) else _lib.F32
        a.y = act.data_ptr() if gelu else z.data_ptr()
        a.aux_out = z.data_ptr() if gelu else None
        a.epilogue = _lib.EPI_GELU if gelu else _lib.EPI_NONE
        a.M = M
        a.out_dtype = _lib.BF16 if out_dtype == torch.bfloat16 else _lib.F32
        with torch.cuda.device(self.device):
            check(self.lib.psb_gemm_allgather_bf16(
                C.byref(a), torch.cuda.current_stream(self.device).cuda_stream),
                "psb_gemm_allgather_bf16")
        return z, act

    def __call__(self, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor],
                 gelu_out: bool = False):
        """``(z, gelu(z) or None)`` like :func:`..ops.linear.link`."""
        z, act = _FusedLinearFn.apply(x, weight, bias, self, gelu_out)
        return z, (act if gelu_out else None)


class _FusedLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, op: FusedFirstLinear, gelu_out: bool):
        x2d = x.reshape(-1, x.shape[-1]).contiguous()
        b = bias.detach() if bias is not None else None
        if b is not None and (b.dtype not in (torch.bfloat16, torch.float32) or not b.is_contiguous()):
            b = b.float().contiguous()
        y, act = op.launch(x2d, b, gelu=gelu_out)
        if act is None:
            act = y.new_empty(0)
        ctx.mark_non_differentiable(act)  # the next link owns gelu'
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(x2d, weight)
        ctx.has_bias = bias is not None
        ctx.sink_w = linear.take_sink(weight, ctx.needs_input_grad[1])
        ctx.sink_b = linear.take_sink(bias, bias is not None and ctx.needs_input_grad[2])
        ctx.x_shape = x.shape
        lead = x.shape[:-1]
        return (y.view(*lead, weight.shape[0]),
                act.view(*lead, weight.shape[0]) if gelu_out else act)

    @staticmethod
    def backward(ctx, gy, _g_act):
        if gy is None:
            return None, None, None, None, None
        x2d, weight = ctx.saved_tensors
        gy2d = gy.reshape(-1, gy.shape[-1])
        if not gy2d.is_contiguous():
            gy2d = gy2d.contiguous()
        gx = gw = gb = None
        own = linear.supported(gy2d, weight)  # dW / dX on the tcgen05 kernel, in place
        if ctx.needs_input_grad[0]:  # readers of W before any commit (see grad_to_sink)
            gx = (native.gemm(gy2d, weight, trans_b=True) if own else gy2d @ weight).view(ctx.x_shape)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = linear.grad_to_sink(ctx.sink_b, lambda out: (
                native.colsum(gy2d, out=out) if own else torch.sum(gy2d, 0, out=out)))
        if ctx.needs_input_grad[1]:
            if own:
                gw = linear.grad_to_sink(ctx.sink_w, lambda out: native.gemm(
                    gy2d, x2d, trans_a=True, trans_b=True, out=out))
            else:
                gw = gy2d.t() @ x2d
        return gx, gw, gb, None, None
