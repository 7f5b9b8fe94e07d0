"""Linear layers whose three GEMMs (forward, dX, dW) run on the repo's own tcgen05 kernel.

The reference has no model code (its worker fabricates a 10x10 gradient,
src/worker.cpp:316-320, 346-353); this is the compute path of the flagship MLP.

One autograd function covers a whole ``[GELU ->] Linear [-> GELU]`` link so that the
activation never costs a pass over HBM of its own:

* forward: ``Z = A · Wᵀ + b`` with ``gelu(Z)`` produced by the same epilogue (Z itself is
  stored once, in bf16, for the backward pass);
* backward: ``dW = dZᵀ · A`` reads both operands in place as MN-major tiles (no transposed
  copy), ``dX = dZ · W`` reads ``W`` in place as an MN-major B operand, and when the input of
  this layer was ``gelu(Z_prev)`` the epilogue of the dX GEMM multiplies by ``gelu'(Z_prev)``,
  i.e. it returns the gradient of the previous layer's pre-activation directly.

CPU tensors (and shapes whose rows are not 16-byte multiples) take the plain PyTorch path
with the same rounding points, which is also the oracle of the GPU tests.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import native


# PSB_TCGEN05_LINEAR=0 (or ``linear.ENABLED = False``) sends every link to the library GEMMs
# of PyTorch: the A/B switch of ``bench.py --mlp-gemm``.
ENABLED = os.environ.get("PSB_TCGEN05_LINEAR", "1") != "0"


def supported(x: torch.Tensor, weight: torch.Tensor) -> bool:
    """Can this link run on the tcgen05 kernel?  bf16 CUDA operands, all row lengths multiples of 8."""
    return (ENABLED and x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
            and weight.dim() == 2 and weight.shape[0] % 8 == 0 and weight.shape[1] % 8 == 0)


def _gelu(z: torch.Tensor) -> torch.Tensor:
    return F.gelu(z, approximate="tanh")


def _dgelu(z: torch.Tensor) -> torch.Tensor:
    z = z.float()
    k0, k1 = 0.7978845608028654, 0.044715
    t = torch.tanh(k0 * (z + k1 * z * z * z))
    return 0.5 * (1 + t) + 0.5 * z * (1 - t * t) * k0 * (1 + 3 * k1 * z * z)


def take_sink(param, needs_grad: bool):
    """In a layer's forward: the parameter's gradient sink, if the engine installed one."""
    sink = getattr(param, "_psb_sink", None) if (param is not None and needs_grad) else None
    return sink.acquire() if sink is not None else None


def grad_to_sink(sink, produce):
    """Run ``produce(out)``; with a sink the result goes into the gradient bucket, is committed
    and ``None`` is returned to autograd, otherwise a fresh tensor is returned.

    The commit may queue the push of the bucket, which updates the parameters IN PLACE on the
    push stream: call this only after every kernel of the layer that reads the parameter (the
    dX GEMM) has been queued on the current stream."""
    if sink is None or not sink.exclusive or len(sink.parts()) > 1:
        return produce(None)  # (a row-split tensor delivered whole is cut up by the engine's hook)
    produce(sink.view())
    sink.commit()
    return None


class _LinkFn(torch.autograd.Function):
    """``(z_in, act_in, W, b) -> (z_out, act_out)``.

    ``gelu_in``: the GEMM input is ``act_in = gelu(z_in)`` (computed by the previous link and
    passed along, not differentiable here); the gradient is returned for ``z_in``.  Without
    ``gelu_in`` the GEMM input is ``z_in`` itself and ``act_in`` is ignored.
    ``gelu_out``: also return ``gelu(z_out)`` (not differentiable: the next link owns that
    derivative).
    """

    @staticmethod
    def forward(ctx, z_in, act_in, weight, bias, gelu_in: bool, gelu_out: bool):
        a = act_in if gelu_in else z_in
        a2d = a.reshape(-1, a.shape[-1])
        if not a2d.is_contiguous():
            a2d = a2d.contiguous()
        w = weight if weight.is_contiguous() else weight.contiguous()
        n_out = w.shape[0]
        b = bias.detach() if bias is not None else None  # bf16 or fp32, read by the epilogue
        if b is not None and (b.dtype not in (torch.bfloat16, torch.float32) or not b.is_contiguous()):
            b = b.float().contiguous()
        z_out = torch.empty((a2d.shape[0], n_out), dtype=torch.bfloat16, device=a2d.device)
        if gelu_out:
            act_out = native.gemm(a2d, w, bias=b, epilogue="gelu", aux=z_out)
        else:
            native.gemm(a2d, w, bias=b, out=z_out)
            act_out = z_out.new_empty(0)
        z_in2d = z_in.reshape(-1, z_in.shape[-1]) if gelu_in else None
        ctx.save_for_backward(a2d, w, z_in2d if gelu_in else None)
        ctx.gelu_in = gelu_in
        ctx.has_bias = bias is not None
        # parameter-server gradient sinks (parallel/engine.py: GradSink), if the engine set them
        ctx.sink_w = take_sink(weight, ctx.needs_input_grad[2])
        ctx.sink_b = take_sink(bias, bias is not None and ctx.needs_input_grad[3])
        ctx.in_shape = z_in.shape
        ctx.mark_non_differentiable(act_out)
        ctx.set_materialize_grads(False)  # no zero tensor for the gradient of act_out
        lead = z_in.shape[:-1]
        return z_out.view(*lead, n_out), (act_out.view(*lead, n_out) if gelu_out else act_out)

    @staticmethod
    def backward(ctx, gz, _g_act):
        if gz is None:
            return None, None, None, None, None, None
        a2d, w, z_in2d = ctx.saved_tensors
        gz2d = gz.reshape(-1, gz.shape[-1])
        if not gz2d.is_contiguous():
            gz2d = gz2d.contiguous()
        g_in = g_w = g_b = None
        def input_grad():
            if not ctx.needs_input_grad[0]:
                return None
            if ctx.gelu_in:
                zc = z_in2d if z_in2d.is_contiguous() else z_in2d.contiguous()
                g = native.gemm(gz2d, w, trans_b=True, epilogue="dgelu", aux=zc)
            else:
                g = native.gemm(gz2d, w, trans_b=True)
            return g.view(ctx.in_shape)

        def param_grads():
            # with a sink the parameter gradients land in the gradient bucket: no staging copy
            gb = gw = None
            if ctx.has_bias and ctx.needs_input_grad[3]:
                gb = grad_to_sink(ctx.sink_b, lambda out: native.colsum(gz2d, out=out))
            if ctx.needs_input_grad[2]:
                sink = ctx.sink_w
                if sink is not None and sink.exclusive and len(sink.parts()) > 1:
                    # the plan cut this weight into row blocks with a bucket each: dW block by
                    # block (dY[:, lo:hi]^T · X, operands in place), each committed at once so
                    # that its push runs under the next block's GEMM
                    for lo, hi, out, commit in sink.parts():
                        native.gemm(gz2d[:, lo:hi], a2d, trans_a=True, trans_b=True, out=out)
                        commit()
                else:
                    gw = grad_to_sink(sink, lambda out: native.gemm(
                        gz2d, a2d, trans_a=True, trans_b=True, out=out))
            return gw, gb

        # A commit may start the push that rewrites W in place, so normally every reader of W
        # (the dX GEMM) is queued first.  With sinks marked `early` (the push writes the other
        # parameter replica) dW goes first and its push overlaps the dX GEMM.
        sinks = [s for s in (ctx.sink_w, ctx.sink_b) if s is not None and s.exclusive]
        if sinks and all(s.early for s in sinks):
            g_w, g_b = param_grads()
            g_in = input_grad()
        else:
            g_in = input_grad()
            g_w, g_b = param_grads()
        return g_in, None, g_w, g_b, None, None


def link(z_in: torch.Tensor, act_in: Optional[torch.Tensor], weight: torch.Tensor,
         bias: Optional[torch.Tensor], gelu_in: bool, gelu_out: bool
         ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """One ``[GELU ->] Linear [-> GELU]`` link; returns ``(z_out, gelu(z_out) or None)``.

    ``act_in`` must be ``gelu(z_in)`` when ``gelu_in`` (pass ``None`` to have it computed).
    """
    if supported(z_in, weight):
        if gelu_in and act_in is None:
            act_in = _gelu(z_in.detach())
        if act_in is None:
            act_in = z_in.new_empty(0)
        z_out, act_out = _LinkFn.apply(z_in, act_in.detach(), weight, bias, gelu_in, gelu_out)
        return z_out, (act_out if gelu_out else None)
    a = _gelu(z_in) if gelu_in else z_in
    return F.linear(a, weight, bias), None  # the next link applies its own GELU


def link_reference(z_in, weight, bias, gelu_in: bool, gelu_out: bool):
    """PyTorch statement of :func:`link` (autograd does the backward): the test oracle."""
    a = _gelu(z_in) if gelu_in else z_in
    z_out = F.linear(a, weight, bias)
    return z_out, (_gelu(z_out) if gelu_out else None)
