"""Tensor-level wrappers over the sm_100a kernels in ``libpsb_kernels.so``.

Every function launches on ``torch.cuda.current_stream()`` of the tensor's device and
raises :class:`PsbError` on failure.  CUDA tensors never fall back to PyTorch here: the
plain-PyTorch restatements used as test oracles and by the CPU plumbing path live in
:mod:`.reference`.

Kernel IDs refer to SURVEY.md §8; reference citations are paths below /root/reference.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import torch

from .. import _lib
from .._lib import BF16, F32, PsbGatherList, PsbOptim, check

_KINDS = {"sgd": _lib.OPT_SGD, "momentum": _lib.OPT_MOMENTUM, "adam": _lib.OPT_ADAM,
          "adamw": _lib.OPT_ADAMW}


@dataclass
class OptimConfig:
    """Server-side update rule.  ``kind='sgd', lr=1`` is the reference's ``p -= mean(g)``
    (src/parameter_server.cpp:83-90)."""
    kind: str = "sgd"
    lr: float = 1.0
    momentum: float = 0.9
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 0.0
    grad_scale: float = 1.0

    def needs_m(self) -> bool:
        return self.kind != "sgd"

    def needs_v(self) -> bool:
        return self.kind in ("adam", "adamw")

    def to_c(self, grad_scale: Optional[float] = None) -> PsbOptim:
        if self.kind not in _KINDS:
            raise ValueError(f"unknown optimizer kind {self.kind!r}")
        return PsbOptim(_KINDS[self.kind], self.lr, self.momentum, self.beta1, self.beta2, self.eps,
                        self.weight_decay, self.grad_scale if grad_scale is None else grad_scale)


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported dtype {t.dtype}; expected float32 or bfloat16")


def _need_cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise TypeError("native ops need CUDA tensors; use ops.reference on the host")
        if t is not None and not t.is_contiguous():
            raise ValueError("native ops need contiguous tensors")


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# ---------------------------------------------------------------- K7
def fill_(t: torch.Tensor, value: float) -> torch.Tensor:
    """K7: constant fill (device-side replacement of src/worker.cpp:317-320)."""
    _need_cuda(t)
    lib = _lib.load()
    with torch.cuda.device(t.device):
        if t.dtype == torch.float32:
            check(lib.psb_fill_f32(t.data_ptr(), t.numel(), value, _stream(t)), "psb_fill_f32")
        elif t.dtype == torch.bfloat16:
            check(lib.psb_fill_bf16(t.data_ptr(), t.numel(), value, _stream(t)), "psb_fill_bf16")
        else:
            raise TypeError(f"fill_: unsupported dtype {t.dtype}")
    return t


# ---------------------------------------------------------------- K6
def cast_scale(src: torch.Tensor, dst: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """K6: dst = cast(src * scale) between fp32 and bf16 (src/worker.cpp:436-438)."""
    _need_cuda(src, dst)
    if src.numel() != dst.numel():
        raise ValueError("cast_scale: size mismatch")
    lib = _lib.load()
    with torch.cuda.device(src.device):
        if src.dtype == torch.float32 and dst.dtype == torch.bfloat16:
            check(lib.psb_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), scale,
                                           _stream(src)), "psb_cast_f32_to_bf16")
        elif src.dtype == torch.bfloat16 and dst.dtype == torch.float32:
            check(lib.psb_cast_bf16_to_f32(src.data_ptr(), dst.data_ptr(), src.numel(), scale,
                                           _stream(src)), "psb_cast_bf16_to_f32")
        else:
            raise TypeError(f"cast_scale: unsupported {src.dtype} -> {dst.dtype}")
    return dst


def gather_to_bucket(srcs: Sequence[torch.Tensor], bucket: torch.Tensor,
                     offsets: Sequence[int], scale: float = 1.0) -> None:
    """K6 multi-tensor: copy gradient tensors into their slots of a flat bucket, one launch per
    96 tensors.  ``offsets`` are element offsets into ``bucket`` (multiples of 8)."""
    if not srcs:
        return
    _need_cuda(bucket, *srcs)
    lib = _lib.load()
    sdt = _dt(srcs[0])
    with torch.cuda.device(bucket.device):
        stream = _stream(bucket)
        for lo in range(0, len(srcs), _lib.MAX_GATHER):
            chunk = srcs[lo:lo + _lib.MAX_GATHER]
            gl = PsbGatherList()
            for i, t in enumerate(chunk):
                if _dt(t) != sdt:
                    raise TypeError("gather_to_bucket: mixed source dtypes")
                gl.src[i] = t.data_ptr()
                gl.dst_offset[i] = int(offsets[lo + i])
                gl.numel[i] = t.numel()
            gl.count = len(chunk)
            gl.src_dtype = sdt
            gl.dst_dtype = _dt(bucket)
            gl.scale = scale
            check(lib.psb_gather_to_bucket(C.byref(gl), bucket.data_ptr(), stream),
                  "psb_gather_to_bucket")


# ---------------------------------------------------------------- K3
def optim_step(cfg: OptimConfig, master: torch.Tensor, grad: torch.Tensor,
               m: Optional[torch.Tensor] = None, v: Optional[torch.Tensor] = None,
               out_bf16: Optional[torch.Tensor] = None, clip_coef: Optional[torch.Tensor] = None,
               step: Optional[torch.Tensor] = None) -> None:
    """K3: flat-shard SGD / momentum / Adam(W) on fp32 master weights, optional bf16 emit.
    ``step`` is an int32/uint32 device scalar holding the number of completed steps."""
    _need_cuda(master, grad, m, v, out_bf16, clip_coef, step)
    if master.dtype != torch.float32:
        raise TypeError("optim_step: master weights must be float32")
    n = master.numel()
    if grad.numel() != n or (out_bf16 is not None and out_bf16.numel() != n):
        raise ValueError("optim_step: size mismatch")
    opt = cfg.to_c()
    lib = _lib.load()
    with torch.cuda.device(master.device):
        check(lib.psb_optim_step(C.byref(opt), master.data_ptr(), _ptr(m), _ptr(v), grad.data_ptr(),
                                 _dt(grad), _ptr(out_bf16), n, _ptr(clip_coef), _ptr(step),
                                 _stream(master)), "psb_optim_step")


def counter_add_(counter: torch.Tensor, delta: int = 1) -> None:
    _need_cuda(counter)
    with torch.cuda.device(counter.device):
        check(_lib.load().psb_counter_add(counter.data_ptr(), delta, _stream(counter)),
              "psb_counter_add")


# ---------------------------------------------------------------- K4
def sumsq(x: torch.Tensor, out: Optional[torch.Tensor] = None,
          scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K4: Σx² as a 1-element fp32 tensor (deterministic two-phase reduction)."""
    _need_cuda(x, out, scratch)
    lib = _lib.load()
    need = lib.psb_sumsq_scratch_floats()
    if scratch is None:
        scratch = torch.empty(need, dtype=torch.float32, device=x.device)
    elif scratch.numel() < need or scratch.dtype != torch.float32:
        raise ValueError(f"sumsq: scratch must hold {need} float32")
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.psb_sumsq(x.data_ptr(), _dt(x), x.numel(), scratch.data_ptr(), out.data_ptr(),
                            _stream(x)), "psb_sumsq")
    return out


def clip_coef(sumsq_parts: torch.Tensor, max_norm: float, eps: float = 1e-6):
    """K4: (coef, norm) with coef = min(1, max_norm / (sqrt(Σ parts) + eps))."""
    _need_cuda(sumsq_parts)
    coef = torch.empty(1, dtype=torch.float32, device=sumsq_parts.device)
    norm = torch.empty(1, dtype=torch.float32, device=sumsq_parts.device)
    with torch.cuda.device(sumsq_parts.device):
        check(_lib.load().psb_clip_coef(sumsq_parts.data_ptr(), sumsq_parts.numel(), max_norm, eps,
                                        coef.data_ptr(), norm.data_ptr(), _stream(sumsq_parts)),
              "psb_clip_coef")
    return coef, norm


# ---------------------------------------------------------------- K2f (local operands)
def gemm_bf16(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out_dtype: torch.dtype = torch.bfloat16,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Y = X · Wᵀ (+ bias) on tcgen05 tensor cores; X [M,K], W [N,K] bf16 row-major."""
    _need_cuda(x, w, bias, out)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise TypeError("gemm_bf16: operands must be bfloat16")
    if x.dim() != 2 or w.dim() != 2 or x.shape[1] != w.shape[1]:
        raise ValueError(f"gemm_bf16: bad shapes {tuple(x.shape)} x {tuple(w.shape)}^T")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != w.shape[0]):
        raise ValueError("gemm_bf16: bias must be float32 [N]")
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    with torch.cuda.device(x.device):
        check(_lib.load().psb_gemm_bf16_tn(x.data_ptr(), w.data_ptr(), _ptr(bias), out.data_ptr(),
                                           M, N, K, _dt(out), _stream(x)), "psb_gemm_bf16_tn")
    return out


_EPILOGUES = {None: _lib.EPI_NONE, "none": _lib.EPI_NONE, "gelu": _lib.EPI_GELU,
              "dgelu": _lib.EPI_DGELU}


def gemm(a: torch.Tensor, b: torch.Tensor, *, trans_a: bool = False, trans_b: bool = False,
         bias: Optional[torch.Tensor] = None, epilogue: Optional[str] = None,
         aux: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16,
         out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``Y[M,N] = op(A) · op(B) (+ bias)`` on the CTA-pair tcgen05 kernel, no transposed copies.

    ``trans_a``: A is stored ``[K,M]`` (else ``[M,K]``); ``trans_b``: B is stored ``[K,N]``
    (else ``[N,K]``, i.e. a Linear weight).  Operands may be column blocks of wider matrices
    (``t[:, lo:hi]``, lo a multiple of 8).  ``epilogue="gelu"`` returns ``gelu(Z)`` and
    writes ``Z = acc + bias`` (bf16) into ``aux`` when given; ``epilogue="dgelu"`` returns
    ``acc * gelu'(aux)``.  (include/psb_kernels.h: psb_gemm_bf16)
    """
    _need_cuda(bias, out, aux)
    if not (a.is_cuda and b.is_cuda):
        raise TypeError("native ops need CUDA tensors; use ops.reference on the host")
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise TypeError("gemm: operands must be bfloat16")
    if a.dim() != 2 or b.dim() != 2 or a.stride(1) != 1 or b.stride(1) != 1:
        raise ValueError("gemm: operands must be 2-D with unit stride along their rows")
    K, M = (a.shape if trans_a else a.shape[::-1])
    Kb, N = (b.shape if trans_b else b.shape[::-1])
    if K != Kb:
        raise ValueError(f"gemm: reduction sizes differ ({K} vs {Kb})")
    if bias is not None and (bias.dtype not in (torch.float32, torch.bfloat16) or bias.numel() != N):
        raise ValueError("gemm: bias must be float32 or bfloat16 [N]")
    epi = _EPILOGUES[epilogue]
    if aux is not None and (aux.dtype != torch.bfloat16 or tuple(aux.shape) != (M, N)
                            or not aux.is_contiguous()):
        raise ValueError("gemm: aux must be a contiguous bfloat16 [M,N] tensor")
    if epi == _lib.EPI_DGELU and aux is None:
        raise ValueError("gemm: epilogue='dgelu' needs aux (the pre-activation)")
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    elif tuple(out.shape) != (M, N) or not out.is_contiguous():
        raise ValueError("gemm: out must be a contiguous [M,N] tensor")
    g = _lib.PsbGemm()
    g.a, g.b, g.y, g.bias = a.data_ptr(), b.data_ptr(), out.data_ptr(), _ptr(bias)
    g.aux_in = aux.data_ptr() if epi == _lib.EPI_DGELU else None
    g.aux_out = aux.data_ptr() if (epi == _lib.EPI_GELU and aux is not None) else None
    g.M, g.N, g.K = M, N, K
    g.trans_a, g.trans_b = int(trans_a), int(trans_b)
    g.out_dtype, g.epilogue = _dt(out), epi
    g.bias_dtype = _dt(bias) if bias is not None else F32
    # a column block of a wider matrix (row stride > row length) is used in place
    g.lda = 0 if a.is_contiguous() else a.stride(0)
    g.ldb = 0 if b.is_contiguous() else b.stride(0)
    with torch.cuda.device(a.device):
        check(_lib.load().psb_gemm_bf16(C.byref(g), _stream(a)), "psb_gemm_bf16")
    return out


# ---------------------------------------------------------------- training-step helpers
_SCRATCH: dict = {}


def _scratch(key, nbytes: int, device) -> torch.Tensor:
    """Zero-initialised per-device scratch that the kernels leave clean (arrival counters)."""
    t = _SCRATCH.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _SCRATCH[key] = t
    return t


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None,
           out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """``x.sum(0)`` of a contiguous bf16 ``[M,N]`` matrix (the bias gradient), fp32 accumulation
    in a fixed order.  (include/psb_kernels.h: psb_colsum_bf16)"""
    _need_cuda(x, out)
    if x.dtype != torch.bfloat16 or x.dim() != 2 or not x.is_contiguous():
        raise ValueError("colsum: x must be a contiguous bfloat16 [M,N] tensor")
    M, N = x.shape
    if out is None:
        out = torch.empty(N, dtype=out_dtype, device=x.device)
    elif out.numel() != N or not out.is_contiguous():
        raise ValueError("colsum: out must be a contiguous tensor of N elements")
    lib = _lib.load()
    stream = _stream(x)
    scratch = _scratch(("colsum", x.device.index, stream), lib.psb_colsum_scratch_bytes(M, N), x.device)
    with torch.cuda.device(x.device):
        check(lib.psb_colsum_bf16(x.data_ptr(), out.data_ptr(), _dt(out), M, N, scratch.data_ptr(),
                                  stream), "psb_colsum_bf16")
    return out


def mse_fwd(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``mean((a-b)**2)`` of two contiguous bf16 tensors as an fp32 scalar, one kernel."""
    _need_cuda(a, b)
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16 or a.shape != b.shape \
            or not a.is_contiguous() or not b.is_contiguous():
        raise ValueError("mse_fwd: need two contiguous bfloat16 tensors of one shape")
    loss = torch.empty((), dtype=torch.float32, device=a.device)
    stream = _stream(a)
    scratch = _scratch(("mse", a.device.index, stream), 4 * 304, a.device)
    with torch.cuda.device(a.device):
        check(_lib.load().psb_mse_fwd_bf16(a.data_ptr(), b.data_ptr(), a.numel(), scratch.data_ptr(),
                                           loss.data_ptr(), stream), "psb_mse_fwd_bf16")
    return loss


def mse_bwd(a: torch.Tensor, b: torch.Tensor, gout: Optional[torch.Tensor]) -> torch.Tensor:
    """Gradient of :func:`mse_fwd` w.r.t. ``a``: ``2 (a-b) / n * gout`` in bf16."""
    _need_cuda(a, b, gout)
    if gout is not None and (gout.dtype != torch.float32 or gout.numel() != 1):
        raise ValueError("mse_bwd: gout must be a float32 scalar")
    da = torch.empty_like(a)
    with torch.cuda.device(a.device):
        check(_lib.load().psb_mse_bwd_bf16(a.data_ptr(), b.data_ptr(), _ptr(gout), a.numel(),
                                           da.data_ptr(), _stream(a)), "psb_mse_bwd_bf16")
    return da
