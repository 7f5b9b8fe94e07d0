"""CUDA backend of the sharded parameter server: symmetric buffers + fused kernels.

Holds, per rank: the symmetric gradient space ``G`` and bf16 parameter replica ``P`` (both
peer-mapped), the local shard state (fp32 master, m, v), per-bucket epoch / completion words
and the pre-filled ``PsbPush`` / ``PsbPull`` argument blocks.  All launches go to
``libpsb_kernels.so`` (K1, K2, K9 of SURVEY.md §8); nothing here calls NCCL.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from .. import _lib
from .._lib import PsbPull, PsbPush, check
from ..ops.native import OptimConfig
from .buckets import BucketPlan
from .fabric import Fabric


class DeviceBackend:
    def __init__(self, fabric: Fabric, plan: BucketPlan, optim: OptimConfig,
                 grad_dtype: torch.dtype = torch.bfloat16, broadcast: bool = True,
                 clip: bool = False, push_grid: int = 0, grad_ring: int = 1,
                 push_tma: bool = False, alt_params: bool = False):
        if len(plan.buckets) > _lib.MAX_BUCKETS:
            raise ValueError(f"{len(plan.buckets)} buckets exceed MAX_BUCKETS={_lib.MAX_BUCKETS}")
        self.fabric, self.plan, self.optim = fabric, plan, optim
        self.lib = _lib.load()
        self.device = fabric.device
        self.world, self.rank = fabric.world, fabric.rank
        self.grad_dtype = grad_dtype
        self.broadcast = broadcast
        self.push_grid = push_grid
        self.push_tma = push_tma
        import os
        self.split_wait = os.environ.get("PSB_SPLIT_WAIT", "1") != "0"
        self._emulated = not fabric._owns  # ranks emulated in one process: launches are serial
        fabric.reset_flags()  # epoch words below start at 0: drop a previous engine's flags
        gsz = 2 if grad_dtype == torch.bfloat16 else 4
        total = plan.total_elems
        self.grad_ring = grad_ring
        self.ring_stride_bytes = (total * gsz + 4095) // 4096 * 4096
        self.G = fabric.alloc(self.ring_stride_bytes * grad_ring, align=4096)
        self.P = fabric.alloc(total * 2, align=4096)
        self.grad_flat = self.G.view(grad_dtype)[:total]
        self.param_flat = self.P.view(torch.bfloat16)
        # second replica: a push may write the new parameters of a bucket while kernels of the
        # same iteration still read the old ones (the engine swaps the two after every step)
        self.P2 = fabric.alloc(total * 2, align=4096) if alt_params else None
        self.param_flat2 = self.P2.view(torch.bfloat16) if alt_params else None
        # The heap hands out ranges that an earlier engine may have used (free list): the padding
        # between tensors and behind the last one is reduced, normed and updated like everything
        # else, so it must start at zero — stale parameter values there once showed up as a
        # gradient norm of 35 instead of 0.02.
        for buf in (self.G, self.P, self.P2):
            if buf is not None:
                buf.local.zero_()
        shard = plan.shard_elems
        dev = self.device
        self.master = torch.zeros(shard, dtype=torch.float32, device=dev)
        self.m = torch.zeros(shard, dtype=torch.float32, device=dev) if optim.needs_m() else None
        self.v = torch.zeros(shard, dtype=torch.float32, device=dev) if optim.needs_v() else None
        self.reduced = torch.zeros(shard, dtype=torch.float32, device=dev) if clip else None
        nb = len(plan.buckets)
        self.epoch = torch.zeros(nb, dtype=torch.int32, device=dev)
        self.done = torch.zeros(nb, dtype=torch.int32, device=dev)
        self.step_word = torch.zeros(1, dtype=torch.int32, device=dev)
        self.sumsq_acc = torch.zeros(1, dtype=torch.float32, device=dev)
        self.clip_coef = torch.ones(1, dtype=torch.float32, device=dev)
        self.grad_norm = torch.zeros(1, dtype=torch.float32, device=dev)
        self.norm_seq = torch.zeros(1, dtype=torch.int32, device=dev)
        self.peer_mask = torch.zeros(1, dtype=torch.int32, device=dev)
        self._push_args: List[PsbPush] = [self._make_push(b) for b in range(nb)]
        self._push_args_alt: List[PsbPush] = (
            [self._make_push(b, self.P2) for b in range(nb)] if alt_params else [])
        self._pull_args: List[PsbPull] = [self._make_pull(b) for b in range(nb)]

    def close(self) -> None:
        """Return G / P / P2 to the symmetric heap (in reverse order of allocation)."""
        for name in ("P2", "P", "G"):
            buf = getattr(self, name, None)
            if buf is not None:
                buf.free()
                setattr(self, name, None)
        self.param_flat = self.param_flat2 = self.grad_flat = None

    # ------------------------------------------------------------------ argument blocks
    def _make_push(self, b: int, P=None) -> PsbPush:
        P = self.P if P is None else P
        bk = self.plan.buckets[b]
        gsz = 2 if self.grad_dtype == torch.bfloat16 else 4
        a = PsbPush()
        a.fabric = self.fabric.c_fabric()
        for r in range(self.world):
            a.grad[r] = self.G.peer_ptr(r) + bk.base * gsz
            a.param[r] = P.peer_ptr(r) + bk.base * 2
        # NVLS: one multicast address per buffer names every rank's copy (0 without multicast)
        a.grad_mc = (self.G.mc_ptr + bk.base * gsz) if self.G.mc_ptr else None
        a.param_mc = (P.mc_ptr + bk.base * 2) if P.mc_ptr else None
        sb = bk.shard_base(self.world)
        a.master = self.master[sb:].data_ptr()
        a.m = self.m[sb:].data_ptr() if self.m is not None else None
        a.v = self.v[sb:].data_ptr() if self.v is not None else None
        a.reduced = self.reduced[sb:].data_ptr() if self.reduced is not None else None
        a.sumsq_acc = self.sumsq_acc.data_ptr()
        a.slice_elems = bk.slice_elems
        a.bucket = b
        a.grad_dtype = _lib.BF16 if self.grad_dtype == torch.bfloat16 else _lib.F32
        a.epoch = self.epoch[b:].data_ptr()
        a.step = self.step_word.data_ptr()
        a.clip_coef = None
        a.peer_mask = self.peer_mask.data_ptr()
        a.done_counter = self.done[b:].data_ptr()
        a.flags = 0
        a.grid = self.push_grid
        a.opt = self.optim.to_c()
        return a

    def _make_pull(self, b: int) -> PsbPull:
        bk = self.plan.buckets[b]
        a = PsbPull()
        a.fabric = self.fabric.c_fabric()
        for r in range(self.world):
            a.param[r] = self.P.peer_ptr(r) + bk.base * 2
        a.slice_bytes = bk.slice_elems * 2
        a.bucket = b
        a.epoch = self.epoch[b:].data_ptr()
        a.grid = 0
        a.skip_wait = 0
        return a

    def grad_slot(self, k: int) -> torch.Tensor:
        """Flat gradient space of ring slot ``k`` (async mode keeps ``grad_ring`` clocks alive)."""
        gsz = 2 if self.grad_dtype == torch.bfloat16 else 4
        lo = k * self.ring_stride_bytes // gsz
        return self.G.view(self.grad_dtype)[lo:lo + self.plan.total_elems]

    def set_optim(self, optim: OptimConfig) -> None:
        """Change hyper-parameters (e.g. lr schedule); buffers stay."""
        self.optim = optim
        for a in self._push_args + self._push_args_alt:
            a.opt = optim.to_c()

    # ------------------------------------------------------------------ views
    def grad_bucket(self, b: int) -> torch.Tensor:
        bk = self.plan.buckets[b]
        return self.grad_flat[bk.base:bk.base + bk.numel]

    def param_bucket(self, b: int) -> torch.Tensor:
        bk = self.plan.buckets[b]
        return self.param_flat[bk.base:bk.base + bk.numel]

    def shard_slice(self, t: torch.Tensor, b: int) -> torch.Tensor:
        bk = self.plan.buckets[b]
        sb = bk.shard_base(self.world)
        return t[sb:sb + bk.slice_elems]

    # ------------------------------------------------------------------ launches
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def push(self, b: int, *, reduce_only: bool = False, from_reduced: bool = False,
             no_wait: bool = False, masked: bool = False, clip: bool = False,
             seqlock: bool = False, mask_ptr: Optional[int] = None,
             grad_off_ptr: Optional[int] = None, broadcast: Optional[bool] = None,
             tma: Optional[bool] = None, grid: Optional[int] = None, block: int = 0,
             alt: bool = False, nvls: bool = False) -> None:
        """K1 on the current stream.  ``block``: threads per CTA of the register-path kernel
        (0 = 512; 256 shares an SM with a tcgen05 GEMM CTA).  ``alt``: the new parameters go to
        the second replica.  ``nvls``: the switch reduces and replicates (push_nvls.cu)."""
        a = self._push_args_alt[b] if alt else self._push_args[b]
        bc = self.broadcast if broadcast is None else broadcast
        flags = _lib.PUSH_BROADCAST if bc else 0
        if seqlock:
            flags |= _lib.PUSH_SEQLOCK
        if nvls:
            if not a.grad_mc:
                raise RuntimeError("NVLS push needs a VMM heap bound to a multicast object")
            flags |= _lib.PUSH_NVLS
        elif self.push_tma if tma is None else tma:
            flags |= _lib.PUSH_TMA
        a.grid = self.push_grid if grid is None else grid
        a.peer_mask = mask_ptr if mask_ptr is not None else self.peer_mask.data_ptr()
        a.grad_off = grad_off_ptr
        a.block = block
        if reduce_only:
            flags |= _lib.PUSH_REDUCE_ONLY
        if from_reduced:
            flags |= _lib.PUSH_FROM_REDUCED
        if no_wait:
            flags |= _lib.PUSH_NO_WAIT
        if masked:
            flags |= _lib.PUSH_MASKED
        # READY handshake as a one-warp kernel of its own in front of the data kernel, so that the
        # data kernel's CTAs never hold SMs while the slowest rank is still producing gradients
        ahead = (self.split_wait and self.world > 1 and not self._emulated
                 and not (no_wait or masked or from_reduced))
        if ahead:
            flags |= _lib.PUSH_NO_WAIT
        a.flags = flags
        a.clip_coef = self.clip_coef.data_ptr() if clip else None
        with torch.cuda.device(self.device):
            if ahead:
                check(self.lib.psb_push_ready(C.byref(a), self._stream()), "psb_push_ready")
            check(self.lib.psb_push_fused(C.byref(a), self._stream()), "psb_push_fused")

    def pull(self, b: int, skip_wait: bool = False, seqlock: bool = False) -> None:
        """K2 on the current stream."""
        a = self._pull_args[b]
        a.skip_wait = int(skip_wait)
        a.seqlock = int(seqlock)
        with torch.cuda.device(self.device):
            check(self.lib.psb_pull_allgather(C.byref(a), self._stream()), "psb_pull_allgather")

    @property
    def nvls_available(self) -> bool:
        return bool(self.G.mc_ptr) and self.grad_dtype == torch.bfloat16 and self.grad_ring == 1

    def bcast_slice_nvls(self, b: int, publish: bool = False, grid: int = 0) -> None:
        """K2 over NVLS: this rank's slice of bucket ``b`` goes into every replica with
        multimem.st (push-mode all-gather)."""
        bk = self.plan.buckets[b]
        fab = self.fabric.c_fabric()
        with torch.cuda.device(self.device):
            check(self.lib.psb_bcast_slice_nvls(
                C.byref(fab), self.P.peer_ptr(self.rank) + bk.base * 2,
                self.P.mc_ptr + bk.base * 2, bk.slice_elems * 2, b,
                self.epoch[b:].data_ptr() if publish else None, self.done[b:].data_ptr(), grid,
                self._stream()), "psb_bcast_slice_nvls")

    def wait_versions(self, lo: int = 0, hi: Optional[int] = None) -> None:
        """K9: block the current stream until every owner published buckets [lo, hi)."""
        hi = len(self.plan.buckets) if hi is None else hi
        fab = self.fabric.c_fabric()
        with torch.cuda.device(self.device):
            check(self.lib.psb_wait_versions(C.byref(fab), self.epoch.data_ptr(), lo, hi,
                                             self._stream()), "psb_wait_versions")

    def norm_allreduce(self, max_norm: float, eps: float = 1e-6) -> None:
        """K4: cross-shard Σg² through the signal pads, clip coefficient into ``clip_coef``."""
        fab = self.fabric.c_fabric()
        with torch.cuda.device(self.device):
            check(self.lib.psb_norm_allreduce(C.byref(fab), self.sumsq_acc.data_ptr(),
                                              self.norm_seq.data_ptr(), max_norm, eps,
                                              self.clip_coef.data_ptr(), self.grad_norm.data_ptr(),
                                              self._stream()), "psb_norm_allreduce")

    def advance_step(self) -> None:
        with torch.cuda.device(self.device):
            check(self.lib.psb_counter_add(self.step_word.data_ptr(), 1, self._stream()),
                  "psb_counter_add")
