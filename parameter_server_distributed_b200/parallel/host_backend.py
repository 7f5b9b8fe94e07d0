"""Host (CPU tensor) backend of the sharded parameter server.

Same bucket / shard layout and the same update math as the CUDA backend, but built from
``torch.distributed`` (``gloo``) collectives and the fp32 PyTorch oracles in
:mod:`..ops.reference`.  It exists so the engine logic — bucket hooks, shard ownership,
modes, checkpoints — is testable without a GPU (world_size > 1 over gloo), and as the
numerical oracle of the fused kernels.  It is never selected for CUDA tensors.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import reference as ref
from ..ops.native import OptimConfig
from .buckets import BucketPlan


class HostBackend:
    def __init__(self, plan: BucketPlan, optim: OptimConfig, rank: int = 0, world: int = 1,
                 group=None, param_dtype: torch.dtype = torch.float32,
                 grad_dtype: torch.dtype = torch.float32, broadcast: bool = True,
                 clip: bool = False, alt_params: bool = False):
        self.plan, self.optim = plan, optim
        self.rank, self.world, self.group = rank, world, group
        self.device = torch.device("cpu")
        self.broadcast = broadcast
        self.grad_dtype = grad_dtype
        total, shard = plan.total_elems, plan.shard_elems
        self.grad_flat = torch.zeros(total, dtype=grad_dtype)
        self.param_flat = torch.zeros(total, dtype=param_dtype)
        # second replica (see DeviceBackend): pushes with alt=True write here
        self.param_flat2 = torch.zeros(total, dtype=param_dtype) if alt_params else None
        self.master = torch.zeros(shard, dtype=torch.float32)
        self.m = torch.zeros(shard, dtype=torch.float32) if optim.needs_m() else None
        self.v = torch.zeros(shard, dtype=torch.float32) if optim.needs_v() else None
        self.reduced = torch.zeros(shard, dtype=torch.float32) if clip else None
        self.epoch = torch.zeros(len(plan.buckets), dtype=torch.int32)
        self.step_word = torch.zeros(1, dtype=torch.int32)
        self.sumsq_acc = torch.zeros(1, dtype=torch.float32)
        self.clip_coef = torch.ones(1, dtype=torch.float32)
        self.grad_norm = torch.zeros(1, dtype=torch.float32)
        self.peer_mask = torch.zeros(1, dtype=torch.int32)

    def set_optim(self, optim: OptimConfig) -> None:
        self.optim = optim

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

    # ------------------------------------------------------------------ collectives
    def _gather_slices(self, mine: torch.Tensor) -> List[torch.Tensor]:
        if self.world == 1:
            return [mine]
        out = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(out, mine.contiguous(), group=self.group)
        return out

    def push(self, b: int, *, reduce_only: bool = False, from_reduced: bool = False,
             no_wait: bool = False, masked: bool = False, clip: bool = False,
             broadcast=None, alt: bool = False, **_unused) -> None:
        bk = self.plan.buckets[b]
        s = bk.slice_elems
        if from_reduced:
            g = self.shard_slice(self.reduced, b).clone()
        else:
            # every rank contributes its full bucket; the owner sums its slice in rank order
            contribs = self._gather_slices(self.grad_bucket(b))
            mask = int(self.peer_mask.item()) if masked else -1
            acc = torch.zeros(s, dtype=torch.float32)
            for w, c in enumerate(contribs):
                if (mask >> w) & 1:
                    acc += c[self.rank * s:(self.rank + 1) * s].to(torch.float32)
            g = acc * self.optim.grad_scale
        if reduce_only:
            self.shard_slice(self.reduced, b).copy_(g)
            self.sumsq_acc += ref.sumsq(g)
            return
        cfg = OptimConfig(**{**self.optim.__dict__, "grad_scale": 1.0})
        ref.optim_step(cfg, self.shard_slice(self.master, b), g,
                       self.shard_slice(self.m, b) if self.m is not None else None,
                       self.shard_slice(self.v, b) if self.v is not None else None,
                       clip_coef=float(self.clip_coef.item()) if clip else 1.0,
                       step=int(self.step_word.item()))
        mine = self.shard_slice(self.master, b).to(self.param_flat.dtype)
        pb = (self.param_flat2[bk.base:bk.base + bk.numel] if alt else self.param_bucket(b))
        pb[self.rank * s:(self.rank + 1) * s].copy_(mine)
        if self.broadcast if broadcast is None else broadcast:
            for r, sl in enumerate(self._gather_slices(mine)):
                pb[r * s:(r + 1) * s].copy_(sl)
        self.epoch[b] += 1

    def pull(self, b: int, skip_wait: bool = False) -> None:
        bk = self.plan.buckets[b]
        s = bk.slice_elems
        pb = self.param_bucket(b)
        mine = pb[self.rank * s:(self.rank + 1) * s].clone()
        for r, sl in enumerate(self._gather_slices(mine)):
            pb[r * s:(r + 1) * s].copy_(sl)

    def wait_versions(self, lo: int = 0, hi: Optional[int] = None) -> None:
        return  # collectives above are blocking

    def norm_allreduce(self, max_norm: float, eps: float = 1e-6) -> None:
        total = self.sumsq_acc.clone()
        if self.world > 1:
            parts = [torch.empty_like(total) for _ in range(self.world)]
            dist.all_gather(parts, total, group=self.group)
            total = torch.stack(parts).sum(0)
        self.grad_norm.copy_(total.sqrt())
        coef = ref.clip_coef(float(total.item()), max_norm, eps) if max_norm > 0 else 1.0
        self.clip_coef.fill_(coef)
        self.sumsq_acc.zero_()

    def advance_step(self) -> None:
        self.step_word += 1
