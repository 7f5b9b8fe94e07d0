"""AsyncParameterServer — asynchronous and bounded-staleness (SSP) training on versioned
shard buffers.

The reference is BSP only; its wire schema merely carries the fields SSP needs
(``PullRequest.iteration`` is ignored by the server, src/parameter_server.cpp:93-97;
``ParameterUpdate.ready`` is ignored by the worker, src/worker.cpp:249-251).  On the host path
those fields now drive ``parameter_server --mode=async|ssp``.  This module is the HBM path:

* every rank is a worker with its own **clock** and owns one slice of every bucket;
* gradients of clock ``c`` are staged into slot ``c % ring`` of a per-worker gradient **ring**
  in symmetric memory and announced with ``ready[b][w] = c`` — nobody waits for anybody;
* at each of its own steps an owner applies, per bucket, the oldest unapplied clock of every
  worker that has one (``psb_async_plan`` → K1 with ``MASKED | SEQLOCK``), each scaled by 1/N,
  and acknowledges the consumed clocks so the ring slots can be reused;
* parameter slices are **versioned**: odd while an update is in flight, even when stable; the
  pull (K2 with ``seqlock``) copies a slice between two reads of its version and retries on
  change, so a worker never trains on a torn slice;
* the gate before clock ``c`` enforces (a) ring-slot reuse safety — a worker runs at most
  ``ring`` clocks ahead of the slowest owner — and (b) in SSP mode the staleness bound
  ``c - 1 - min_w clock[w] <= staleness``.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from .. import _lib
from .._lib import PsbAsync, check
from ..ops import native
from ..ops import reference as ref
from ..ops.native import OptimConfig
from .engine import ShardedParameterServer


class AsyncParameterServer(ShardedParameterServer):
    def __init__(self, named_params: Sequence[Tuple[str, torch.nn.Parameter]], optim: OptimConfig,
                 fabric=None, *, world: Optional[int] = None, rank: Optional[int] = None,
                 group=None, bucket_mb: float = 64.0, staleness: Optional[int] = None,
                 ring: Optional[int] = None):
        """``staleness=None``: fully asynchronous (only the ring bounds run-ahead);
        ``staleness=s >= 0``: stale-synchronous with bound ``s`` (``s = 0`` is lock-step)."""
        self.staleness = -1 if staleness is None else int(staleness)
        self.ring = ring if ring is not None else max(2, self.staleness + 1, 2 if staleness is not None else 4)
        if not 1 <= self.ring <= 8:
            raise ValueError("ring must be in [1, 8]")
        super().__init__(named_params, optim, fabric, world=world, rank=rank, group=group,
                         bucket_mb=bucket_mb, broadcast=False, clip_norm=None, overlap=False,
                         grad_ring=self.ring)
        self.clock = 0
        if self.is_cuda:
            be = self.backend
            nb = len(self.plan.buckets)
            dev = self.device
            self._applied = torch.zeros(nb * _lib.MAX_PEERS, dtype=torch.int32, device=dev)
            self._masks = torch.zeros(nb, dtype=torch.int32, device=dev)
            self._grad_off = torch.zeros(nb * _lib.MAX_PEERS, dtype=torch.int64, device=dev)
            a = PsbAsync()
            a.fabric = fabric.c_fabric()
            a.num_buckets = nb
            a.ring = self.ring
            a.ring_stride_bytes = be.ring_stride_bytes
            a.applied = self._applied.data_ptr()
            a.masks = self._masks.data_ptr()
            a.grad_off = self._grad_off.data_ptr()
            a.epoch = be.epoch.data_ptr()
            self._args = a
            self.lib = _lib.load()
        else:
            self._host_applied = 0

    # hooks only collect; everything else happens in step()
    def _make_hook(self, name: str):
        slot = self.plan.slots[name]

        def hook(p: torch.nn.Parameter):
            g = p.grad
            p.grad = None
            self._pending[slot.bucket].append((slot, g))

        return hook

    def _s(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _owner_round(self) -> None:
        """Apply what is pending on this owner, acknowledge, refresh the local replica."""
        be = self.backend
        a = self._args
        check(self.lib.psb_async_plan(C.byref(a), self._s()), "psb_async_plan")
        for b in range(len(self.plan.buckets)):
            be.push(b, masked=True, seqlock=True, no_wait=True,
                    mask_ptr=self._masks[b:].data_ptr(),
                    grad_off_ptr=self._grad_off[b * _lib.MAX_PEERS:].data_ptr())
        check(self.lib.psb_async_ack(C.byref(a), self._s()), "psb_async_ack")
        be.advance_step()
        for b in range(len(self.plan.buckets)):
            be.pull(b, seqlock=True)

    def step(self) -> None:
        clock = self.clock + 1
        be = self.backend
        nb = len(self.plan.buckets)
        if not self.is_cuda:
            return self._host_step(clock)
        with torch.cuda.device(self.device):
            # gate: blocks the stream (not the host) until clock `clock` may start — its ring
            # slot is free on every owner and, in SSP mode, nobody is too far behind
            check(self.lib.psb_async_gate(C.byref(self._args), clock, self.staleness, self._s()),
                  "psb_async_gate")
            slot = be.grad_slot(clock % self.ring)
            for b, bk in enumerate(self.plan.buckets):
                pend = self._pending[b]
                bucket = slot[bk.base:bk.base + bk.numel]
                if len(pend) < len(bk.slots):
                    bucket.zero_()
                native.gather_to_bucket([self._flat_grad(s, g) for s, g in pend], bucket,
                                        [s.offset for s, _ in pend])
                self._pending[b] = []
            check(self.lib.psb_async_signal(C.byref(self._args), clock, self._s()),
                  "psb_async_signal")
            self._owner_round()
        self.clock = clock
        self.steps_done += 1

    def drain(self, poll_s: float = 0.002) -> None:
        """Collective, after the last step: keep serving as an owner until every worker has
        finished and all of its gradients are applied here, then refresh the replica — so all
        ranks end with the same parameters and each gradient was applied exactly once.  A rank
        that finishes early must not stop consuming: slower peers still wait for its acks."""
        if not self.is_cuda:
            return
        import time
        with torch.cuda.device(self.device):
            fab = self.fabric.c_fabric()
            check(self.lib.psb_pad_post(C.byref(fab), _lib.PAD_DONE_OFF, self.clock, self._s()),
                  "psb_pad_post")
            pad = self.fabric.pad.view(torch.int32)
            while True:
                self._owner_round()
                torch.cuda.synchronize(self.device)
                done = pad[_lib.PAD_DONE_OFF:_lib.PAD_DONE_OFF + self.world]
                if bool((done > 0).all()):
                    applied = self._applied.view(-1, _lib.MAX_PEERS)[:, :self.world]
                    if bool((applied >= done.unsqueeze(0)).all()):
                        break
                time.sleep(poll_s)
            self.fabric.barrier(4)  # every owner has applied everything
            for b in range(len(self.plan.buckets)):
                self.backend.pull(b, seqlock=True)
            self.fabric.barrier(4)
            torch.cuda.synchronize(self.device)

    def applied_clocks(self) -> torch.Tensor:
        """[num_buckets, world] last applied clock per worker on this owner."""
        return self._applied.view(-1, _lib.MAX_PEERS)[:, :self.world].clone()

    # ------------------------------------------------------------------ host emulation
    def _host_step(self, clock: int) -> None:
        """CPU tensors / gloo: collectives are blocking, so every worker's gradient of this
        clock arrives at once; owners apply them one after the other in rank order, each scaled
        by 1/N — the arithmetic of the device path when nobody is ahead."""
        be = self.backend
        cfg = OptimConfig(**{**self.optim.__dict__, "grad_scale": 1.0})
        for b, bk in enumerate(self.plan.buckets):
            bucket = be.grad_bucket(b)
            bucket.zero_()
            for s, g in self._pending[b]:
                bucket[s.offset:s.offset + s.numel].copy_(self._flat_grad(s, g).reshape(-1))
            self._pending[b] = []
            contribs = be._gather_slices(bucket)
            n = bk.slice_elems
            for w, c in enumerate(contribs):
                g = c[self.rank * n:(self.rank + 1) * n].to(torch.float32) * self.optim.grad_scale
                ref.optim_step(cfg, be.shard_slice(be.master, b), g,
                               be.shard_slice(be.m, b) if be.m is not None else None,
                               be.shard_slice(be.v, b) if be.v is not None else None,
                               step=int(be.step_word.item()))
            mine = be.shard_slice(be.master, b).to(be.param_flat.dtype)
            pb = be.param_bucket(b)
            for r, sl in enumerate(be._gather_slices(mine)):
                pb[r * n:(r + 1) * n].copy_(sl)
            be.epoch[b] += 1
        be.advance_step()
        self.clock = clock
        self.steps_done += 1
