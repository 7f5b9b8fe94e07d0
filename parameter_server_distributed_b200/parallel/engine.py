"""ShardedParameterServer — worker + parameter-server shard of one rank.

What the reference spreads over three processes and two gRPC hops per iteration
(``pull -> compute -> push -> wait-for-sync``, src/worker.cpp:331-406, aggregation in
src/parameter_server.cpp:18-91) happens here inside one process per GPU:

* the model's parameters are views of the local bf16 replica ``P``;
* backward hooks copy each finished bucket of gradients into the symmetric space ``G`` and
  enqueue the fused push (K1) on a high-priority communication stream, so the push of
  bucket *b* overlaps the backward of the layers in front of it;
* K1 on rank *r* reduces slice *r* over all workers, updates the fp32 master shard and writes
  the new bf16 slice into every replica (push-mode all-gather) or only locally (pull mode,
  followed by K2 / the fused all-gather GEMM);
* ``step()`` closes the iteration: leftover buckets, optional global-norm clipping, the
  version wait that replaces the reference's CheckSyncStatus polling.
"""
from __future__ import annotations

import os

from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import torch

from ..ops import native
from ..ops.native import OptimConfig
from ..utils.tracing import Metrics, nvtx_range
from .buckets import BucketPlan, ParamSlot, plan_buckets


class GradSink:
    """Where a layer may write one parameter's gradient itself (attribute ``_psb_sink`` of the
    parameter).  ``view()`` is the parameter-shaped window of the gradient bucket, ``commit()``
    tells the engine that the kernel producing it is queued on the current stream; the layer
    then returns ``None`` as that parameter's gradient so autograd does not deliver it twice."""

    __slots__ = ("ps", "slot", "uses")

    def __init__(self, ps: "ShardedParameterServer", slot: ParamSlot):
        self.ps, self.slot = ps, slot
        self.uses = 0  # forward passes that took this sink since the last step()

    def acquire(self) -> "GradSink":
        """Called by a layer's forward.  A parameter used by more than one layer call in a step
        (tied weights) needs its gradients ADDED, which autograd does: `exclusive` is then
        false and the layers return their gradients the ordinary way."""
        self.uses += 1
        return self

    @property
    def exclusive(self) -> bool:
        return self.uses == 1

    def view(self) -> torch.Tensor:
        s = self.slot
        if s.pieces:
            raise RuntimeError(f"{s.name} is split over {len(s.pieces)} buckets: use parts()")
        return self.ps.backend.grad_bucket(s.bucket)[s.offset:s.offset + s.numel].view(s.shape)

    def commit(self) -> None:
        if self.slot.pieces:
            raise RuntimeError(f"{self.slot.name} is split: commit its parts")
        self.ps._arrive(self.slot, None)

    def parts(self):
        """``[(row_lo, row_hi, view, commit)]``: the row blocks in which the gradient may be
        produced, each with its window of a gradient bucket and its own commit.  A tensor that
        the plan did not split has one part covering all rows.  Committing part *i* before
        computing part *i+1* lets the push of *i* overlap that computation."""
        s = self.slot
        if not s.pieces:
            return [(0, s.shape[0] if s.shape else 1, self.view(), self.commit)]
        be = self.ps.backend
        return [(p.row_lo, p.row_hi,
                 be.grad_bucket(p.bucket)[p.offset:p.offset + p.numel].view(p.shape),
                 (lambda p=p: self.ps._arrive(p, None))) for p in s.pieces]

    @property
    def early(self) -> bool:
        """May the layer commit BEFORE it has queued the kernels that still read the parameter
        (its dX GEMM)?  True when the push of this bucket cannot touch the replica those kernels
        read: it writes the other replica, or (clipping) it only reduces."""
        return self.ps._early_ok(self.slot.bucket)


class ShardedParameterServer:
    def __init__(self, named_params: Sequence[Tuple[str, torch.nn.Parameter]], optim: OptimConfig,
                 fabric=None, *, world: Optional[int] = None, rank: Optional[int] = None,
                 group=None, bucket_mb: float = 64.0, broadcast: bool = True,
                 clip_norm: Optional[float] = None, dedicated: Iterable[str] = (),
                 overlap: bool = True, push_grid: int = 0, sync_init: bool = True,
                 grad_ring: int = 1, push_tma: Optional[bool] = None, lazy_dedicated: bool = False,
                 overlap_grid: int = 32, alt_params: Optional[bool] = None,
                 split_mb: Optional[float] = None, push_nvls: Optional[bool] = None):
        """``broadcast``: owners write new slices into every replica inside K1 (push-mode
        all-gather).  ``lazy_dedicated``: buckets named in ``dedicated`` are the exception — their
        owners publish locally only and the consumer gathers them itself (the fused all-gather
        GEMM of the first layer), which halves the un-overlappable tail of the last push.
        ``alt_params`` (None = on for the synchronous CUDA engine with more than one rank): keep
        two parameter replicas and alternate between them, so that a bucket can be pushed as
        soon as its gradients exist — before the layer's dX GEMM, which still reads the old
        weights — instead of after it.
        ``push_nvls`` (``PSB_PUSH_NVLS``): reduce and replicate inside the NVSwitch
        (multimem.ld_reduce / multimem.st) where the heap is bound to a multicast object.  The
        switch rounds the sum to bf16 once (as a bf16 NCCL all-reduce does), so results match the
        peer-load kernels to bf16 tolerance rather than bit for bit.  None = automatic: used from
        4 ranks up, where it moves S bytes per GPU and direction instead of 1.75 S (measured at
        8 GPUs: 64 MB bucket 0.19 vs 0.23 ms, 256 MB 0.60 vs 0.84 ms); ``False`` keeps the
        deterministic peer-load kernels.
        ``split_mb`` (None = 16 for the synchronous engine from 4 ranks up, 0 = off;
        ``PSB_SPLIT_MB``): the weight matrices whose gradients come LAST in backward — nothing is
        left to hide their push behind — are cut into row blocks of about this size, one bucket
        each; a layer that uses ``GradSink.parts()`` then computes dW block by block and the push
        of a block overlaps the GEMM of the next, so only the last block's push is exposed."""
        named_params = [(n, p) for n, p in named_params if p.requires_grad]
        if not named_params:
            raise ValueError("no trainable parameters")
        self.named_params = named_params
        dev = named_params[0][1].device
        self.device = dev
        self.is_cuda = dev.type == "cuda"
        if self.is_cuda:
            if fabric is None:
                raise ValueError("CUDA parameters need a Fabric (symmetric heap)")
            world, rank = fabric.world, fabric.rank
        else:
            world = 1 if world is None else world
            rank = 0 if rank is None else rank
        self.world, self.rank = world, rank
        self.clip_norm = clip_norm
        self.overlap = overlap and self.is_cuda
        self.broadcast = broadcast
        self.lazy_dedicated = lazy_dedicated
        # K1 variant policy.  None = auto: buckets pushed while backward is still running use the
        # bulk-copy kernel on `overlap_grid` SMs (the rest keep computing); the last bucket has
        # nothing left to overlap with and takes the whole GPU with the register-path kernel.
        self.push_tma = push_tma
        if "PSB_PUSH_NVLS" in os.environ:
            push_nvls = os.environ["PSB_PUSH_NVLS"] != "0"
        self._want_nvls = ((world >= 4 and push_tma is None) if push_nvls is None
                           else bool(push_nvls))  # an explicit K1 variant request wins over auto
        self.overlap_grid = int(os.environ.get("PSB_OVERLAP_GRID", overlap_grid))
        self._user_grid = push_grid
        optim = OptimConfig(**{**optim.__dict__, "grad_scale": optim.grad_scale / world})
        self.optim = optim

        elem_bytes = 2
        if "PSB_SPLIT_MB" in os.environ:
            split_mb = float(os.environ["PSB_SPLIT_MB"])
        if split_mb is None:
            split_mb = 16.0 if (world >= 4 and self.is_cuda
                                and type(self) is ShardedParameterServer) else 0.0
        self.plan: BucketPlan = plan_buckets(
            [(n, tuple(p.shape)) for n, p in named_params], world,
            bucket_elems=max(1, int(bucket_mb * (1 << 20) / elem_bytes)), dedicated=dedicated,
            split=self._tail_split(named_params, world, split_mb, bucket_mb, set(dedicated)))
        if self.is_cuda:
            from .device_backend import DeviceBackend
            gdt = named_params[0][1].dtype
            if gdt not in (torch.bfloat16, torch.float32):
                raise TypeError(f"unsupported parameter dtype {gdt}")
            if any(p.dtype != torch.bfloat16 for _, p in named_params):
                raise TypeError("the CUDA backend keeps bf16 replicas: cast the model to bfloat16")
            if alt_params is None:
                alt_params = (world > 1 and broadcast and clip_norm is None
                              and type(self) is ShardedParameterServer)
            mk = lambda alt: DeviceBackend(fabric, self.plan, optim, grad_dtype=torch.bfloat16,  # noqa: E731
                                           broadcast=broadcast, clip=clip_norm is not None,
                                           push_grid=push_grid, grad_ring=grad_ring,
                                           push_tma=bool(push_tma), alt_params=alt)
            # a second replica only makes sense where pushes write parameters during backward
            alt_params = (bool(alt_params) and broadcast and clip_norm is None
                          and type(self) is ShardedParameterServer)
            if "PSB_ALT_PARAMS" in os.environ:  # A/B knob
                alt_params = alt_params and os.environ["PSB_ALT_PARAMS"] != "0"
            if alt_params:
                # gradient ring + two replicas must fit in what is left of the symmetric heap
                need = (grad_ring + 2) * (self.plan.total_elems * 2 + 8192)
                if fabric.free_bytes() < need:
                    alt_params = False
            self.backend = mk(alt_params)
            self.comm_stream = torch.cuda.Stream(device=dev, priority=-1)
        else:
            from .host_backend import HostBackend
            pdt = named_params[0][1].dtype
            self.backend = HostBackend(self.plan, optim, rank, world, group, param_dtype=pdt,
                                       grad_dtype=pdt, broadcast=broadcast,
                                       clip=clip_norm is not None,
                                       alt_params=bool(alt_params) and broadcast
                                       and clip_norm is None
                                       and type(self) is ShardedParameterServer)
            self.comm_stream = None
        self.fabric = fabric

        # ---- adopt the model: parameters become views of the replica
        be = self.backend
        self._params: Dict[str, torch.nn.Parameter] = dict(named_params)
        # 4-D weights kept in channels-last order (cuDNN's preferred layout) stay that way:
        # the flat space stores them NHWC and the parameter is a permuted view of it
        self._nhwc = {name for name, p in named_params
                      if p.dim() == 4 and not p.is_contiguous()
                      and p.is_contiguous(memory_format=torch.channels_last)}
        with torch.no_grad():
            for name, p in named_params:
                slot = self.plan.slots[name]
                bk = self.plan.buckets[slot.bucket]
                view = be.param_flat[bk.base + slot.offset: bk.base + slot.offset + slot.numel]
                if name in self._nhwc:
                    n, c, h, w = p.shape
                    view.copy_(p.detach().permute(0, 2, 3, 1).reshape(-1))
                    p.data = view.view(n, h, w, c).permute(0, 3, 1, 2)
                else:
                    view.copy_(p.detach().reshape(-1))
                    p.data = view.view(p.shape)
            for b, bk in enumerate(self.plan.buckets):
                s = bk.slice_elems
                be.shard_slice(be.master, b).copy_(
                    be.param_bucket(b)[rank * s:(rank + 1) * s].to(torch.float32))
        if sync_init and world > 1:
            self._sync_replicas()

        # ---- backward hooks
        self._pending: List[List[Tuple[ParamSlot, torch.Tensor]]] = [[] for _ in self.plan.buckets]
        self._flushed = [False] * len(self.plan.buckets)
        self._events = ([torch.cuda.Event() for _ in self.plan.buckets] if self.is_cuda else None)
        self._hooks = []
        for name, p in named_params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(name)))
        # gradient sinks: layers that know about them (ops/linear.py) write dW / db straight into
        # the bucket and report it the moment the GEMM is queued, before their dX GEMM, so the
        # push of that bucket starts one GEMM earlier and no staging copy is made
        self._alt = getattr(self.backend, "param_flat2", None) is not None
        self._area = [0] * len(self.plan.buckets)   # replica the model reads, per bucket
        self._wrote = [None] * len(self.plan.buckets)  # replica the push of this step writes
        self._sinks: List[GradSink] = []
        if self._sinks_supported():
            for name, p in named_params:
                if name not in self._nhwc:
                    p._psb_sink = GradSink(self, self.plan.slots[name])
                    self._sinks.append(p._psb_sink)
        self._sink_params = [p for _, p in named_params if hasattr(p, "_psb_sink")]
        self.steps_done = 0
        self.metrics = Metrics()

    # ------------------------------------------------------------------ setup helpers
    @staticmethod
    def _tail_split(named_params, world: int, split_mb: float, bucket_mb: float, skip) -> dict:
        """``name -> pieces``.  The 2-D tensors among the last ``bucket_mb`` megabytes that
        backward produces (= the first ones in model order) are cut into blocks of ``split_mb``:
        only the last block's push stays exposed.  Other matrices much larger than a bucket are
        cut into bucket-sized blocks, so that their push starts while their own dW GEMM is still
        running instead of after it (the push stream is the critical path at 8 GPUs)."""
        from .buckets import split_rows
        out = {}
        if split_mb <= 0:
            return out
        seen = 0.0
        for name, p in named_params:
            mb = p.numel() * 2 / (1 << 20)
            tail = seen < bucket_mb
            seen += mb
            target = split_mb if tail else bucket_mb
            if p.dim() != 2 or name in skip or mb < (target * 1.5 if not tail else target + 1e-9):
                continue
            if tail and mb <= split_mb:
                continue
            want = int(-(-mb // target))
            for k in range(max(want, 2), 2 * want + 1):
                if split_rows(tuple(p.shape), k, world) is not None:
                    out[name] = k
                    break
        return out

    def _sync_replicas(self) -> None:
        """Make every replica the concatenation of the owners' slices (owners are
        authoritative), so ranks agree even if their initialisation seeds differed."""
        be = self.backend
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
            self.fabric.barrier(1)
            for b in range(len(self.plan.buckets)):
                be.pull(b, skip_wait=True)
            self.fabric.barrier(1)
            torch.cuda.synchronize(self.device)
        else:
            for b in range(len(self.plan.buckets)):
                be.pull(b)

    def _make_hook(self, name: str):
        slot = self.plan.slots[name]

        def hook(p: torch.nn.Parameter):
            g = p.grad
            if g is None:
                return  # the layer delivered this gradient through the sink (autograd still
                        # runs the hook when a backward function returns None)
            p.grad = None
            if slot.pieces:  # delivered whole by autograd: hand each row block to its bucket
                for piece in slot.pieces:
                    self._arrive(piece, g[piece.row_lo:piece.row_hi])
            else:
                self._arrive(slot, g)

        return hook

    def _arrive(self, slot: ParamSlot, g: Optional[torch.Tensor]) -> None:
        """One parameter's gradient is ready (``g is None``: already written in the bucket)."""
        self._pending[slot.bucket].append((slot, g))
        if len(self._pending[slot.bucket]) == len(self.plan.buckets[slot.bucket].slots):
            self._flush(slot.bucket)

    def _alt_ok(self, b: int) -> bool:
        """Is bucket ``b`` double-buffered?  Broadcast buckets only; the bucket of the fused
        first layer stays single (its GEMM reads the peers' first replicas directly)."""
        return (self._alt and self._bcast(b)
                and not (self.plan.buckets[b].dedicated and self._fused_first_layer))

    def _early_ok(self, b: int) -> bool:
        return self.clip_norm is not None or self._alt_ok(b)

    def _param_view(self, flat: torch.Tensor, name: str, shape) -> torch.Tensor:
        slot = self.plan.slots[name]
        bk = self.plan.buckets[slot.bucket]
        view = flat[bk.base + slot.offset: bk.base + slot.offset + slot.numel]
        if name in self._nhwc:
            n, c, h, w = shape
            return view.view(n, h, w, c).permute(0, 3, 1, 2)
        return view.view(shape)

    def _swap_replicas(self) -> None:
        """After the pushes of a step: the model reads, from now on, the replica they wrote."""
        be = self.backend
        changed = set()
        for b, area in enumerate(self._wrote):
            if area is not None and area != self._area[b]:
                self._area[b] = area
                changed.add(b)
            self._wrote[b] = None
        if not changed:
            return
        for name, p in self.named_params:
            b = self.plan.slots[name].bucket
            if b in changed:
                flat = be.param_flat2 if self._area[b] == 1 else be.param_flat
                p.data = self._param_view(flat, name, p.shape)

    def _sinks_supported(self) -> bool:
        """In-place gradient buckets need the bucket to be free for the whole backward pass:
        true for the synchronous engine (``step`` joins the push stream before the next forward)."""
        return type(self) is ShardedParameterServer

    # ------------------------------------------------------------------ per-bucket push
    def _stage(self, b: int) -> None:
        """Copy the bucket's gradients into the symmetric space (zeros where none arrived)."""
        be = self.backend
        pend = self._pending[b]
        bucket = be.grad_bucket(b)
        slots = self.plan.buckets[b].slots
        if len(pend) < len(slots):  # parameters that took no part in this backward pass
            have = {s.name for s, _ in pend}
            for s in slots:
                if s.name not in have:
                    bucket[s.offset:s.offset + s.numel].zero_()
        pend = [(s, g) for s, g in pend if g is not None]  # None: written in place by a sink
        if self.is_cuda:
            if pend:
                native.gather_to_bucket([self._flat_grad(s, g) for s, g in pend], bucket,
                                        [s.offset for s, _ in pend])
        else:
            for s, g in pend:
                bucket[s.offset:s.offset + s.numel].copy_(self._flat_grad(s, g).reshape(-1))
        self._pending[b] = []

    def _flat_grad(self, slot: ParamSlot, g: torch.Tensor) -> torch.Tensor:
        """Gradient in the element order of the flat space (NHWC for channels-last weights)."""
        if slot.name in self._nhwc:
            return g.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)
        return g.contiguous()

    def _bcast(self, b: int) -> bool:
        """Does the owner of bucket ``b`` write its new slice into every replica?"""
        if self.lazy_dedicated and self.plan.buckets[b].dedicated and self._fused_first_layer:
            return False
        return self.broadcast

    def _variant(self, b: int) -> dict:
        """Kernel variant / grid of K1 for bucket ``b`` (see ``push_tma`` above)."""
        if not self.is_cuda:
            return {}
        last = b == len(self.plan.buckets) - 1
        if (self._want_nvls and self.world > 1 and self.clip_norm is None
                and getattr(self.backend, "nvls_available", False)
                and type(self) is ShardedParameterServer):
            # one load + one store per 16 bytes: few CTAs fill the link even next to a GEMM
            return {"nvls": True, "grid": self._user_grid or (0 if last else self.overlap_grid)}
        if self.push_tma is not None:
            return {"tma": self.push_tma, "grid": self._user_grid}
        # the last bucket has nothing left to overlap with: a big one takes the whole GPU with
        # the register-path kernel; a small one (a row block of a split tensor) is latency-bound
        # and the bulk-copy kernel is the faster of the two there (bw_sweep: 16 MiB 73 vs 82 us)
        big_last = last and self.plan.buckets[b].numel * 2 > (32 << 20)
        if self.world > 1 and self.overlap and not big_last and self.clip_norm is None:
            return {"tma": True, "grid": self._user_grid or self.overlap_grid}
        # register-path kernel, 512 threads per CTA.  (Measured on B200, 1 GPU, MLP step: 256-thread
        # CTAs that share the SMs with the tcgen05 GEMM cost 0.08-0.15 ms per step more than
        # 512-thread CTAs that take turns with it — profiles/overlap_matrix_1gpu.txt.)
        block = int(os.environ.get("PSB_PUSH_BLOCK", "0"))  # experiment knob
        return {"tma": False, "grid": self._user_grid, "block": block}

    def _flush(self, b: int) -> None:
        self._stage(b)
        self._flushed[b] = True
        clip = self.clip_norm is not None
        kw = dict(reduce_only=clip, broadcast=self._bcast(b), **self._variant(b))
        if self._alt_ok(b):
            self._wrote[b] = 1 - self._area[b]
            kw["alt"] = self._wrote[b] == 1
        bk = self.plan.buckets[b]
        self.metrics.inc("pushes")
        # bytes entering this GPU over NVLink for the bucket: peers' gradient slices of my shard
        # plus (broadcast) the other owners' refreshed slices of my replica
        link = (self.world - 1) * bk.slice_elems * 2
        self.metrics.inc("nvlink_in_bytes", link * (2 if self._bcast(b) else 1))
        if not self.is_cuda:
            self.backend.push(b, **kw)
            return
        if self.overlap:
            ev = self._events[b]
            ev.record(torch.cuda.current_stream(self.device))
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream), nvtx_range(f"psb.push[{b}]"):
                self.backend.push(b, **kw)
        else:
            with nvtx_range(f"psb.push[{b}]"):
                self.backend.push(b, **kw)

    # ------------------------------------------------------------------ end of iteration
    def step(self) -> None:
        """Finish the iteration: after this returns (stream-ordered on CUDA) the model's
        parameters are the updated ones on every rank."""
        be = self.backend
        nb = len(self.plan.buckets)
        for b in range(nb):
            if not self._flushed[b]:
                self._flush(b)
        if self.is_cuda and self.overlap:
            ctx = torch.cuda.stream(self.comm_stream)
        else:
            import contextlib
            ctx = contextlib.nullcontext()
        with ctx:
            if self.clip_norm is not None:
                be.norm_allreduce(self.clip_norm)
                for b in range(nb):
                    be.push(b, from_reduced=True, clip=True, broadcast=self._bcast(b))
            be.advance_step()
            for b in range(nb):
                fused = self.plan.buckets[b].dedicated and self._fused_first_layer
                if not self._bcast(b) and not fused:
                    be.pull(b)  # pull-mode bucket without a fused consumer: K2
        if self.is_cuda:
            cur = torch.cuda.current_stream(self.device)
            if self.overlap:
                cur.wait_stream(self.comm_stream)
            # broadcast buckets: wait until every owner's slice has landed in the replica
            lo = None
            for b in range(nb + 1):
                bc = b < nb and self._bcast(b)
                if bc and lo is None:
                    lo = b
                if not bc and lo is not None:
                    be.wait_versions(lo, b)
                    lo = None
        if self._alt:
            self._swap_replicas()
        for s in self._sinks:
            s.uses = 0
        self._flushed = [False] * nb
        self.steps_done += 1
        self.metrics.inc("steps")

    _fused_first_layer = False

    # ------------------------------------------------------------------ shard state <-> names
    def gather_state(self, which: str = "master", group=None) -> Dict[str, torch.Tensor]:
        """Reassemble a sharded fp32 state (``master`` / ``m`` / ``v``) into named tensors on
        every rank.  Control-plane path (checkpoints, tests).  On CUDA the shards are read from
        the peers through the symmetric heap; the host backend uses its gloo group."""
        be = self.backend
        shard = getattr(be, which)
        if shard is None:
            return {}
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
        if self.world > 1 and self.is_cuda:
            parts = self.fabric.all_gather(shard)
        elif self.world > 1:
            import torch.distributed as dist
            parts = [torch.empty_like(shard) for _ in range(self.world)]
            dist.all_gather(parts, shard.contiguous(), group=group)
        else:
            parts = [shard]
        out = {}
        for name, slot in self.plan.slots.items():
            flat = torch.empty(slot.numel, dtype=torch.float32, device=shard.device)
            for r, a, b, lo in self.plan.shard_ranges(slot):
                flat[a:b].copy_(parts[r][lo:lo + (b - a)])
            if name in self._nhwc:
                n, c, h, w = slot.shape
                out[name] = flat.view(n, h, w, c).permute(0, 3, 1, 2)
            else:
                out[name] = flat.view(slot.shape)
        return out

    def scatter_state(self, tensors: Dict[str, torch.Tensor], which: str = "master") -> None:
        """Inverse of :meth:`gather_state` for this rank's shard; ``master`` also refreshes the
        local replica (every rank must call it with the same tensors)."""
        be = self.backend
        shard = getattr(be, which)
        if shard is None:
            return
        with torch.no_grad():
            for name, t in tensors.items():
                slot = self.plan.slots.get(name)
                if slot is None:
                    continue
                bk = self.plan.buckets[slot.bucket]
                t = t.detach()
                if name in self._nhwc:
                    t = t.permute(0, 2, 3, 1)
                flat = t.reshape(-1).to(device=shard.device, dtype=torch.float32)
                for r, a, b, lo in self.plan.shard_ranges(slot):
                    if r == self.rank:
                        shard[lo:lo + (b - a)].copy_(flat[a:b])
                if which == "master":
                    for pf in (be.param_flat, getattr(be, "param_flat2", None)):
                        if pf is not None:
                            view = pf[bk.base + slot.offset: bk.base + slot.offset + slot.numel]
                            view.copy_(flat.to(view.dtype))
        if self.is_cuda:
            torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ introspection
    def grad_norm(self) -> torch.Tensor:
        return self.backend.grad_norm

    def set_lr(self, lr: float) -> None:
        self.optim = OptimConfig(**{**self.optim.__dict__, "lr": lr})
        self.backend.set_optim(self.optim)

    def close(self) -> None:
        """Detach from the model and give the symmetric buffers back to the heap: the
        parameters become ordinary tensors holding their current values.  Collective (every
        rank closes its engines in the same order)."""
        for h in self._hooks:
            h.remove()
        self._hooks = []
        if self.is_cuda and getattr(self.backend, "P", None) is not None:
            torch.cuda.synchronize(self.device)
            with torch.no_grad():
                for _, p in self.named_params:
                    p.data = p.data.clone()
            self.backend.close()
        for p in self._sink_params:
            if hasattr(p, "_psb_sink"):
                del p._psb_sink
        self._sink_params = []
