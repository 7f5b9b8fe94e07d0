"""Public training API: ``init()`` + ``Trainer``.

``Trainer.step(*host_batch)`` is the call a user makes per iteration.  It is the whole
reference iteration (``Worker::run_iteration``, src/worker.cpp:331-406: pull -> compute
gradients -> push -> wait for sync) for one worker/shard pair: H2D copy of the batch from
pinned host memory, forward, backward with bucket-wise fused pushes, the version wait that
makes the updated parameters visible, and a D2H read of the loss.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Callable, Optional

import torch
import torch.distributed as dist

from .ops.native import OptimConfig
from .parallel.engine import ShardedParameterServer
from .parallel.fabric import Fabric


@dataclass
class Context:
    """One rank of a job: identity, device, symmetric heap, and the host-side control plane
    (``coord``: the C++ coordinator; else a ``torch.distributed`` group; else single rank)."""
    rank: int
    world: int
    local_rank: int
    device: torch.device
    fabric: Optional[Fabric]
    coord: Optional[object] = None   # control.client.CoordinatorClient
    abort: Optional[object] = None   # utils.elastic.AbortWord
    numa: Optional[dict] = None      # utils.numa.bind_to_gpu_node: where this process was bound

    # ---- host-side collectives of the control plane (small, off the training path)
    def barrier(self, name: str = "host") -> None:
        if self.world == 1:
            return
        if self.coord is not None:
            self.coord.barrier(name)
        elif dist.is_initialized():
            dist.barrier()

    def all_gather_object(self, obj, key: str = "obj") -> list:
        if self.world == 1:
            return [obj]
        if self.coord is not None:
            return self.coord.allgather_object(key, obj)
        out = [None] * self.world
        dist.all_gather_object(out, obj)
        return out

    def reduce_max(self, x: float, key: str = "max") -> float:
        return max(self.all_gather_object(float(x), key))

    @property
    def peer_lost(self) -> bool:
        """A peer was evicted by the coordinator (missed heartbeats) since ``init``."""
        return self.coord is not None and self.coord.peer_lost

    def close(self) -> None:
        if self.coord is not None:
            self.coord.watch(0)
        if self.fabric is not None:
            self.fabric.close()
            self.fabric = None
        if self.abort is not None:
            self.abort.close()
            self.abort = None
        if self.coord is not None:
            self.coord.close()
            self.coord = None
        if dist.is_initialized():
            dist.destroy_process_group()


def init(heap_bytes: int = 4 << 30, backend: Optional[str] = None,
         spin_timeout_s: Optional[float] = None, control: Optional[str] = None) -> Context:
    """Join the job described by RANK / LOCAL_RANK / WORLD_SIZE (torchrun, or
    scripts/deploy_box.sh) and build the symmetric heap.

    Control plane (``control`` / ``PSB_CONTROL``):

    * ``"coordinator"`` — default for CUDA jobs with more than one rank: every rank registers
      with the C++ coordinator (``PSB_COORDINATOR=host:port``; under plain torchrun rank 0
      launches ``build/bin/coordinator`` itself), heartbeats from a native thread, registers its
      PS shard, exchanges the heap handles and synchronises on the host through it — the
      bootstrap of the reference's workers (/root/reference/src/worker.cpp:108-122, 141-186,
      216-238).  No ``torch.distributed`` process group exists in this mode.
    * ``"dist"`` — ``torch.distributed`` (gloo on CPU: the host backend moves its data through
      it; nccl on CUDA) as control plane.

    ``spin_timeout_s`` (or ``PSB_SPIN_TIMEOUT_S``; kernel default 300 s): how long a kernel may
    wait on a flag of a peer before it traps — a dead peer then surfaces as a CUDA error on this
    rank instead of a GPU that spins for ever.  With the coordinator the heartbeat thread
    normally gets there first: an evicted peer raises the abort word and waiting kernels
    return (``Context.peer_lost``)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    cuda = torch.cuda.is_available()
    control = control or os.environ.get("PSB_CONTROL")
    if control is None:
        from .control import client as _client
        want = cuda and (world > 1 or "PSB_COORDINATOR" in os.environ)
        control = "coordinator" if want and _client.available() else "dist"
    if control not in ("coordinator", "dist"):
        raise ValueError(f"unknown control plane {control!r}")
    numa = None
    if cuda:
        torch.cuda.set_device(local_rank)
        device = torch.device("cuda", local_rank)
        from .utils.numa import bind_to_gpu_node
        numa = bind_to_gpu_node(local_rank)  # before the first pinned allocation (PSB_NUMA_BIND=0: off)
        if spin_timeout_s is None and "PSB_SPIN_TIMEOUT_S" in os.environ:
            spin_timeout_s = float(os.environ["PSB_SPIN_TIMEOUT_S"])
        if spin_timeout_s is not None:
            from . import _lib
            _lib.check(_lib.load().psb_set_spin_timeout_ms(max(1, int(spin_timeout_s * 1000))),
                       "psb_set_spin_timeout_ms")
    else:
        device = torch.device("cpu")
    coord = abort = None
    if control == "coordinator":
        from .control.client import ensure_coordinator
        coord = ensure_coordinator(rank, world)
        if cuda:
            coord.register_shard(rank, f"hbm://{os.uname().nodename}/gpu{local_rank}", local_rank)
            fabric = Fabric.from_coordinator(coord, device, heap_bytes)
            from .utils.elastic import AbortWord
            abort = AbortWord()           # host-mapped word polled by every peer spin-wait
            coord.barrier("init/registered")
            coord.watch(world, abort.host_ptr)
        else:
            fabric = None
        return Context(rank, world, local_rank, device, fabric, coord, abort, numa)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if cuda else "gloo"
        kw = {"device_id": device} if (cuda and backend == "nccl") else {}
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    fabric = Fabric.from_process_group(device, heap_bytes) if cuda else None
    return Context(rank, world, local_rank, device, fabric, numa=numa)


class LossHandle:
    """Loss of one step, copied to pinned host memory asynchronously."""

    def __init__(self, host: torch.Tensor, event):
        self._host, self._event = host, event

    def item(self) -> float:
        if self._event is not None:
            self._event.synchronize()
        return float(self._host.item())


class Trainer:
    def __init__(self, model: torch.nn.Module, loss_fn: Callable, ctx: Context,
                 optim: OptimConfig, *, mode: str = "sync", bucket_mb: float = 64.0,
                 pull: str = "broadcast", clip_norm: Optional[float] = None,
                 fused_first_layer: Optional[bool] = None, overlap: bool = True,
                 staleness: int = 0, push_grid: int = 0, push_tma: Optional[bool] = None,
                 checkpoint_every: int = 0, checkpoint_path: Optional[str] = None,
                 split_mb: Optional[float] = None, push_nvls: Optional[bool] = None,
                 tables=()):
        """``tables``: sparse embedding tables the model looks rows up in
        (``parallel.sparse.ShardedEmbedding``); after every backward pass their collected row
        gradients are pushed (``apply_gradients``) next to the dense push.

        ``pull``: ``"broadcast"`` — owners write new slices into every replica inside K1;
        ``"gather"`` — owners publish locally, workers pull (K2, and K2f for the first layer);
        ``"hybrid"`` — broadcast for every bucket except the first layer's, whose all-gather is
        fused into the first GEMM of the next forward (K2f).

        ``checkpoint_every`` / ``checkpoint_path``: write a checkpoint (reference format +
        sidecar, atomically) every so many steps; it is what the surviving ranks restart from
        after a :class:`~parameter_server_distributed_b200.utils.elastic.PeerLost`."""
        if mode not in ("sync", "async", "ssp"):
            raise ValueError(f"unknown mode {mode!r}")
        self.model, self.loss_fn, self.ctx = model, loss_fn, ctx
        self.device = ctx.device
        self.tables = list(tables)
        dedicated = ()
        first = getattr(model, "first_weight_name", None)
        if pull not in ("broadcast", "gather", "hybrid"):
            raise ValueError(f"unknown pull mode {pull!r}")
        want_fused = (pull in ("gather", "hybrid")) if fused_first_layer is None else fused_first_layer
        if want_fused and first is not None and ctx.device.type == "cuda":
            from .parallel.fused_linear import FusedFirstLinear
            shape = tuple(dict(model.named_parameters())[first()].shape)
            if FusedFirstLinear.supported(shape, ctx.world):
                dedicated = (first(),)
        if mode == "sync":
            self.ps = ShardedParameterServer(
                list(model.named_parameters()), optim, ctx.fabric, world=ctx.world, rank=ctx.rank,
                bucket_mb=bucket_mb, broadcast=(pull != "gather"), clip_norm=clip_norm,
                dedicated=dedicated, overlap=overlap, push_grid=push_grid,
                push_tma=push_tma, lazy_dedicated=(pull == "hybrid"), split_mb=split_mb,
                push_nvls=push_nvls,
                # two parameter replicas pay off only for layers that commit their gradients
                # before their dX GEMM (ops/linear.py); other models keep one replica
                alt_params=None if getattr(model, "grad_sink_layers", False) else False)
        else:
            from .parallel.async_engine import AsyncParameterServer
            self.ps = AsyncParameterServer(
                list(model.named_parameters()), optim, ctx.fabric, world=ctx.world, rank=ctx.rank,
                bucket_mb=bucket_mb, staleness=(staleness if mode == "ssp" else None))
        if dedicated and pull in ("gather", "hybrid"):
            from .parallel.fused_linear import FusedFirstLinear
            model.first_linear_fn = FusedFirstLinear(self.ps, dedicated[0])
        self.checkpoint_every = int(checkpoint_every)
        self.checkpoint_path = checkpoint_path
        if self.checkpoint_every and not checkpoint_path:
            raise ValueError("checkpoint_every needs a checkpoint_path")
        self.steps = 0
        self._prefetched = None
        self._dev_inputs = None
        self._loss_ring = None
        self._ring_i = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 4

    # ------------------------------------------------------------------ checkpoints
    def save_checkpoint(self, path: str, epoch: int = 0) -> None:
        """Collective.  Writes the fp32 master weights in the reference's checkpoint format
        (loadable by the reference's ``LoadCheckpoint``, /root/reference/src/parameter_server.cpp:
        146-188) plus a ``.psb`` sidecar with the optimizer state.  In async / ssp mode the
        gradients still in flight are applied first."""
        from .utils.checkpoint import save_checkpoint
        if hasattr(self.ps, "drain"):
            self.ps.drain()
        # kernels give up their waits only once the abort word is set, so state gathered while
        # it is still clear is sound; anything else must not replace the last good file
        save_checkpoint(self.ps, path, epoch=epoch, validate=self._guard)
        for i, t in enumerate(self.tables):  # each table as one tensor next to the model's file
            t.save(f"{path}.table{i}", name=f"table{i}", epoch=epoch)

    def load_checkpoint(self, path: str):
        """Collective.  Restores master weights (and optimizer state when the sidecar exists);
        returns ``(epoch, iteration)``."""
        from .utils.checkpoint import load_checkpoint
        out = load_checkpoint(self.ps, path)
        for i, t in enumerate(self.tables):
            if os.path.exists(f"{path}.table{i}"):
                t.load(f"{path}.table{i}")
        return out

    # ------------------------------------------------------------------ failure handling
    def _guard(self) -> None:
        """Before any launch of a step: has the coordinator evicted a peer?  Its heartbeat
        thread has then already raised the abort word, so kernels that were waiting for the dead
        rank have returned; nothing computed since is trustworthy."""
        if self.ctx.peer_lost or (self.ctx.abort is not None and self.ctx.abort.is_set):
            from .utils.elastic import PeerLost
            live = self.ctx.coord.live_workers() if self.ctx.coord is not None else "?"
            raise PeerLost(f"rank {self.ctx.rank}: a peer left the job (coordinator reports "
                           f"{live} of {self.ctx.world} workers alive); "
                           f"restart from the last checkpoint"
                           + (f" ({self.checkpoint_path})" if self.checkpoint_path else ""))

    def _after_step(self) -> None:
        self.steps += 1
        if self.checkpoint_every and self.steps % self.checkpoint_every == 0:
            self._guard()  # never publish a checkpoint of a step that a lost peer left half done
            self.save_checkpoint(self.checkpoint_path, epoch=self.steps // self.checkpoint_every)

    # ------------------------------------------------------------------ device-resident step
    def step_device(self, *inputs: torch.Tensor) -> torch.Tensor:
        """Forward + backward + push/pull on tensors already on the device; returns the loss
        tensor (on device)."""
        self._guard()
        *xs, target = inputs
        out = self.model(*xs)
        loss = self.loss_fn(out, target)
        loss.backward()
        self.ps.step()
        for t in self.tables:  # sparse rows: collective, once per step on every rank
            t.apply_gradients()
        self._after_step()
        return loss.detach()

    # ------------------------------------------------------------------ end-to-end step
    def _stage(self, host_inputs) -> tuple:
        """Queue the H2D copy of one batch into the next device buffer set (copy stream);
        returns ``(slot, copied_event)``."""
        if self._dev_inputs is None:
            self._dev_inputs = [[torch.empty(t.shape, dtype=t.dtype, device=self.device)
                                 for t in host_inputs] for _ in range(2)]
            self._consumed = [None, None]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._loss_stream = torch.cuda.Stream(device=self.device)
            self._loss_ring = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(4)]
            self.h2d_bytes = sum(t.numel() * t.element_size() for t in host_inputs)
        slot = self._ring_i & 1
        with torch.cuda.stream(self._copy_stream):
            if self._consumed[slot] is not None:
                self._copy_stream.wait_event(self._consumed[slot])  # step i-2 is done with it
            for d, h in zip(self._dev_inputs[slot], host_inputs):
                d.copy_(h, non_blocking=True)
            copied = torch.cuda.Event()
            copied.record(self._copy_stream)
        return slot, copied

    def prefetch(self, *host_inputs: torch.Tensor) -> None:
        """Start the H2D copy of the NEXT step's batch now; the following ``step()`` (called
        without arguments, or with these same tensors) uses it.  Call it right after ``step``
        of the current batch: the copy then has the whole current step to finish in, instead of
        the part of it that is left once the caller has read the previous loss — with eight
        ranks sharing the host's memory system a 40 MB batch takes longer than that part."""
        if self.device.type != "cuda":
            self._prefetched = (None, None, host_inputs)
            return
        if self._prefetched is not None:
            raise RuntimeError("prefetch() called twice without a step() in between")
        # the buffer set the next step will use is the one after the step that is in flight
        slot, copied = self._stage(host_inputs)
        self._prefetched = (slot, copied, host_inputs)

    def step(self, *host_inputs: torch.Tensor) -> LossHandle:
        """End-to-end iteration from (pinned) host tensors; returns a handle on the loss.

        The H2D copy of step i runs on a copy stream into one of two device buffer sets, so it
        overlaps the tail of step i-1 (all of it after :meth:`prefetch`); the loss goes back
        through a pinned ring."""
        pre, self._prefetched = self._prefetched, None
        if self.device.type != "cuda":
            if pre is not None and not host_inputs:
                host_inputs = pre[2]
            loss = self.step_device(*host_inputs)
            return LossHandle(loss.reshape(()).float(), None)
        self._guard()
        if pre is not None:
            if host_inputs and any(a is not b for a, b in zip(host_inputs, pre[2])):
                raise ValueError("step() got a different batch than the one that was prefetched")
            slot, copied = pre[0], pre[1]
        else:
            if not host_inputs:
                raise ValueError("step() without arguments needs a prefetch() first")
            slot, copied = self._stage(host_inputs)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(copied)
        *xs, target = self._dev_inputs[slot]
        loss = self.loss_fn(self.model(*xs), target)
        # the loss goes home as soon as the forward pass has produced it (its own stream), not
        # after backward and the pushes: a caller that reads it every step then waits for a
        # third of a step instead of a whole one and keeps the launch queue fuller
        host = self._loss_ring[self._ring_i % len(self._loss_ring)]
        self._ring_i += 1
        loss_val = loss.detach().float()
        ready = torch.cuda.Event()
        ready.record(cur)
        with torch.cuda.stream(self._loss_stream):
            self._loss_stream.wait_event(ready)
            host.copy_(loss_val, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._loss_stream)
        loss_val.record_stream(self._loss_stream)
        loss.backward()
        self.ps.step()
        for t in self.tables:
            t.apply_gradients()
        done = torch.cuda.Event()
        done.record(cur)
        self._consumed[slot] = done
        self._after_step()
        return LossHandle(host, ev)
