"""K5 — sparse embedding push / pull: a row-sharded embedding table on the parameter server.

The reference moves the whole model on every push and pull (``Worker::push_gradients`` /
``pull_parameters``, /root/reference/src/worker.cpp:240-312) and its ``Tensor`` is dense
(/root/reference/proto/parameter_server.proto:19-24), so an embedding table — where a batch
touches a few thousand of millions of rows — has no workable path there.  Here:

* rank ``r`` owns rows ``[r * rows_per_rank, (r + 1) * rows_per_rank)``: fp32 master rows and
  optimizer state stay with the owner, a bf16 copy of the shard lives in the symmetric heap;
* ``pull(indices)`` reads exactly the requested rows from their owners' HBM over NVLink peer
  loads (``csrc/kernels/sparse.cu``: ``sparse_pull_kernel``);
* ``push(indices, grads)`` (collective, sync mode) stages the worker's rows in its own heap
  range; after a device barrier every owner marks, accumulates (fp32) and applies the rows it
  owns — mean over the workers, *lazy* SGD / momentum / Adam(W): rows without a gradient keep
  their state — and a second barrier makes the new rows visible to every puller.

CPU tensors take the host path (plain PyTorch, ``gloo`` for world > 1) with the same layout and
arithmetic; it is the oracle of the kernels and keeps the logic testable without a GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from .. import _lib
from .._lib import PsbFabric, PsbSparse, check
from ..ops import reference as ref
from ..ops.native import OptimConfig
from ..utils.tracing import Metrics, nvtx_range

_DEVICE_PATH_ON_HOST = False
_BARRIER_SLOT = 8      # device-barrier slot of the sparse path (0-4 belong to the engines)
_INIT_BLOCK = 4096     # rows per seeded block of the initialiser (layout independent of world)


def _init_rows(lo: int, hi: int, dim: int, std: float, seed: int) -> torch.Tensor:
    """Rows [lo, hi) of the N(0, std^2) initial table; the same values at every world size."""
    out = torch.empty(hi - lo, dim, dtype=torch.float32)
    b = lo // _INIT_BLOCK
    while b * _INIT_BLOCK < hi:
        g = torch.Generator().manual_seed(seed * 1000003 + b)
        blk = torch.randn(_INIT_BLOCK, dim, generator=g) * std
        s, e = max(lo, b * _INIT_BLOCK), min(hi, (b + 1) * _INIT_BLOCK)
        out[s - lo:e - lo] = blk[s - b * _INIT_BLOCK:e - b * _INIT_BLOCK]
        b += 1
    return out


class ShardedEmbedding:
    """One rank's view of a ``vocab x dim`` embedding table sharded over ``world`` ranks.

    ``capacity``: the most rows one worker may push per step (staging size).  ``fabric``: the
    symmetric heap of a multi-GPU job (``Context.fabric``); without it a CUDA table is
    single-rank.  ``group``: the ``torch.distributed`` group of the host path."""

    def __init__(self, vocab: int, dim: int, optim: OptimConfig, capacity: int, *,
                 fabric=None, rank: int = 0, world: int = 1, group=None,
                 device: Optional[torch.device] = None, grad_dtype: torch.dtype = torch.bfloat16,
                 init_std: float = 0.02, seed: int = 0):
        if fabric is not None:
            rank, world, device = fabric.rank, fabric.world, fabric.device
        self.device = torch.device(device if device is not None else "cpu")
        if dim % 8:
            raise ValueError("embedding width must be a multiple of 8 (16-byte bf16 vectors)")
        if grad_dtype not in (torch.bfloat16, torch.float32):
            raise TypeError("grad_dtype: bfloat16 or float32")
        self.vocab, self.dim, self.capacity = int(vocab), int(dim), int(capacity)
        self.rank, self.world, self.group, self.fabric = rank, world, group, fabric
        self.optim, self.grad_dtype = optim, grad_dtype
        self.rows_per_rank = (self.vocab + world - 1) // world
        self.lo = min(self.vocab, rank * self.rows_per_rank)
        self.hi = min(self.vocab, self.lo + self.rows_per_rank)
        self.steps = 0
        self.metrics = Metrics("psb_sparse")  # rows / bytes by construction, Prometheus text
        self._pending: List[Tuple[torch.Tensor, torch.Tensor]] = []
        R, D, dev = self.rows_per_rank, self.dim, self.device
        self.master = torch.zeros(R, D, dtype=torch.float32, device=dev)
        self.master[:self.hi - self.lo] = _init_rows(self.lo, self.hi, D, init_std, seed).to(dev)
        self.m = torch.zeros(R, D, dtype=torch.float32, device=dev) if optim.needs_m() else None
        self.v = torch.zeros(R, D, dtype=torch.float32, device=dev) if optim.needs_v() else None
        # _DEVICE_PATH_ON_HOST: test hook (tests/test_sparse_device_path_emulated.py) — the device
        # code path below on host tensors, against the kernels compiled for the host
        self.cuda = self.device.type == "cuda" or _DEVICE_PATH_ON_HOST
        if self.cuda:
            self._build_device()
        else:
            self.table = self.master.to(torch.bfloat16)
        self.sync_table()

    @classmethod
    def from_context(cls, ctx, vocab: int, dim: int, optim: OptimConfig, capacity: int,
                     **kw) -> "ShardedEmbedding":
        """The table of a job started with :func:`parameter_server_distributed_b200.init`:
        on GPUs in the job's symmetric heap, without one on the host path (``gloo``)."""
        if ctx.fabric is not None:
            return cls(vocab, dim, optim, capacity, fabric=ctx.fabric, **kw)
        return cls(vocab, dim, optim, capacity, rank=ctx.rank, world=ctx.world,
                   device=ctx.device, **kw)

    # ------------------------------------------------------------------ device set-up
    def _build_device(self) -> None:
        R, D, cap, W = self.rows_per_rank, self.dim, self.capacity, self.world
        gsz = 2 if self.grad_dtype == torch.bfloat16 else 4
        self.lib = _lib.load()
        if self.fabric is not None:
            self._bufs = [self.fabric.alloc(R * D * 2), self.fabric.alloc(cap * 8),
                          self.fabric.alloc(cap * D * gsz), self.fabric.alloc(256)]
            tb, ib, gb, cb = self._bufs
            self.table = tb.view(torch.bfloat16).view(R, D)
            self.stage_idx = ib.view(torch.int64)
            self.stage_grad = gb.view(self.grad_dtype).view(cap, D)
            self.stage_cnt = cb.view(torch.int32)
            ptrs = [b.peer_ptrs() for b in self._bufs]
            cfab = self.fabric.c_fabric()
        else:
            if W != 1:
                raise ValueError("a CUDA table over several ranks needs the symmetric heap (fabric=)")
            dev = self.device
            self._bufs = []
            self.table = torch.zeros(R, D, dtype=torch.bfloat16, device=dev)
            self.stage_idx = torch.zeros(cap, dtype=torch.int64, device=dev)
            self.stage_grad = torch.zeros(cap, D, dtype=self.grad_dtype, device=dev)
            self.stage_cnt = torch.zeros(64, dtype=torch.int32, device=dev)
            ptrs = [[t.data_ptr()] for t in (self.table, self.stage_idx, self.stage_grad,
                                             self.stage_cnt)]
            cfab = PsbFabric(0, 1)
        dev = self.device
        self.slot = torch.empty(R, dtype=torch.int32, device=dev)
        self.touched = torch.zeros(W * cap, dtype=torch.int32, device=dev)
        self.n_touched = torch.zeros(1, dtype=torch.int32, device=dev)
        self.acc = torch.zeros(W * cap, D, dtype=torch.float32, device=dev)
        self.step_word = torch.zeros(1, dtype=torch.int32, device=dev)
        self.stage_cnt.zero_()
        with torch.cuda.device(dev):
            check(self.lib.psb_sparse_init(self.slot.data_ptr(), R, self._stream()),
                  "psb_sparse_init")
        s = PsbSparse()
        s.fabric = cfab
        for w in range(W):
            s.table[w], s.stage_idx[w], s.stage_grad[w], s.stage_cnt[w] = (
                ptrs[0][w], ptrs[1][w], ptrs[2][w], ptrs[3][w])
        s.vocab, s.rows_per_rank, s.dim, s.cap = self.vocab, R, D, cap
        s.grad_dtype = _lib.BF16 if self.grad_dtype == torch.bfloat16 else _lib.F32
        s.master = self.master.data_ptr()
        s.m = self.m.data_ptr() if self.m is not None else None
        s.v = self.v.data_ptr() if self.v is not None else None
        s.slot, s.touched = self.slot.data_ptr(), self.touched.data_ptr()
        s.n_touched, s.acc = self.n_touched.data_ptr(), self.acc.data_ptr()
        s.step = self.step_word.data_ptr()
        self._c = s
        self.set_optim(self.optim)

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    @property
    def _emulated(self) -> bool:
        """Several ranks of a ``Fabric.local_group`` on one GPU (kernel tests): no spinning
        barriers — the caller runs every rank's stage, then every rank's apply."""
        return self.fabric is not None and not self.fabric._owns

    def _barrier(self) -> None:
        if self.cuda:
            if self.world > 1 and not self._emulated:
                self.fabric.barrier(_BARRIER_SLOT)
        elif self.world > 1:
            dist.barrier(group=self.group)

    # ------------------------------------------------------------------ configuration
    def set_optim(self, optim: OptimConfig) -> None:
        """New hyper-parameters (learning-rate schedules); the kind cannot change."""
        if optim.kind != self.optim.kind:
            raise ValueError("the optimizer kind is fixed at construction (state layout)")
        self.optim = optim
        if self.cuda:
            self._c.opt = optim.to_c(optim.grad_scale / self.world)

    def sync_table(self) -> None:
        """bf16 shard := master (after construction or ``load_state_dict``)."""
        self.table.copy_(self.master.to(torch.bfloat16))
        if self.cuda:
            torch.cuda.current_stream(self.device).synchronize()
        self._barrier()

    # ------------------------------------------------------------------ pull
    def pull(self, indices: torch.Tensor) -> torch.Tensor:
        """Rows ``indices`` (int64, any shape) -> bf16 ``[*indices.shape, dim]``; an index
        outside ``[0, vocab)`` gives a zero row (padding)."""
        idx = indices.reshape(-1).to(device=self.device, dtype=torch.int64).contiguous()
        n = idx.numel()
        self.metrics.inc("pulls")
        self.metrics.inc("rows_pulled", n)
        self.metrics.inc("pull_bytes", n * self.dim * 2)
        if self.cuda:
            out = torch.empty(n, self.dim, dtype=torch.bfloat16, device=self.device)
            if n:
                with torch.cuda.device(self.device), nvtx_range("psb.sparse_pull"):
                    check(self.lib.psb_sparse_pull(C.byref(self._c), idx.data_ptr(), n,
                                                   out.data_ptr(), self._stream()),
                          "psb_sparse_pull")
        else:
            out = self._host_pull(idx)
        return out.view(*indices.shape, self.dim)

    def _host_pull(self, idx: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            shards = [self.table]
        else:
            shards = [torch.empty_like(self.table) for _ in range(self.world)]
            dist.all_gather(shards, self.table, group=self.group)  # test-sized tables only
        full = torch.cat(shards, 0)
        ok = (idx >= 0) & (idx < self.vocab)
        out = torch.zeros(idx.numel(), self.dim, dtype=torch.bfloat16)
        out[ok] = full[idx[ok]]
        return out

    # ------------------------------------------------------------------ push
    def stage(self, indices: torch.Tensor, grads: torch.Tensor) -> None:
        """Worker half of a push: this rank's (row, gradient) pairs into its staging range."""
        idx = indices.reshape(-1).to(device=self.device, dtype=torch.int64)
        g = grads.reshape(-1, self.dim)
        n = idx.numel()
        if g.shape[0] != n:
            raise ValueError("one gradient row per index")
        if n > self.capacity:
            raise ValueError(f"{n} rows pushed, staging capacity is {self.capacity}")
        self.metrics.inc("rows_pushed", n)
        self.metrics.inc("push_bytes", n * (8 + self.dim * g.element_size()))
        if self.cuda:
            # one launch: ids, rows (cast to the staging dtype on the way) and the count
            g = g.to(self.device)
            if g.dtype not in (torch.bfloat16, torch.float32):
                g = g.float()
            g, idx = g.contiguous(), idx.contiguous()
            if g.data_ptr() % 16:  # a row slice of a wider tensor: the kernel wants 16-byte vectors
                g = g.clone()
            with torch.cuda.device(self.device), nvtx_range("psb.sparse_stage"):
                check(self.lib.psb_sparse_stage(
                    C.byref(self._c), idx.data_ptr() if n else None, g.data_ptr() if n else None,
                    _lib.BF16 if g.dtype == torch.bfloat16 else _lib.F32, n, self._stream()),
                    "psb_sparse_stage")
        else:
            self._staged = (idx.clone(), g.to(self.grad_dtype))

    def apply(self) -> None:
        """Owner half of a push: mean of the staged gradients of the rows this rank owns
        through the optimizer (``psb_sparse_apply``: mark, accumulate, apply)."""
        self.metrics.inc("pushes")
        if self.cuda:
            with torch.cuda.device(self.device), nvtx_range("psb.sparse_apply"):
                check(self.lib.psb_sparse_apply(C.byref(self._c), self._stream()),
                      "psb_sparse_apply")
                check(self.lib.psb_counter_add(self.step_word.data_ptr(), 1, self._stream()),
                      "psb_counter_add")
        else:
            self._host_apply()
        self.steps += 1

    def _host_apply(self) -> None:
        idx, g = self._staged
        if self.world == 1:
            parts = [(idx, g)]
        else:
            parts = [None] * self.world
            dist.all_gather_object(parts, (idx, g), group=self.group)
        cfg = OptimConfig(**{**self.optim.__dict__, "grad_scale": self.optim.grad_scale / self.world})
        acc = torch.zeros(self.rows_per_rank, self.dim, dtype=torch.float32)
        hit = torch.zeros(self.rows_per_rank, dtype=torch.bool)
        for pidx, pg in parts:                       # fixed order w = 0..N-1
            own = (pidx >= self.lo) & (pidx < self.hi)
            rows = pidx[own] - self.lo
            acc.index_add_(0, rows, pg[own].to(torch.float32))
            hit[rows] = True
        rows = hit.nonzero().flatten()
        if rows.numel():
            p = self.master[rows]
            m = self.m[rows] if self.m is not None else None
            v = self.v[rows] if self.v is not None else None
            ref.optim_step(cfg, p, acc[rows], m, v, step=self.steps)
            self.master[rows] = p
            if m is not None:
                self.m[rows] = m
            if v is not None:
                self.v[rows] = v
            self.table[rows] = p.to(torch.bfloat16)

    def push(self, indices: torch.Tensor, grads: torch.Tensor) -> None:
        """Collective sync-mode push: every rank calls it once per step (an empty ``indices``
        is fine).  Duplicate indices — within one worker or across workers — are summed; the
        sum is divided by the number of workers, like a dense data-parallel mean."""
        if self._emulated:
            raise RuntimeError("emulated ranks: call stage() on every rank, then apply() on every rank")
        self.stage(indices, grads)
        self._barrier()          # every worker's staging is complete and visible
        self.apply()
        self._barrier()          # every owner has rewritten its rows; staging may be reused

    # ------------------------------------------------------------------ autograd front end
    def lookup(self, indices: torch.Tensor) -> torch.Tensor:
        """``pull`` as a differentiable op: the gradient of the returned rows is kept until
        :meth:`apply_gradients` pushes it (one push per step, however many lookups)."""
        return _Lookup.apply(self, indices, torch.empty(0, requires_grad=True, device=self.device))

    def apply_gradients(self) -> None:
        """Collective: push what the backward passes since the last call collected."""
        if self._pending:
            idx = torch.cat([i for i, _ in self._pending])
            g = torch.cat([g for _, g in self._pending])
        else:
            idx = torch.empty(0, dtype=torch.int64, device=self.device)
            g = torch.empty(0, self.dim, dtype=self.grad_dtype, device=self.device)
        self._pending = []
        self.push(idx, g)

    # ------------------------------------------------------------------ state
    def full_table(self) -> torch.Tensor:
        """Every row (bf16) on this rank — checkpoints and tests, not the training path."""
        return self.pull(torch.arange(self.vocab, device=self.device))

    def state_dict(self) -> dict:
        n = self.hi - self.lo
        sd = {"rows": (self.lo, self.hi), "vocab": self.vocab, "dim": self.dim,
              "steps": self.steps, "master": self.master[:n].detach().cpu().clone()}
        if self.m is not None:
            sd["m"] = self.m[:n].detach().cpu().clone()
        if self.v is not None:
            sd["v"] = self.v[:n].detach().cpu().clone()
        return sd

    def load_state_dict(self, sd: dict) -> None:
        if tuple(sd["rows"]) != (self.lo, self.hi) or sd["dim"] != self.dim:
            raise ValueError("state of another shard layout (rows %s, here %s)"
                             % (tuple(sd["rows"]), (self.lo, self.hi)))
        n = self.hi - self.lo
        self.master[:n].copy_(sd["master"])
        if self.m is not None:
            self.m[:n].copy_(sd["m"])
        if self.v is not None:
            self.v[:n].copy_(sd["v"])
        self.steps = int(sd["steps"])
        if self.cuda:
            self.step_word.fill_(self.steps)
        self.sync_table()

    # ------------------------------------------------------------------ whole-table checkpoints
    def _gather_rows(self, t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """Every rank's shard of ``t`` -> the full ``[vocab, dim]`` fp32 tensor on the host."""
        if t is None:
            return None
        if self.world == 1:
            parts = [t]
        elif self.cuda:
            if self._emulated:
                raise RuntimeError("emulated ranks have no collective gather")
            parts = self.fabric.all_gather(t)
        else:
            parts = [torch.empty_like(t) for _ in range(self.world)]
            dist.all_gather(parts, t.contiguous(), group=self.group)
        return torch.cat([p.detach().cpu() for p in parts], 0)[:self.vocab]

    def save(self, path: str, name: str = "embedding", epoch: int = 0) -> None:
        """Collective.  The whole table as ONE tensor in the reference's checkpoint format
        (readable by the reference's ``load_checkpoint`` and by
        ``utils.checkpoint.read_reference_file``) plus the ``.psb`` sidecar with m / v and the
        step count; rank 0 writes, atomically.  Full tensors: loads at any world size."""
        from ..utils.checkpoint import _write_sidecar, write_reference_file
        full = [self._gather_rows(t) for t in (self.master, self.m, self.v)]
        if self.rank == 0:
            write_reference_file(path, epoch, self.steps,
                                 [(name, (self.vocab, self.dim), full[0].numpy())])
            empty = torch.empty(0).numpy()
            _write_sidecar(path + ".psb", self.optim.kind, self.steps,
                           [(full[1].numpy() if full[1] is not None else empty,
                             full[2].numpy() if full[2] is not None else empty)])
        self._barrier()

    def load(self, path: str) -> int:
        """Every rank reads the file and keeps its rows (any world size); optimizer state comes
        back when the sidecar exists and was written by the same optimizer kind.  Returns the
        step count."""
        import os
        from ..utils.checkpoint import _read_sidecar, read_reference_file
        _, steps, tensors = read_reference_file(path)
        name, shape, data = tensors[0]
        if tuple(shape) != (self.vocab, self.dim):
            raise ValueError(f"{path}: table {name!r} is {tuple(shape)}, expected "
                             f"{(self.vocab, self.dim)}")
        n = self.hi - self.lo
        full = torch.from_numpy(data).view(self.vocab, self.dim)
        self.master[:n].copy_(full[self.lo:self.hi])
        if os.path.exists(path + ".psb"):
            kind, steps, states = _read_sidecar(path + ".psb")
            if kind == self.optim.kind and states:
                for mine, saved in zip((self.m, self.v), states[0]):
                    if mine is not None and saved.size == self.vocab * self.dim:
                        mine[:n].copy_(torch.from_numpy(saved).view(self.vocab, self.dim)
                                       [self.lo:self.hi])
        self.steps = int(steps)
        if self.cuda:
            self.step_word.fill_(self.steps)
        self.sync_table()
        return self.steps

    def close(self) -> None:
        for b in getattr(self, "_bufs", []):
            b.free()
        self._bufs = []


class _Lookup(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb: ShardedEmbedding, indices: torch.Tensor, _anchor: torch.Tensor):
        ctx.emb, ctx.indices = emb, indices.reshape(-1)
        return emb.pull(indices)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        emb = ctx.emb
        emb._pending.append((ctx.indices.to(emb.device),
                             grad_out.reshape(-1, emb.dim).to(emb.grad_dtype)))
        return None, None, None
