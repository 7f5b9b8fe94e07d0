"""Flat bucketed parameter space: ``name -> (bucket, offset, shape)``.

The reference keeps the model as an ordered ``vector<tensor>`` on ONE server and matches
gradients to parameters positionally (src/parameter_server.cpp:83-85).  Here every bucket is
split evenly across all shards (ZeRO-style): shard ``r`` owns elements
``[r * slice, (r + 1) * slice)`` of every bucket, which balances NVLink ingress per GPU.
The name table preserves the reference's named-tensor view for the gRPC services and the
checkpoint format (App. B of SURVEY.md).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

ALIGN_ELEMS = 128       # every tensor starts on a 256-byte boundary (bf16)
SLICE_QUANTUM = 512     # slice_elems is a multiple of this (1 KB of bf16, 16 B vectors)


@dataclass
class ParamSlot:
    name: str
    shape: Tuple[int, ...]
    numel: int
    bucket: int
    offset: int          # elements from the start of the bucket
    index: int           # position in the model's own parameter order
    # Row-split tensors: the tensor's own slot (``plan.slots[name]``) lists its ``pieces``; each
    # piece is a slot of its own bucket (``bucket.slots``) covering rows [row_lo, row_hi) and
    # naming the tensor in ``parent``.  The pieces' buckets are consecutive and unpadded, so the
    # tensor stays one contiguous range of the flat space.
    parent: Optional[str] = None
    row_lo: int = 0
    row_hi: int = 0
    pieces: List["ParamSlot"] = field(default_factory=list)


@dataclass
class Bucket:
    index: int
    base: int            # element offset of the bucket in the flat space
    numel: int           # padded: world * slice_elems
    slice_elems: int
    slots: List[ParamSlot] = field(default_factory=list)
    dedicated: bool = False

    def shard_base(self, world: int) -> int:
        """Offset of this bucket's slice inside a shard's flat state (master / m / v)."""
        return self.base // world


@dataclass
class BucketPlan:
    world: int
    buckets: List[Bucket]
    slots: Dict[str, ParamSlot]

    @property
    def total_elems(self) -> int:
        return sum(b.numel for b in self.buckets)

    @property
    def shard_elems(self) -> int:
        return self.total_elems // self.world

    def shard_ranges(self, slot: ParamSlot) -> List[Tuple[int, int, int, int]]:
        """Where a tensor lives: ``(rank, start, stop, shard_offset)`` — tensor elements
        ``[start, stop)`` are elements ``shard_offset ...`` of rank's flat shard state."""
        out = []
        row_elems = slot.numel // slot.shape[0] if slot.shape else slot.numel
        for p in (slot.pieces or [slot]):
            b = self.buckets[p.bucket]
            sb = b.shard_base(self.world)
            t0 = p.row_lo * row_elems if p.parent else 0
            pos, end = p.offset, p.offset + p.numel
            while pos < end:
                r = pos // b.slice_elems
                stop = min(end, (r + 1) * b.slice_elems)
                out.append((r, t0 + pos - p.offset, t0 + stop - p.offset,
                            sb + pos - r * b.slice_elems))
                pos = stop
        return out

    def owner_ranges(self, slot: ParamSlot) -> List[Tuple[int, int, int]]:
        """Pieces of a tensor by owning shard: ``(rank, start, stop)`` in tensor elements."""
        return [(r, a, b) for r, a, b, _ in self.shard_ranges(slot)]


def _round_up(x: int, q: int) -> int:
    return (x + q - 1) // q * q


def split_rows(shape: Sequence[int], pieces: int, world: int) -> Optional[int]:
    """Rows per piece if a tensor of ``shape`` can be cut into ``pieces`` row blocks that each
    fill a bucket exactly (no padding between them), else None."""
    if pieces < 2 or len(shape) != 2 or shape[0] % pieces:
        return None
    rows = shape[0] // pieces
    if rows % 8 or (rows * shape[1]) % (world * SLICE_QUANTUM):
        return None
    return rows


def plan_buckets(named_shapes: Sequence[Tuple[str, Sequence[int]]], world: int,
                 bucket_elems: int = 32 << 20, dedicated: Iterable[str] = (),
                 reverse: bool = True, split: Optional[Dict[str, int]] = None) -> BucketPlan:
    """Assign tensors to buckets.

    ``reverse=True`` walks the model back to front so bucket 0 holds the tensors whose
    gradients are produced first by backward.  Names in ``dedicated`` get a bucket of their
    own (the first-layer weight, whose row-slices the fused all-gather GEMM reads).
    ``split``: ``name -> k`` cuts a 2-D tensor into ``k`` row blocks, one bucket each, so that
    the layer can produce and push its gradient block by block (the push of block *i* overlaps
    the dW GEMM of block *i+1*); requests that do not divide evenly are ignored."""
    if world < 1:
        raise ValueError("world must be >= 1")
    dedicated = set(dedicated)
    order = list(enumerate(named_shapes))
    if reverse:
        order = order[::-1]
    buckets: List[Bucket] = []
    slots: Dict[str, ParamSlot] = {}
    cur: Optional[Bucket] = None
    cur_fill = 0
    base = 0
    quantum = world * SLICE_QUANTUM

    def close():
        nonlocal cur, cur_fill, base
        if cur is None:
            return
        cur.numel = _round_up(max(cur_fill, 1), quantum)
        cur.slice_elems = cur.numel // world
        base += cur.numel
        buckets.append(cur)
        cur, cur_fill = None, 0

    for index, (name, shape) in order:
        numel = 1
        for d in shape:
            numel *= int(d)
        if name in slots:
            raise ValueError(f"duplicate tensor name {name!r}")
        if name in dedicated:
            close()
            cur = Bucket(len(buckets), base, 0, 0, dedicated=True)
            slot = ParamSlot(name, tuple(int(d) for d in shape), numel, cur.index, 0, index)
            cur.slots.append(slot)
            slots[name] = slot
            cur_fill = numel
            close()
            continue
        rows = split_rows(shape, (split or {}).get(name, 0), world)
        tiny = bucket_elems // 64  # a bucket this small is not worth a push of its own
        if rows is not None:
            # a few small tensors waiting in the open bucket (the layer's bias) ride with the
            # LAST row block instead of costing a push of their own
            carried = []
            if cur is not None and 0 < cur_fill <= tiny:
                carried, cur, cur_fill = cur.slots, None, 0
            close()
            shape_t = tuple(int(d) for d in shape)
            whole = ParamSlot(name, shape_t, numel, len(buckets), 0, index)
            n_pieces = shape_t[0] // rows
            for k in range(n_pieces):
                cur = Bucket(len(buckets), base, 0, 0)
                piece = ParamSlot(f"{name}[{k * rows}:{(k + 1) * rows}]", (rows, shape_t[1]),
                                  rows * shape_t[1], cur.index, 0, index, parent=name,
                                  row_lo=k * rows, row_hi=(k + 1) * rows)
                cur.slots.append(piece)
                whole.pieces.append(piece)
                cur_fill = piece.numel
                if k == n_pieces - 1:
                    for c in carried:  # after the block: padding behind it harms nobody
                        c.bucket, c.offset = cur.index, _round_up(cur_fill, ALIGN_ELEMS)
                        cur.slots.append(c)
                        cur_fill = c.offset + c.numel
                close()
            slots[name] = whole
            continue
        if cur is not None and cur_fill > tiny and cur_fill + numel > bucket_elems:
            close()
        if cur is None:
            cur = Bucket(len(buckets), base, 0, 0)
        off = _round_up(cur_fill, ALIGN_ELEMS)
        slot = ParamSlot(name, tuple(int(d) for d in shape), numel, cur.index, off, index)
        cur.slots.append(slot)
        slots[name] = slot
        cur_fill = off + numel
    close()
    return BucketPlan(world, buckets, slots)
