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

    def owner_ranges(self, slot: ParamSlot) -> List[Tuple[int, int, int]]:
        """Pieces of a tensor by owning shard: ``(rank, start, stop)`` in tensor elements."""
        b = self.buckets[slot.bucket]
        out = []
        pos = slot.offset
        end = slot.offset + slot.numel
        while pos < end:
            r = pos // b.slice_elems
            stop = min(end, (r + 1) * b.slice_elems)
            out.append((r, pos - slot.offset, stop - slot.offset))
            pos = stop
        return out


def _round_up(x: int, q: int) -> int:
    return (x + q - 1) // q * q


def plan_buckets(named_shapes: Sequence[Tuple[str, Sequence[int]]], world: int,
                 bucket_elems: int = 32 << 20, dedicated: Iterable[str] = (),
                 reverse: bool = True) -> BucketPlan:
    """Assign tensors to buckets.

    ``reverse=True`` walks the model back to front so bucket 0 holds the tensors whose
    gradients are produced first by backward.  Names in ``dedicated`` get a bucket of their
    own (the first-layer weight, whose row-slices the fused all-gather GEMM reads)."""
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
        if cur is not None and cur_fill > 0 and cur_fill + numel > bucket_elems:
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
