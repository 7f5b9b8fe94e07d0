"""Bucket plan: name -> (bucket, offset) table, alignment, ownership ranges."""
import pytest

from parameter_server_distributed_b200.parallel.buckets import (ALIGN_ELEMS, SLICE_QUANTUM,
                                                                plan_buckets)


def shapes():
    return [("a.weight", (300, 70)), ("a.bias", (300,)), ("b.weight", (1000, 300)),
            ("b.bias", (1000,)), ("c.weight", (10, 1000))]


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_plan_invariants(world):
    plan = plan_buckets(shapes(), world, bucket_elems=200_000)
    seen = set()
    base = 0
    for b in plan.buckets:
        assert b.base == base
        assert b.numel == b.slice_elems * world
        assert b.slice_elems % SLICE_QUANTUM == 0
        base += b.numel
        end = 0
        for s in b.slots:
            assert s.offset % ALIGN_ELEMS == 0
            assert s.offset >= end
            end = s.offset + s.numel
            seen.add(s.name)
        assert end <= b.numel
    assert seen == {n for n, _ in shapes()}
    assert plan.total_elems == base and plan.shard_elems * world == base


def test_reverse_order_and_dedicated():
    plan = plan_buckets(shapes(), 2, bucket_elems=10**9, dedicated=["a.weight"])
    # back-to-front: the last layer lands in bucket 0
    assert plan.slots["c.weight"].bucket == 0
    d = plan.buckets[plan.slots["a.weight"].bucket]
    assert d.dedicated and len(d.slots) == 1 and plan.slots["a.weight"].offset == 0


def test_owner_ranges_cover_tensor():
    plan = plan_buckets(shapes(), 4, bucket_elems=10**9)
    for slot in plan.slots.values():
        pieces = plan.owner_ranges(slot)
        assert pieces[0][1] == 0 and pieces[-1][2] == slot.numel
        for (r0, a0, b0), (r1, a1, b1) in zip(pieces, pieces[1:]):
            assert b0 == a1 and r1 == r0 + 1


def test_duplicate_name_rejected():
    with pytest.raises(ValueError):
        plan_buckets([("w", (4,)), ("w", (4,))], 1)


def test_row_split_gives_contiguous_unpadded_buckets():
    world = 2
    shp = [("first.weight", (4096, 1024)), ("first.bias", (4096,)), ("last.weight", (64, 4096))]
    plan = plan_buckets(shp, world, bucket_elems=10**9, split={"first.weight": 4, "last.weight": 3})
    whole = plan.slots["first.weight"]
    assert len(whole.pieces) == 4 and not plan.slots["last.weight"].pieces  # 64 % 3 != 0: ignored
    pos = plan.buckets[whole.bucket].base
    for k, piece in enumerate(whole.pieces):
        b = plan.buckets[piece.bucket]
        assert b.slots[0] is piece and piece.offset == 0
        if k < 3:
            assert b.slots == [piece] and b.numel == piece.numel  # no padding between the blocks
        else:  # the small tensors that were waiting ride behind the last block
            assert [s.name for s in b.slots[1:]] == ["last.weight", "first.bias"]
            assert all(s.bucket == b.index and s.offset >= piece.numel for s in b.slots[1:])
        assert b.base == pos and (piece.row_lo, piece.row_hi) == (k * 1024, (k + 1) * 1024)
        assert piece.parent == "first.weight"
        pos += b.numel
    # ownership covers the tensor exactly once, in order, over the pieces
    rngs = plan.shard_ranges(whole)
    assert rngs[0][1] == 0 and rngs[-1][2] == whole.numel
    assert all(b0 == a1 for (_, _, b0, _), (_, a1, _, _) in zip(rngs, rngs[1:]))
    assert [r for r, *_ in rngs] == [0, 1] * 4
    seen = {s.name for b in plan.buckets for s in b.slots}
    assert "first.bias" in seen and "last.weight" in seen and "first.weight" not in seen


def test_heap_allocator_is_deterministic_and_reuses_freed_ranges():
    """The symmetric heap's sub-allocator (offsets only): first fit, merge on free, shrink at the
    end — and two instances fed the same calls agree, which is what keeps the heap symmetric."""
    from parameter_server_distributed_b200.parallel.fabric import HeapAllocator
    a, b = HeapAllocator(1 << 20), HeapAllocator(1 << 20)
    log = []
    for h in (a, b):
        p = h.alloc(1000, 4096)
        g = h.alloc(100_000, 4096)
        q = h.alloc(50_000, 256)
        h.free(g, 100_000)
        r = h.alloc(60_000, 4096)          # fits into the hole g left
        s = h.alloc(30_000, 256)           # so does this, behind r
        t = h.alloc(200_000, 4096)         # this does not: goes to the end
        log.append((p, g, q, r, s, t, h.cursor))
    assert log[0] == log[1]
    p, g, q, r, s, t, top = log[0]
    assert p == 0 and g == 4096 and r == g and g + 60_000 <= s < g + 100_000 and t >= q + 50_000
    spans = sorted([(p, 1000), (q, 50_000), (r, 60_000), (s, 30_000), (t, 200_000)])
    assert all(x + n <= y for (x, n), (y, _) in zip(spans, spans[1:]))  # no overlap
    a.free(t, 200_000)                     # trailing range: the high-water mark drops
    assert a.cursor < top and a.largest_free() >= (1 << 20) - a.cursor
    for off, n in ((s, 30_000), (r, 60_000), (q, 50_000), (p, 1000)):
        a.free(off, n)
    assert a.cursor == 0 and a.holes == []
    with pytest.raises(MemoryError):
        a.alloc((1 << 20) + 1)
