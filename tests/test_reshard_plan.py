"""Host side of the peer-to-peer reshard (reshard.plan_owners): pure integer planning, CPU only."""

import _cases
from torchsnapshot_b200.manifest import Shard, ShardedTensorEntry, TensorEntry
from torchsnapshot_b200.reshard import cross_rank_bytes, pieces_of, plan_owners


def _entry(boxes, dtype="torch.float32"):
    return ShardedTensorEntry(shards=[
        Shard(offsets=list(o), sizes=list(s),
              tensor=TensorEntry(location=f"sharded/t_{'_'.join(map(str, o))}", serializer="buffer_protocol", dtype=dtype,
                                 shape=list(s), replicated=False))
        for o, s in boxes])


def test_row_wise_8_to_4_needs_no_peer_traffic():
    shape = [800, 64]
    pieces = pieces_of("t", _entry(_cases.chunk_boxes(shape, 0, 8)))
    new = _cases.chunk_boxes(shape, 0, 4)
    boxes_by_rank = [{"t": [b]} for b in new]
    owners = plan_owners(pieces, boxes_by_rank)
    assert owners == [0, 0, 1, 1, 2, 2, 3, 3]  # every saved piece lies inside one new shard
    assert cross_rank_bytes(pieces, owners, boxes_by_rank) == 0


def test_row_to_column_is_balanced_all_to_all():
    shape = [800, 64]
    pieces = pieces_of("t", _entry(_cases.chunk_boxes(shape, 0, 8)))
    boxes_by_rank = [{"t": [b]} for b in _cases.chunk_boxes(shape, 1, 4)]
    owners = plan_owners(pieces, boxes_by_rank)
    assert sorted(owners) == [0, 0, 1, 1, 2, 2, 3, 3]  # ties are spread, not piled on rank 0
    total = 800 * 64 * 4
    assert cross_rank_bytes(pieces, owners, boxes_by_rank) == total * 3 // 4


def test_unneeded_pieces_have_no_owner_and_plan_is_deterministic():
    shape = [100, 10]
    pieces = pieces_of("t", _entry(_cases.chunk_boxes(shape, 0, 5)))
    boxes_by_rank = [{"t": [([0, 0], [30, 10])]}, {"t": [([30, 0], [10, 10])]}]
    owners = plan_owners(pieces, boxes_by_rank)
    # rows 20..40 are needed half by rank 0, half by rank 1: a tie, so the less loaded rank 1 owns them
    assert owners == [0, 1, None, None, None]
    assert plan_owners(pieces, boxes_by_rank) == owners


def _emulate(mem, regions):
    """Byte-level emulation of tsb_copy_regions on a numpy buffer (addresses are offsets into `mem`)."""
    import itertools

    for src, dst, run, sizes, sstr, dstr in regions:
        for idx in itertools.product(*[range(n) for n in sizes]):
            so = src + sum(i * s for i, s in zip(idx, sstr))
            do = dst + sum(i * s for i, s in zip(idx, dstr))
            mem[do : do + run] = mem[so : so + run]


def test_scatter_units_are_the_partitioned_ones():
    from torchsnapshot_b200.reshard import scatter_keys

    replicated, resharded, lone = ("slab", 0, 100), ("piece", -1, -1), ("piece2", -1, -1)
    needs = [
        {replicated: (100, 100), resharded: (60, 100), lone: (100, 100)},
        {replicated: (100, 100), resharded: (40, 100)},
        {replicated: (100, 100)},
    ]
    assert scatter_keys(needs) == sorted([resharded, lone])


def test_packing_per_destination_then_unpacking_equals_the_direct_strided_copy():
    """The owner packs a destination's boxes back to back (packed_layout on the REQUEST), the destination
    unpacks them with its own strides (packed_layout on the same request): together they must move exactly
    the bytes the one-step region would."""
    import numpy as np

    from torchsnapshot_b200.regions import box_regions
    from torchsnapshot_b200.reshard import packed_layout, packed_strides

    rs = np.random.RandomState(7)
    mem = rs.randint(0, 256, size=1 << 17).astype(np.uint8)
    unit_a, unit_b, dst, buf = 0, 30000, 70000, 100000
    # destination tensor [40, 8] fp32 = columns 8..16 of unit A ([40, 32] fp32) ...
    regs_a = box_regions(unit_a + 8 * 4, [32 * 4, 4], dst, [8 * 4, 4], [40, 8], 4)
    # ... and a [3, 5, 6] box of unit B ([4, 9, 10] int16) into a permuted destination view
    regs_b = box_regions(unit_b + (1 * 90 + 2 * 10 + 3) * 2, [180, 20, 2], dst + 4000, [2, 6 * 2 * 3, 6], [3, 5, 6], 2)
    direct = mem.copy()
    _emulate(direct, regs_a + regs_b)
    key_a, key_b = ("a", -1, -1), ("b", 0, 720)
    request = {k: [(r[0] - base, r[2], tuple(r[3]), tuple(r[4])) for r in regs]
               for k, regs, base in ((key_a, regs_a, unit_a), (key_b, regs_b, unit_b))}
    offsets, total = packed_layout(request)
    assert total % 512 == 0 and len(offsets) == len(regs_a) + len(regs_b)
    staged = {key_a: unit_a, key_b: unit_b}
    pack = [(staged[k] + off, buf + offsets[(k, i)], run, sizes, sstr, packed_strides(run, sizes))
            for k, reqs in request.items() for i, (off, run, sizes, sstr) in enumerate(reqs)]
    unpack = [(buf + offsets[(k, i)], r[1], r[2], tuple(r[3]), packed_strides(r[2], r[3]), tuple(r[5]))
              for k, regs in ((key_a, regs_a), (key_b, regs_b)) for i, r in enumerate(regs)]
    two_step = mem.copy()
    _emulate(two_step, pack)
    _emulate(two_step, unpack)
    assert np.array_equal(two_step[dst:buf], direct[dst:buf])
    assert np.array_equal(two_step[:dst], mem[:dst])  # sources untouched
