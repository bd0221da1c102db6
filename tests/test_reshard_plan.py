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
