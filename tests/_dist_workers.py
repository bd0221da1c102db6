"""Worker bodies for the world_size-2 gloo tests (spawned processes import this module by name)."""

import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch
import torch.distributed as dist

import tsb200_loader

tsb200_loader.load()


def _init(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)


def run(rank, world, port, fn_name, tmp, errq):
    try:
        _init(rank, world, port)
        globals()[fn_name](rank, world, tmp)
        dist.barrier()
        dist.destroy_process_group()
    except Exception:
        errq.put(f"rank {rank}:\n{traceback.format_exc()}")
        raise


def _state(seed):
    import _cases

    return {
        "w": _cases.raw_tensor(seed, [300, 40], torch.float32),
        "b": _cases.raw_tensor(seed + 1, [77], torch.bfloat16 if False else torch.float16),
        "big": _cases.raw_tensor(seed + 2, [64, 100], torch.float64),
        "n": 7,
    }


def replicated_partition(rank, world, tmp):
    """DDP-style replicated state: every replicated byte is written exactly once, by some rank."""
    import json

    from torchsnapshot_b200 import knobs
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    shared = StateDict(_state(100))
    own = StateDict(r=torch.full((10,), float(rank)))
    path = os.path.join(tmp, "snap")
    with knobs.override_max_chunk_size_bytes(20000), knobs.override_slab_size_threshold_bytes(9000):
        snap = Snapshot.take(path, {"shared": shared, "own": own}, replicated=["shared/**"])
    dist.barrier()
    meta = json.load(open(os.path.join(path, ".snapshot_metadata")))["manifest"]
    assert meta["0/shared/w"]["type"] == "ChunkedTensor" and meta["0/shared/w"]["replicated"] is True
    assert "1/shared/w" not in meta  # replicated entries are listed once, under rank 0
    assert meta["0/own/r"]["replicated"] is False and "1/own/r" in meta
    # every chunk of the replicated tensor exists exactly once on disk, written by either rank
    chunks = meta["0/shared/w"]["chunks"]
    assert [c["offsets"] for c in chunks] == sorted(c["offsets"] for c in chunks)
    for c in chunks:
        assert os.path.exists(os.path.join(path, c["tensor"]["location"]))
    # both ranks restore the replicated state and their own state
    out_shared = StateDict({k: (torch.zeros_like(v) if isinstance(v, torch.Tensor) else 0) for k, v in shared.items()})
    out_own = StateDict(r=torch.zeros(10))
    snap.restore({"shared": out_shared, "own": out_own})
    for k, v in shared.items():
        if isinstance(v, torch.Tensor):
            assert torch.equal(out_shared[k], v), k
    assert out_shared["n"] == 7 and torch.equal(out_own["r"], own["r"])
    # the write load was split: both ranks wrote payload files of the replicated state
    if rank == 0:
        import glob

        files = glob.glob(os.path.join(path, "replicated", "**", "*"), recursive=True) + glob.glob(
            os.path.join(path, "batched", "*"))
        assert len([f for f in files if os.path.isfile(f)]) >= 2


def sharded_reshard(rank, world, tmp):
    """ShardedTensor saved with rows split over the ranks, restored with columns split over them
    (every rank needs bytes of every persisted piece), and read back as a dense tensor."""
    import _cases
    from torch.distributed._shard import sharded_tensor
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec

    from torchsnapshot_b200 import knobs
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    full = _cases.raw_tensor(500, [48, 36], torch.float32)
    src_spec = ChunkShardingSpec(dim=0, placements=[f"rank:{r}/cpu" for r in range(world)])
    dst_spec = ChunkShardingSpec(dim=1, placements=[f"rank:{r}/cpu" for r in range(world)])
    src = sharded_tensor.empty(src_spec, 48, 36)
    for shard in src.local_shards():
        shard.tensor.copy_(_cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes)))
    path = os.path.join(tmp, "snap")
    with knobs.override_max_shard_size_bytes(5 * 36 * 4):
        snap = Snapshot.take(path, {"emb": StateDict(t=src)})
    m = snap.get_manifest()
    pieces = [s for r in range(world) for s in m[f"{r}/emb/t"].shards]
    assert len(pieces) == 10 and all(s.tensor.location.startswith("sharded/emb/t_") or s.tensor.location.startswith("batched/")
                                     for s in pieces)
    dst = sharded_tensor.empty(dst_spec, 48, 36)
    snap.restore({"emb": StateDict(t=dst)})
    for shard in dst.local_shards():
        assert torch.equal(shard.tensor, _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes)))
    dense = torch.zeros(48, 36)
    got = snap.read_object("0/emb/t", obj_out=dense)
    assert torch.equal(got, full)


def async_commit(rank, world, tmp):
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    state = StateDict(t=torch.arange(1000, dtype=torch.float32) + rank)
    pending = Snapshot.async_take(os.path.join(tmp, "snap"), {"s": state})
    state["t"].zero_()
    snap = pending.wait()
    out = StateDict(t=torch.zeros(1000))
    snap.restore({"s": out})
    assert torch.equal(out["t"], torch.arange(1000, dtype=torch.float32) + rank)
    assert os.path.exists(os.path.join(tmp, "snap", ".snapshot_metadata"))


def _flat_sharded(world, rank_boxes_only=False):
    """FSDP-style flat parameter: a 1-D ShardedTensor chunked over the ranks (BASELINE config 3 layout)."""
    from torch.distributed._shard import sharded_tensor
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec

    spec = ChunkShardingSpec(dim=0, placements=[f"rank:{r}/cpu" for r in range(world)])
    return sharded_tensor.empty(spec, 10007, dtype=torch.float32)


def elastic_take(rank, world, tmp):
    import _cases
    from torchsnapshot_b200 import knobs
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    full = _cases.raw_tensor(900, [10007], torch.float32)
    st = _flat_sharded(world)
    for shard in st.local_shards():
        o, n = shard.metadata.shard_offsets[0], shard.metadata.shard_sizes[0]
        shard.tensor.copy_(full[o : o + n])
    rep = StateDict(w=_cases.raw_tensor(901, [33, 7], torch.float32), step=5)
    with knobs.override_max_shard_size_bytes(4000):
        Snapshot.take(os.path.join(tmp, "snap"), {"flat": StateDict(p=st), "rep": rep}, replicated=["rep/**"])


def elastic_restore(rank, world, tmp):
    """Restore at a different world size than the snapshot was taken with (8->4 / 4->8 in miniature)."""
    import _cases
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    full = _cases.raw_tensor(900, [10007], torch.float32)
    st = _flat_sharded(world)
    rep = StateDict(w=torch.zeros(33, 7), step=0)
    snap = Snapshot(os.path.join(tmp, "snap"))
    assert snap.metadata.world_size == 2
    snap.restore({"flat": StateDict(p=st), "rep": rep})
    for shard in st.local_shards():
        o, n = shard.metadata.shard_offsets[0], shard.metadata.shard_sizes[0]
        assert torch.equal(shard.tensor, full[o : o + n]), (rank, o, n)
    assert torch.equal(rep["w"], _cases.raw_tensor(901, [33, 7], torch.float32)) and rep["step"] == 5


def object_collectives(rank, world, tmp):
    """PGWrapper.scatter_object_list (reference pg_wrapper.py:58-89) and the collective form of the
    replicated-entry consolidation (partitioner.py:306-316) over a real 2-rank group."""
    from torchsnapshot_b200.manifest import ChunkedTensorEntry, Shard, TensorEntry
    from torchsnapshot_b200.partitioner import consolidate_replicated_entries_dist
    from torchsnapshot_b200.pg_wrapper import PGWrapper

    pg = PGWrapper()
    out = [None]
    pg.scatter_object_list(out, [{"for": r, "payload": list(range(r + 3))} for r in range(world)] if rank == 0 else None)
    assert out[0] == {"for": rank, "payload": list(range(rank + 3))}
    if rank == 0:
        try:
            pg.scatter_object_list([None], [1])  # wrong length on the source rank
        except RuntimeError:
            pass
        else:
            raise AssertionError("a short input_list must be refused")

    def chunk(i):
        t = TensorEntry(location=f"replicated/w_{i}_0", serializer="buffer_protocol", dtype="torch.float32",
                        shape=[2, 4], replicated=False)
        return Shard(offsets=[2 * i, 0], sizes=[2, 4], tensor=t)

    # each rank wrote one chunk of a replicated chunked tensor, plus a private entry of its own
    mine = {
        "m/w": ChunkedTensorEntry(dtype="torch.float32", shape=[4, 4], chunks=[chunk(rank)], replicated=True),
        f"m/private_{rank}": TensorEntry(location=f"{rank}/m/private", serializer="buffer_protocol",
                                         dtype="torch.int64", shape=[3], replicated=False),
    }
    merged = consolidate_replicated_entries_dist(mine, pg)
    assert f"m/private_{rank}" in merged and f"m/private_{1 - rank}" not in merged
    if rank == 0:  # replicated entries are listed once, under rank 0, with every rank's chunks in offset order
        assert [c.offsets for c in merged["m/w"].chunks] == [[0, 0], [2, 0]]
    else:
        assert "m/w" not in merged
    keep_all = consolidate_replicated_entries_dist(mine, pg, dedup=False)
    assert [c.offsets for c in keep_all["m/w"].chunks] == [[0, 0], [2, 0]]


def slot_gather(rank, world, tmp):
    """PGWrapper.all_gather_object: small objects in one fixed-slot collective, anything longer -- on ANY rank --
    through dist.all_gather_object on every rank, the same answers either way; a sub-group has slots of its own."""
    import os
    import pickle

    import torch.distributed as dist

    from torchsnapshot_b200 import pg_wrapper
    from torchsnapshot_b200.pg_wrapper import PGWrapper

    pg = PGWrapper()
    calls = {"fallback": 0}
    plain = dist.all_gather_object

    def counting(*a, **k):
        calls["fallback"] += 1
        return plain(*a, **k)

    dist.all_gather_object = counting
    try:
        def gather(obj, **kw):
            out = [None] * world
            pg.all_gather_object(out, obj, **kw)
            return out

        assert gather(("hello", rank, {"x": [rank] * 3})) == [("hello", r, {"x": [r] * 3}) for r in range(world)]
        assert gather(None) == [None] * world and gather(b"") == [b""] * world
        assert calls["fallback"] == 0
        # the longest pickle that fits, and one byte more, on the last rank only
        room = pg_wrapper._SLOT_BYTES - 8
        overhead = len(pickle.dumps(b"", protocol=pickle.HIGHEST_PROTOCOL)) + 8  # generous: framing grows with length
        fits = b"x" * (room - overhead)
        assert len(pickle.dumps(fits, protocol=pickle.HIGHEST_PROTOCOL)) <= room
        assert gather(fits if rank == world - 1 else rank) == list(range(world - 1)) + [fits]
        assert calls["fallback"] == 0
        big = b"y" * (room + 1)
        assert gather(big if rank == world - 1 else rank) == list(range(world - 1)) + [big]
        assert calls["fallback"] == 1  # every rank fell back, because one rank's object was long
        assert gather({"rank": rank}, large=True) == [{"rank": r} for r in range(world)]
        assert calls["fallback"] == 2
        os.environ["TSB200_SLOT_GATHER"] = "0"
        assert gather(rank) == list(range(world)) and calls["fallback"] == 3
        del os.environ["TSB200_SLOT_GATHER"]
        # many rounds: the buffers are reused
        for i in range(50):
            assert gather((i, rank)) == [(i, r) for r in range(world)]
        assert calls["fallback"] == 3
        if world >= 4:
            sub = dist.new_group(ranks=[1, 3])
            if rank in (1, 3):
                spg = PGWrapper(sub)
                out = [None, None]
                spg.all_gather_object(out, f"sub{rank}")
                assert out == ["sub1", "sub3"]
            assert gather("after") == ["after"] * world
    finally:
        dist.all_gather_object = plain


def elastic_roundtrip_cpu(rank, world, tmp):
    """Taken by `world` ranks, restored by world/2 (a sub-group of the same job), taken again by those, restored
    by all: flat parameters in two dtypes, a row-wise table, replicated state -- on CPU tensors over gloo."""
    import _elastic

    info = _elastic.elastic_roundtrip(tmp, "cpu")
    assert info["world"] == world and info["half"] == max(world // 2, 1)


def replicated_paths_differ(rank, world, tmp):
    """`replicated=["**"]` with state dicts that differ between the ranks: only the paths every rank has are
    replicated (reference snapshot.py:796-822).  The ranks find out in the partitioner's exchange (the digests of
    their candidates differ), then exchange the candidate lists and plan again."""
    import _cases
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    common = _cases.raw_tensor(31, [50, 9], torch.float32)
    mine = _cases.raw_tensor(32 + rank, [20 + rank], torch.float32)
    state = StateDict(common=common, **{f"only_{rank}": mine})
    snap = Snapshot.take(os.path.join(tmp, "snap"), {"s": state}, replicated=["**"])
    m = snap.get_manifest()
    assert m["0/s/common"].replicated and "1/s/common" not in m
    assert not m[f"{rank}/s/only_{rank}"].replicated
    assert f"{1 - rank}/s/only_{rank}" not in m
    out = StateDict(common=torch.zeros_like(common), **{f"only_{rank}": torch.zeros_like(mine)})
    snap.restore({"s": out})
    assert torch.equal(out["common"], common) and torch.equal(out[f"only_{rank}"], mine)


def repeated_replicated_takes(rank, world, tmp):
    """A job that checkpoints the same replicated tensors again and again repeats its previous plan (no
    planning, same files from the same ranks) as long as every rank can and the split of the writes still
    stands; new values in the same tensors are what gets saved; a replaced tensor on ONE rank makes every rank
    plan afresh.  Every snapshot restores to what the state held when it was taken."""
    import _cases
    from torchsnapshot_b200 import knobs
    from torchsnapshot_b200.snapshot import Snapshot, StateDict, _RepeatTake

    state = StateDict({f"p{i}": _cases.raw_tensor(300 + i, [64 + i, 9], torch.float32) for i in range(30)})
    state["epoch"] = 0
    own = StateDict(r=torch.full((7,), float(rank)))

    def check(snap, want, epoch):
        out = StateDict({k: (torch.zeros_like(v) if isinstance(v, torch.Tensor) else -1) for k, v in want.items()})
        out_own = StateDict(r=torch.zeros(7))
        snap.restore({"m": out, "own": out_own})
        for k, v in want.items():
            if isinstance(v, torch.Tensor):
                assert torch.equal(out[k], v), k
        assert out["epoch"] == epoch and torch.equal(out_own["r"], own["r"])

    snaps, wants = [], []
    with knobs.override_slab_size_threshold_bytes(6000):
        for epoch in range(4):
            state["epoch"] = epoch
            for v in state.values():
                if isinstance(v, torch.Tensor):
                    v.add_(1.0)
            before = _RepeatTake.repeats
            snaps.append(Snapshot.take(os.path.join(tmp, f"s{epoch}"), {"m": state, "own": own}, replicated=["m/**"]))
            wants.append({k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in state.items()})
            # per-rank entries next to the replicated ones: planned once, repeated from the second take on
            assert _RepeatTake.repeats - before == (1 if epoch else 0), epoch
        for epoch, (snap, want) in enumerate(zip(snaps, wants)):
            check(snap, want, epoch)
        # a purely replicated state dict: another plan (first take), then repeats
        took = []
        for epoch in range(4):
            state["epoch"] = 10 + epoch
            for v in state.values():
                if isinstance(v, torch.Tensor):
                    v.mul_(1.5)
            if epoch == 3 and rank == 1:
                state["p5"] = state["p5"].clone()  # one rank moved one tensor: nobody repeats
            before = _RepeatTake.repeats
            snap = Snapshot.take(os.path.join(tmp, f"t{epoch}"), {"m": state}, replicated=["**"])
            took.append(_RepeatTake.repeats - before)
            want = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in state.items()}
            out = StateDict({k: (torch.zeros_like(v) if isinstance(v, torch.Tensor) else -1) for k, v in want.items()})
            snap.restore({"m": out})
            for k, v in want.items():
                assert (torch.equal(out[k], v) if isinstance(v, torch.Tensor) else out[k] == v), (epoch, k)
            m = snap.get_manifest()
            assert m["0/m/p3"].replicated and "1/m/p3" not in m and m["0/m/epoch"].get_value() == 10 + epoch
        assert took == [0, 1, 1, 0], took
        # When no rank's manifest changed, only digests travel and the previous metadata document stands; a rank
        # that has dropped its caches (or a changed value anywhere) makes every rank exchange manifests again.
        from torchsnapshot_b200 import snapshot as snapmod

        def take(name):
            snap = Snapshot.take(os.path.join(tmp, name), {"m": state}, replicated=["**"])
            out = StateDict({k: (torch.zeros_like(v) if isinstance(v, torch.Tensor) else -1) for k, v in state.items()})
            snap.restore({"m": out})
            for k, v in state.items():
                assert (torch.equal(out[k], v) if isinstance(v, torch.Tensor) else out[k] == v), (name, k)
            return snap

        a, b = take("same_a"), take("same_b")
        assert b.metadata is a.metadata
        if rank == 1:
            snapmod.clear_plan_caches()
        c = take("same_c")
        assert c.metadata is not b.metadata and c.get_manifest().keys() == b.get_manifest().keys()
        d = take("same_d")  # (the rank that dropped its caches planned afresh: new slab names, new document)
        e = take("same_e")
        assert e.metadata is d.metadata
        state["epoch"] = 99
        f = take("same_f")
        assert f.metadata is not e.metadata and f.get_manifest()["0/m/epoch"].get_value() == 99
