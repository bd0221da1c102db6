"""Worker bodies for the multi-GPU tests (one process per GPU, NCCL process group)."""

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


def run(rank, world, port, fn_name, tmp, errq):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        globals()[fn_name](rank, world, tmp)
        dist.barrier(device_ids=[rank])
        dist.destroy_process_group()
    except Exception:
        errq.put(f"rank {rank}:\n{traceback.format_exc()}")
        raise


def _sharded(spec_dim, world, shape, full=None):
    import _cases
    from torch.distributed._shard import sharded_tensor
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec

    spec = ChunkShardingSpec(dim=spec_dim, placements=[f"rank:{r}/cuda:{r}" for r in range(world)])
    st = sharded_tensor.empty(spec, *shape, dtype=torch.float32)
    for shard in st.local_shards():
        if full is not None:
            shard.tensor.copy_(_cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes)))
        else:
            shard.tensor.zero_()
    return st


def p2p_reshard(rank, world, tmp):
    """Rows saved, columns restored: every rank needs a slice of every persisted piece.  The P2P path
    (each piece read once, the rest over NVLink) must give the bytes of the storage-only path."""
    import _cases
    from torchsnapshot_b200 import knobs, reshard
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    shape = [1024, 640]
    full = _cases.raw_tensor(77, shape, torch.float32)
    src = _sharded(0, world, shape, full)
    own = StateDict(x=torch.full((5,), float(rank), device="cuda"))
    path = os.path.join(tmp, "snap")
    with knobs.override_max_shard_size_bytes(100 * 640 * 4):
        snap = Snapshot.take(path, {"emb": StateDict(t=src), "own": own})
    # P2P restore into a column-wise layout
    dst = _sharded(1, world, shape)
    out_own = StateDict(x=torch.zeros(5, device="cuda"))
    snap.restore({"emb": StateDict(t=dst), "own": out_own})
    st = reshard.last_stats
    assert st is not None and st.launches == 1 and st.peer_bytes > 0, vars(st) if st else None
    total = 1024 * 640 * 4
    assert st.local_bytes + st.peer_bytes == total // world
    for shard in dst.local_shards():
        want = _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes)).cuda()
        assert torch.equal(shard.tensor, want)
    assert torch.equal(out_own["x"], own["x"])
    # a reshard goes through per-destination buffers (the owner packed this rank's columns for it) ...
    assert st.pack_launches > 0, vars(st)
    assert abs(st.amplification["sharded"] - world) < 1e-6, st.amplification  # every rank would read every piece
    # ... and the staging blocks and peer mappings of this restore serve the next one: nothing new from the driver
    dst0 = _sharded(1, world, shape)
    snap.restore({"emb": StateDict(t=dst0), "own": out_own})
    st0 = reshard.last_stats
    assert st0.new_staging_bytes == 0 and st0.opens == 0 and st0.peer_bytes == st.peer_bytes, vars(st0)
    for a, b in zip(dst.local_shards(), dst0.local_shards()):
        assert torch.equal(a.tensor, b.tensor)
    # ... and with that turned off every destination reads the owner's whole staging block: same bytes
    os.environ["TSB200_P2P_SCATTER"] = "0"
    try:
        dst1 = _sharded(1, world, shape)
        snap.restore({"emb": StateDict(t=dst1), "own": out_own})
        assert reshard.last_stats.pack_launches == 0 and reshard.last_stats.peer_bytes == st.peer_bytes
        for a, b in zip(dst.local_shards(), dst1.local_shards()):
            assert torch.equal(a.tensor, b.tensor)
    finally:
        del os.environ["TSB200_P2P_SCATTER"]
    # the storage-only path (reference semantics) gives the same bytes
    os.environ["TSB200_P2P_RESHARD"] = "0"
    try:
        dst2 = _sharded(1, world, shape)
        reshard.last_stats = None
        snap.restore({"emb": StateDict(t=dst2), "own": out_own})
        assert reshard.last_stats is None
        for a, b in zip(dst.local_shards(), dst2.local_shards()):
            assert torch.equal(a.tensor, b.tensor)
    finally:
        del os.environ["TSB200_P2P_RESHARD"]
    # row-wise -> row-wise at the same world size: the storage-only path reads every byte once as well, so
    # every rank leaves the table to it (no staging, no peer traffic) ...
    dst3 = _sharded(0, world, shape)
    snap.restore({"emb": StateDict(t=dst3), "own": out_own})
    st3 = reshard.last_stats
    assert st3.fell_back and st3.skipped_groups == ["sharded"] and st3.read_bytes == 0, vars(st3)
    for a, b in zip(src.local_shards(), dst3.local_shards()):
        assert torch.equal(a.tensor, b.tensor)
    # ... unless told to share reads regardless: every rank then owns exactly its own pieces
    os.environ["TSB200_P2P_MIN_AMPLIFICATION"] = "0"
    try:
        dst4 = _sharded(0, world, shape)
        snap.restore({"emb": StateDict(t=dst4), "own": out_own})
        assert reshard.last_stats.peer_bytes == 0 and reshard.last_stats.local_bytes == total // world
        for a, b in zip(src.local_shards(), dst4.local_shards()):
            assert torch.equal(a.tensor, b.tensor)
    finally:
        del os.environ["TSB200_P2P_MIN_AMPLIFICATION"]


def ddp_replicated(rank, world, tmp):
    """Replicated CUDA state partitioned over the ranks; every rank restores all of it."""
    import _cases
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    state = StateDict({f"p{i}": _cases.raw_tensor(200 + i, [257, 33 + i], torch.float32).cuda() for i in range(12)})
    state["step"] = torch.tensor(3.0, device="cuda")
    snap = Snapshot.take(os.path.join(tmp, "snap"), {"m": state}, replicated=["**"])
    out = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
    from torchsnapshot_b200 import reshard

    reshard.last_stats = None
    snap.restore({"m": out})
    for k in state:
        assert torch.equal(out[k], state[k]), k
    # replicated restore: every byte is read from storage by one rank only; the other gets it over NVLink
    st = reshard.last_stats
    total = sum(v.numel() * v.element_size() for v in state.values())
    t = torch.tensor([st.read_bytes, st.peer_bytes, st.local_bytes], dtype=torch.float64, device="cuda")
    dist.all_reduce(t)
    assert int(t[0].item()) == total, (t.tolist(), total)
    assert int(t[1].item()) + int(t[2].item()) == total * world and st.peer_bytes > 0
    # and the storage-only path agrees
    os.environ["TSB200_P2P_RESHARD"] = "0"
    try:
        out2 = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
        snap.restore({"m": out2})
        for k in state:
            assert torch.equal(out2[k], state[k]), k
    finally:
        del os.environ["TSB200_P2P_RESHARD"]
    m = snap.get_manifest()
    assert all(m[f"0/m/p{i}"].replicated for i in range(12)) and "1/m/p0" not in m
    # the next checkpoints of the same tensors repeat the plan (new values, same files from the same ranks)
    from torchsnapshot_b200.snapshot import _RepeatTake

    for round_ in range(2):
        for v in state.values():
            v.add_(1.0)
        before = _RepeatTake.repeats
        snap2 = Snapshot.take(os.path.join(tmp, f"snap_again{round_}"), {"m": state}, replicated=["**"])
        assert _RepeatTake.repeats == before + 1, (round_, _RepeatTake.repeats, before)
        out3 = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
        snap2.restore({"m": out3})
        for k in state:
            assert torch.equal(out3[k], state[k]), k
        m2 = snap2.get_manifest()
        assert {k: (e.location, e.byte_range) for k, e in m.items() if hasattr(e, "byte_range")} == {
            k: (e.location, e.byte_range) for k, e in m2.items() if hasattr(e, "byte_range")}


def _flat_cuda(world, rank, n=200003):
    from torch.distributed._shard import sharded_tensor
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec

    spec = ChunkShardingSpec(dim=0, placements=[f"rank:{r}/cuda:{r}" for r in range(world)])
    return sharded_tensor.empty(spec, n, dtype=torch.float32)


def elastic_take(rank, world, tmp):
    """FSDP-style flat parameter (BASELINE config 3 layout) + replicated state, saved by `world` ranks."""
    import _cases
    from torchsnapshot_b200 import knobs
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    full = _cases.raw_tensor(900, [200003], torch.float32)
    st = _flat_cuda(world, rank)
    for shard in st.local_shards():
        o, n = shard.metadata.shard_offsets[0], shard.metadata.shard_sizes[0]
        shard.tensor.copy_(full[o : o + n])
    rep = StateDict(w=_cases.raw_tensor(901, [333, 7], torch.float32).cuda(), step=5)
    with knobs.override_max_shard_size_bytes(100000):
        Snapshot.take(os.path.join(tmp, "snap"), {"flat": StateDict(p=st), "rep": rep}, replicated=["rep/**"])


def elastic_restore(rank, world, tmp):
    """...restored by another number of ranks (8->4 / 4->8 in miniature), on the GPUs."""
    import _cases
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    full = _cases.raw_tensor(900, [200003], torch.float32)
    st = _flat_cuda(world, rank)
    for shard in st.local_shards():
        shard.tensor.zero_()
    rep = StateDict(w=torch.zeros(333, 7, device="cuda"), step=0)
    Snapshot(os.path.join(tmp, "snap")).restore({"flat": StateDict(p=st), "rep": rep})
    for shard in st.local_shards():
        o, n = shard.metadata.shard_offsets[0], shard.metadata.shard_sizes[0]
        assert torch.equal(shard.tensor.cpu(), full[o : o + n]), (rank, o, n)
    assert torch.equal(rep["w"].cpu(), _cases.raw_tensor(901, [333, 7], torch.float32)) and rep["step"] == 5


def elastic_roundtrip_gpu(rank, world, tmp, sub=None):
    """N -> N/2 -> N inside one job (tests/_elastic.py), on the GPUs: once the way the product decides (row-wise
    re-cuts read nothing twice, so they go through storage on every rank; the replicated state is shared), and
    once with every read shared, so that the re-cut shards cross NVLink where saved pieces straddle the new
    boundaries.  Every restored shard is compared with the source pattern."""
    import _elastic

    if sub is None:
        sub = dist.new_group(ranks=list(range(max(world // 2, 1))))
    info = _elastic.elastic_roundtrip(os.path.join(tmp, "a"), "cuda", sub=sub)
    assert info["restore_at_n"] is not None and "sharded" in info["restore_at_n"]["skipped"], info
    os.environ["TSB200_P2P_MIN_AMPLIFICATION"] = "0"
    try:
        info = _elastic.elastic_roundtrip(os.path.join(tmp, "b"), "cuda", sub=sub)
        assert info["restore_at_n"]["p2p"], info
        t = torch.tensor([info["restore_at_n"]["peer_bytes"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        assert t.item() > 0, "some shard boundary falls inside a saved piece, so some bytes cross NVLink"
    finally:
        del os.environ["TSB200_P2P_MIN_AMPLIFICATION"]


def ranks_on_different_hosts(rank, world, tmp):
    """A job that spans hosts (here: the ranks CLAIM different hostnames) must not try CUDA IPC between them:
    restore takes the reference's storage-only path on every rank, for replicated and for sharded state."""
    import _cases
    import torchsnapshot_b200.snapshot as snapmod
    from torchsnapshot_b200 import reshard
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    shape = [512, 64]
    full = _cases.raw_tensor(78, shape, torch.float32)
    src = _sharded(0, world, shape, full)
    rep = StateDict({f"p{i}": _cases.raw_tensor(400 + i, [129, 17], torch.float32).cuda() for i in range(6)})
    snap = Snapshot.take(os.path.join(tmp, "snap"), {"emb": StateDict(t=src), "rep": rep}, replicated=["rep/**"])
    real = snapmod.socket.gethostname
    snapmod.socket.gethostname = lambda: f"node-{rank}"
    try:
        dst = _sharded(1, world, shape)
        out = StateDict({k: torch.zeros_like(v) for k, v in rep.items()})
        reshard.last_stats = None
        snap.restore({"emb": StateDict(t=dst), "rep": out})
        assert reshard.last_stats is None, "no peer-to-peer path between hosts"
    finally:
        snapmod.socket.gethostname = real
    for shard in dst.local_shards():
        want = _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes)).cuda()
        assert torch.equal(shard.tensor, want)
    for k in rep:
        assert torch.equal(out[k], rep[k]), k
