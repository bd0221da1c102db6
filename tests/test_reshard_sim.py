"""The peer-to-peer restore orchestration of reshard.py, end to end on the CPU (tests/_p2p_sim.py): ranks are
threads, device memory is host memory, the copy kernel is the memmove interpreter.  What is checked here is
everything around the CUDA calls -- who owns which unit, what each destination asks for, where the owner
packs it, which regions each rank launches, that replicated units are shared whole while reshards go through
per-destination buffers, and that every rank ends up with the bytes the saved tensor holds.  The same flows
run on real GPUs in tests/test_gpu_multi.py."""

import asyncio
import threading
from types import SimpleNamespace as NS

import pytest
import torch

import _cases
import _p2p_sim
from torchsnapshot_b200 import knobs, reshard
from torchsnapshot_b200.batcher import batch_read_requests
from torchsnapshot_b200.io_preparer import prepare_read
from torchsnapshot_b200.manifest_ops import get_manifest_for_rank
from torchsnapshot_b200.snapshot import Snapshot, StateDict
from torchsnapshot_b200.storage_plugins.fs import FSStoragePlugin

ROWS, COLS = 96, 40


class Sharded:
    """Duck-typed ShardedTensor over CPU tensors: local shards with global offsets."""

    def __init__(self, boxes, full=None):
        self._shards = []
        for off, sizes in boxes:
            t = _cases.box_slice(full, (off, sizes)).contiguous() if full is not None else torch.zeros(sizes)
            self._shards.append(NS(tensor=t, metadata=NS(shard_offsets=list(off), shard_sizes=list(sizes))))

    def local_shards(self):
        return self._shards

    def metadata(self):
        return NS(size=[ROWS, COLS])

    def sharding_spec(self):
        return None


@pytest.fixture
def sharded_duck(monkeypatch):
    import torchsnapshot_b200.io_preparer as iop
    import torchsnapshot_b200.snapshot as snapmod

    monkeypatch.setattr(iop, "_is_sharded_tensor", lambda o: isinstance(o, Sharded))
    monkeypatch.setattr(snapmod, "_is_sharded_tensor", lambda o: isinstance(o, Sharded))


def _take(tmp_path, full, replicated_extra):
    """A snapshot of a [ROWS, COLS] fp32 table saved row-wise by "4 ranks" (four local shards of one process,
    each subdivided into pieces of 8 rows) plus a few dense tensors that play the replicated state."""
    per = ROWS // 4
    src = Sharded([([r * per, 0], [per, COLS]) for r in range(4)], full)
    with knobs.override_max_shard_size_bytes(8 * COLS * 4):
        snap = Snapshot.take(str(tmp_path / "snap"), {"emb": StateDict(t=src), "rep": StateDict(**replicated_extra)})
    return snap


@pytest.fixture(autouse=True)
def _classic_mode(monkeypatch):
    """The tests written against the first version of the path: always take it (no read-amplification
    threshold) and give all staging back at the end of every call.  The tests of the pool, of the threshold
    and of the fall-backs override these."""
    monkeypatch.setenv("TSB200_P2P_MIN_AMPLIFICATION", "0")
    monkeypatch.setenv("TSB200_TRIM_ARENAS", "1")
    monkeypatch.setattr(reshard, "_POOL_GRANULE", 512)  # exact block sizes: the tests below count exported bytes


def _run_ranks(world, root, units_of, monkeypatch, storage_of=None, expect_errors=False, session=None):
    """Run reshard.p2p_restore_units on `world` threads; returns (stats per rank, call log, engines) -- or,
    with expect_errors, the (rank, exception) pairs.  `session` (a dict) keeps engines and the group between
    calls, the way a process keeps them between restores."""
    if session is not None and "engines" in session:
        engines, group, log = session["engines"], session["group"], session["log"]
    else:
        registry, log = {}, []
        engines = [_p2p_sim.FakeEngine(r, registry, log) for r in range(world)]
        group = _p2p_sim.ThreadGroup(world)
        if session is not None:
            session.update(engines=engines, group=group, log=log, views=[group.view(r) for r in range(world)])
    monkeypatch.setattr(reshard, "get_engine", lambda device: engines[device])
    stats, errors = [None] * world, []

    def worker(rank):
        loop = asyncio.new_event_loop()
        storage = storage_of(rank) if storage_of else FSStoragePlugin(root)
        try:
            view = session["views"][rank] if session is not None else group.view(rank)
            stats[rank] = reshard.p2p_restore_units(units_of(rank), storage, view, loop, rank)
        except BaseException as e:  # noqa: B902
            errors.append((rank, e))
            if not expect_errors:
                group._barrier.abort()
        finally:
            storage.sync_close(loop)
            loop.close()

    threads = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert all(not t.is_alive() for t in threads), "a rank is stuck"
    if session is None:
        assert all(not e.ctx.blocks for e in engines), "staging was not freed"
    if expect_errors:
        return errors
    assert not errors, errors
    return stats, log, engines


def _layouts(world):
    per, half, colw = ROWS // world, ROWS // world // 2, COLS // world
    shifted = {}
    for r in range(world):
        boxes = [([(r * per + half) % ROWS, 0], [per - half if r == world - 1 else per, COLS])]
        if r == world - 1:
            boxes.append(([0, 0], [half, COLS]))
        shifted[r] = boxes
    return {
        "row_to_column": {r: [([0, r * colw], [ROWS, colw])] for r in range(world)},
        "rows_shifted": shifted,
        "rows_same_cut": {r: [([r * per, 0], [per, COLS])] for r in range(world)},
    }


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("layout", ["row_to_column", "rows_shifted", "rows_same_cut"])
@pytest.mark.parametrize("scatter", ["1", "0"])
def test_reshard_orchestration_on_cpu(tmp_path, monkeypatch, sharded_duck, world, layout, scatter):
    monkeypatch.setenv("TSB200_P2P_SCATTER", scatter)
    full = _cases.raw_tensor(5, [ROWS, COLS], torch.float32)
    snap = _take(tmp_path, full, {"w": torch.arange(10.0)})
    _manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    entry = merged["emb/t"]
    assert len(entry.shards) == 12  # 4 saved shards x 3 pieces of 8 rows
    dsts = {r: Sharded(_layouts(world)[layout][r]) for r in range(world)}
    stats, log, _engines = _run_ranks(
        world, snap.path, lambda r: reshard.sharded_units([("emb/t", entry, dsts[r])]), monkeypatch
    )
    for r in range(world):
        for shard in dsts[r].local_shards():
            want = _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes))
            assert torch.equal(shard.tensor, want), (r, shard.metadata.shard_offsets)
    total = ROWS * COLS * 4
    assert sum(s.read_bytes for s in stats) == total  # every piece read from storage exactly once
    assert sum(s.local_bytes + s.peer_bytes for s in stats) == total
    assert all(s.launches == 1 for s in stats)
    peer = sum(s.peer_bytes for s in stats)
    if layout == "rows_same_cut":
        assert peer == 0 and not [x for x in log if x[0] == "open"]
    elif layout == "row_to_column":
        assert peer == total * (world - 1) // world
    exports = [x for x in log if x[0] == "export"]
    if scatter == "1" and peer:
        # reshard units never expose the owner's staging: what a peer maps is a buffer of its own, no larger
        # than what it receives (plus alignment), and each is a driver allocation of its own
        assert all(is_own for _k, _r, _n, is_own in exports)
        assert sum(n for _k, _r, n, _o in exports) <= peer + 512 * 12 * world
        assert sum(s.pack_launches for s in stats) > 0
    elif scatter == "0" and peer:
        assert sum(s.pack_launches for s in stats) == 0
        assert sum(n for _k, _r, n, _o in exports) >= total - total // world  # whole staging blocks


@pytest.mark.parametrize("world", [2, 3])
def test_replicated_units_are_shared_whole_and_mixed_with_a_reshard(tmp_path, monkeypatch, sharded_duck, world):
    """Replicated state (every rank needs every byte) is read by one rank per unit and copied by the others
    out of that rank's staging block; a reshard in the SAME call still goes through per-destination buffers."""
    full = _cases.raw_tensor(6, [ROWS, COLS], torch.float32)
    rep = {f"p{i}": _cases.raw_tensor(40 + i, [257 + i, 3], torch.float32) for i in range(6)}
    snap = _take(tmp_path, full, rep)
    manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    entry = merged["emb/t"]
    colw = COLS // world if COLS % world == 0 else None
    boxes = (lambda r: [([0, r * colw], [ROWS, colw])]) if colw else (lambda r: [([0, 0], [ROWS, COLS])] if r == 0 else [])
    dsts = {r: Sharded(boxes(r)) for r in range(world)}
    outs = {r: {k: torch.zeros_like(v) for k, v in rep.items()} for r in range(world)}

    def units_of(r):
        rrs = []
        for k in rep:
            reqs, _fut = prepare_read(manifest[f"rep/{k}"], obj_out=outs[r][k])
            rrs += reqs
        extra = reshard.read_req_units(batch_read_requests(rrs))
        return list(reshard.sharded_units([("emb/t", entry, dsts[r])])) + list(extra)

    stats, log, _engines = _run_ranks(world, snap.path, units_of, monkeypatch)
    for r in range(world):
        for k, v in rep.items():
            assert torch.equal(outs[r][k], v), (r, k)
        for shard in dsts[r].local_shards():
            want = _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes))
            assert torch.equal(shard.tensor, want)
    rep_bytes = sum(v.numel() * 4 for v in rep.values())
    assert sum(s.read_bytes for s in stats) == ROWS * COLS * 4 + rep_bytes  # nothing is read twice
    shared = [x for x in log if x[0] == "export" and x[2] >= rep_bytes // world // 2 and x[3]]
    assert shared, "the replicated staging block is exported as an allocation of its own"


def test_a_failed_read_on_one_rank_raises_on_every_rank(tmp_path, monkeypatch, sharded_duck):
    """Nobody may wait for bytes that will never arrive: the ranks agree on failure before the copy phase
    (reference semantics would be an exception on the failing rank only -- and a hang in a collective here)."""
    full = _cases.raw_tensor(8, [ROWS, COLS], torch.float32)
    snap = _take(tmp_path, full, {"w": torch.arange(10.0)})
    _manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    dsts = {r: Sharded(_layouts(2)["row_to_column"][r]) for r in range(2)}

    class Broken(FSStoragePlugin):
        async def read_into(self, path, byte_range, out):
            raise OSError("disk on fire")

    errors = _run_ranks(
        2, snap.path, lambda r: reshard.sharded_units([("emb/t", merged["emb/t"], dsts[r])]), monkeypatch,
        storage_of=lambda r: Broken(snap.path) if r == 1 else FSStoragePlugin(snap.path), expect_errors=True,
    )
    assert sorted(r for r, _e in errors) == [0, 1]
    for _r, e in errors:
        assert isinstance(e, RuntimeError) and "rank 1" in str(e) and "disk on fire" in str(e)


@pytest.mark.parametrize("layout", ["rows_shifted", "rows_same_cut"])
def test_saved_by_four_restored_by_three(tmp_path, monkeypatch, sharded_duck, layout):
    """A world size that does not divide the saved cut: 4 saved shards (12 pieces of 8 rows) restored into
    3 row shards of 32 rows -- pieces straddle the new boundaries, so most units have two destinations."""
    full = _cases.raw_tensor(9, [ROWS, COLS], torch.float32)
    snap = _take(tmp_path, full, {"w": torch.arange(10.0)})
    _manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    dsts = {r: Sharded(_layouts(3)[layout][r]) for r in range(3)}
    stats, _log, _engines = _run_ranks(
        3, snap.path, lambda r: reshard.sharded_units([("emb/t", merged["emb/t"], dsts[r])]), monkeypatch
    )
    for r in range(3):
        for shard in dsts[r].local_shards():
            want = _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes))
            assert torch.equal(shard.tensor, want), (r, shard.metadata.shard_offsets)
    assert sum(s.read_bytes for s in stats) == ROWS * COLS * 4
    if layout == "rows_same_cut":  # 32-row shards are unions of whole 8-row pieces: every piece has one destination
        assert sum(s.peer_bytes for s in stats) == 0


def _check(dsts, full, world):
    for r in range(world):
        for shard in dsts[r].local_shards():
            want = _cases.box_slice(full, (shard.metadata.shard_offsets, shard.metadata.shard_sizes))
            assert torch.equal(shard.tensor, want), (r, shard.metadata.shard_offsets)


@pytest.mark.parametrize("world", [2, 4])
def test_staging_and_peer_mappings_are_kept_between_restores(tmp_path, monkeypatch, sharded_duck, world):
    """Second restore of the same plan: no new driver allocation, no new mapping, nothing unmapped or freed;
    a changed plan (other columns) still lands in the pooled blocks; the pool is given back collectively."""
    monkeypatch.setenv("TSB200_TRIM_ARENAS", "0")
    full = _cases.raw_tensor(11, [ROWS, COLS], torch.float32)
    rep = {f"p{i}": _cases.raw_tensor(60 + i, [300 + i, 5], torch.float32) for i in range(4)}
    snap = _take(tmp_path, full, rep)
    manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    entry = merged["emb/t"]
    session = {}
    colw = COLS // world

    def make(shift):
        dsts = {r: Sharded([([0, ((r + shift) % world) * colw], [ROWS, colw])]) for r in range(world)}
        outs = {r: {k: torch.zeros_like(v) for k, v in rep.items()} for r in range(world)}

        def units_of(r):
            rrs = []
            for k in rep:
                reqs, _fut = prepare_read(manifest[f"rep/{k}"], obj_out=outs[r][k])
                rrs += reqs
            return list(reshard.sharded_units([("emb/t", entry, dsts[r])])) + list(
                reshard.read_req_units(batch_read_requests(rrs)))

        return dsts, outs, units_of

    dsts, outs, units_of = make(0)
    first, log, engines = _run_ranks(world, snap.path, units_of, monkeypatch, session=session)
    _check(dsts, full, world)
    assert sum(s.new_staging_bytes for s in first) > 0 and sum(s.opens for s in first) > 0
    assert all(e.ctx.blocks for e in engines), "the pool keeps the exportable blocks"
    n_log = len(log)
    dsts, outs, units_of = make(0)
    second, log, engines = _run_ranks(world, snap.path, units_of, monkeypatch, session=session)
    _check(dsts, full, world)
    for r in range(world):
        for k, v in rep.items():
            assert torch.equal(outs[r][k], v)
    assert all(s.new_staging_bytes == 0 and s.opens == 0 for s in second)
    assert not [x for x in log[n_log:] if x[0] in ("open", "close", "alloc_exportable")], log[n_log:]
    # every rank takes the columns of its neighbour: other requests, same sizes -> same blocks, same mappings
    dsts, outs, units_of = make(1)
    third, log, engines = _run_ranks(world, snap.path, units_of, monkeypatch, session=session)
    _check(dsts, full, world)
    assert all(s.new_staging_bytes == 0 for s in third)
    # giving the pool back is collective: unmap first, then free
    threads = [threading.Thread(target=reshard._give_back,
                                args=(reshard._peer_state(engines[r], session["views"][r]), engines[r].ctx, session["views"][r]))
               for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=60)
    assert all(not e.ctx.blocks for e in engines)


def test_a_pool_over_its_bound_is_flushed_by_the_next_call(tmp_path, monkeypatch, sharded_duck):
    monkeypatch.setenv("TSB200_TRIM_ARENAS", "0")
    monkeypatch.setenv("TSB200_P2P_POOL_BYTES", "1")  # any pooled block is over the bound
    full = _cases.raw_tensor(12, [ROWS, COLS], torch.float32)
    snap = _take(tmp_path, full, {"w": torch.arange(10.0)})
    _manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    session = {}
    for _ in range(2):
        dsts = {r: Sharded(_layouts(2)["row_to_column"][r]) for r in range(2)}
        stats, log, engines = _run_ranks(
            2, snap.path, lambda r: reshard.sharded_units([("emb/t", merged["emb/t"], dsts[r])]), monkeypatch,
            session=session)
        _check(dsts, full, 2)
    # the second call began by unmapping and freeing what the first left, then allocated afresh
    assert sum(s.new_staging_bytes for s in stats) > 0
    assert [x for x in log if x[0] == "close"]


@pytest.mark.parametrize("layout,expect_p2p", [("rows_same_cut", False), ("rows_shifted", False), ("row_to_column", True)])
def test_the_path_is_taken_only_when_sharing_reads_saves_reads(tmp_path, monkeypatch, sharded_duck, layout, expect_p2p):
    """Row-wise re-cuts read every byte once on the storage-only path as well (it reads only the rows it
    needs): every rank falls back, nothing is staged.  Row -> column makes every rank read every piece: shared."""
    monkeypatch.setenv("TSB200_P2P_MIN_AMPLIFICATION", "1.25")
    full = _cases.raw_tensor(13, [ROWS, COLS], torch.float32)
    snap = _take(tmp_path, full, {"w": torch.arange(10.0)})
    _manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    dsts = {r: Sharded(_layouts(4)[layout][r]) for r in range(4)}
    stats, log, _engines = _run_ranks(
        4, snap.path, lambda r: reshard.sharded_units([("emb/t", merged["emb/t"], dsts[r])]), monkeypatch)
    assert all(s.fell_back == (not expect_p2p) for s in stats), [(s.fell_back, s.amplification) for s in stats]
    if expect_p2p:
        _check(dsts, full, 4)
        assert all(abs(s.amplification["sharded"] - 4.0) < 1e-6 and not s.skipped_groups for s in stats)
    else:
        assert all(s.amplification["sharded"] <= 1.01 and s.read_bytes == 0 for s in stats)
        assert all(s.skipped_groups == ["sharded"] for s in stats)
        assert not [x for x in log if x[0] in ("alloc_exportable", "export", "open")]


def test_replicated_state_is_shared_while_a_row_recut_in_the_same_call_is_not(tmp_path, monkeypatch, sharded_duck):
    """Two groups in one call: the replicated units (every rank would read all of them: 2x) take the path, the
    sharded table re-cut along rows (1x either way) is left to the caller, on both ranks alike."""
    monkeypatch.setenv("TSB200_P2P_MIN_AMPLIFICATION", "1.25")
    full = _cases.raw_tensor(15, [ROWS, COLS], torch.float32)
    rep = {f"p{i}": _cases.raw_tensor(70 + i, [211 + i, 3], torch.float32) for i in range(5)}
    snap = _take(tmp_path, full, rep)
    manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    dsts = {r: Sharded(_layouts(2)["rows_shifted"][r]) for r in range(2)}
    outs = {r: {k: torch.zeros_like(v) for k, v in rep.items()} for r in range(2)}

    def units_of(r):
        rrs = []
        for k in rep:
            reqs, _fut = prepare_read(manifest[f"rep/{k}"], obj_out=outs[r][k])
            rrs += reqs
        return list(reshard.sharded_units([("emb/t", merged["emb/t"], dsts[r])])) + list(
            reshard.read_req_units(batch_read_requests(rrs)))

    stats, _log, _engines = _run_ranks(2, snap.path, units_of, monkeypatch)
    rep_bytes = sum(v.numel() * 4 for v in rep.values())
    assert all(s.skipped_groups == ["sharded"] and not s.fell_back for s in stats)
    assert sum(s.read_bytes for s in stats) == rep_bytes
    for r in range(2):
        for k, v in rep.items():
            assert torch.equal(outs[r][k], v)
        for shard in dsts[r].local_shards():
            assert not shard.tensor.any(), "the sharded table is the caller's to load"


@pytest.mark.parametrize("fault", ["no_staging", "no_mapping"])
def test_every_rank_falls_back_when_one_cannot_stage_or_map(tmp_path, monkeypatch, sharded_duck, fault):
    """A rank without HBM for its staging, or a peer allocation that cannot be mapped: every rank returns
    `fell_back` (the caller then reads from storage), none hangs in a collective, nothing stays allocated."""
    full = _cases.raw_tensor(14, [ROWS, COLS], torch.float32)
    snap = _take(tmp_path, full, {"w": torch.arange(10.0)})
    _manifest, merged = get_manifest_for_rank(snap.metadata, 0)
    dsts = {r: Sharded(_layouts(2)["row_to_column"][r]) for r in range(2)}
    session = {}
    # build the engines first so that the fault can be planted
    registry, log = {}, []
    engines = [_p2p_sim.FakeEngine(r, registry, log) for r in range(2)]
    group = _p2p_sim.ThreadGroup(2)
    session.update(engines=engines, group=group, log=log, views=[group.view(r) for r in range(2)])
    if fault == "no_staging":
        engines[1].ctx.fail_exportable = True
    else:
        engines[0].ctx.fail_open = True
    stats, log, engines = _run_ranks(
        2, snap.path, lambda r: reshard.sharded_units([("emb/t", merged["emb/t"], dsts[r])]), monkeypatch,
        session=session)
    assert all(s.fell_back for s in stats)
    assert all("rank 1" in s.reason if fault == "no_staging" else "rank 0" in s.reason for s in stats)
    assert all(not e.ctx.blocks for e in engines), "a fall-back leaves no staging behind"
