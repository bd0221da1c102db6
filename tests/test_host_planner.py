"""Host planner of the product (io_preparer / batcher / partitioner) against the oracle and the
reference-made fixtures.  CPU only: plans are integer math, and CPU tensors exercise the same
planner code as CUDA tensors."""

import asyncio
import hashlib
import json
import os

import pytest
import torch

import _cases
import oracle
from oracle import planner
from torchsnapshot_b200 import knobs
from torchsnapshot_b200.batcher import (
    BatchedBufferConsumer,
    BatchedBufferStager,
    batch_read_requests,
    batch_write_requests,
)
from torchsnapshot_b200.io_preparer import (
    ChunkedTensorIOPreparer,
    ShardedTensorIOPreparer,
    TensorIOPreparer,
    overlap_wrt_saved,
    prepare_read,
    prepare_write,
)
from torchsnapshot_b200.io_types import ReadReq
from torchsnapshot_b200.manifest import Shard, ShardedTensorEntry, TensorEntry
from torchsnapshot_b200.partitioner import assign_write_loads


def _ranged(buf, byte_range):
    return buf if byte_range is None else buf[byte_range[0] : byte_range[1]]


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def run(coro):
    loop = asyncio.new_event_loop()
    try:
        return loop.run_until_complete(coro)
    finally:
        loop.close()


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, name)) as f:
        return json.load(f)


@pytest.mark.parametrize("case", ["mixed", "threshold"])
def test_batch_write_requests_matches_reference(golden_dir, case):
    gold = _load(golden_dir, f"slab_{case}.json")
    threshold, items = getattr(_cases, f"slab_case_{case}")()
    entries, wrs = {}, []
    for path, spec in items:
        e, w = prepare_write(_cases.build_view(spec), path, rank=0, replicated=False)
        entries[path] = e
        wrs += w
    n_before = len(wrs)
    _, batched = batch_write_requests(list(entries.values()), wrs, slab_size_threshold_bytes=threshold)
    assert len(batched) < n_before
    staged = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in batched}
    slab_paths = [wr.path for wr in batched if wr.path.startswith("batched/")]
    assert [len(staged[p]) for p in slab_paths] == [s["size"] for s in gold["slabs"]]
    assert [sha(staged[p]) for p in slab_paths] == [s["sha256"] for s in gold["slabs"]]
    for path, e in entries.items():
        g = gold["objects"][path]
        assert e.byte_range == g["byte_range"] and e.dtype == g["dtype"] and e.shape == g["shape"]
        assert e.serializer == g["serializer"]
        if g["slab"] is None:
            assert e.location == g["location"]
        else:
            assert e.location == slab_paths[g["slab"]]
        buf = staged[e.location]
        payload = buf if e.byte_range is None else buf[e.byte_range[0] : e.byte_range[1]]
        assert sha(payload) == g["sha256"]


def test_batching_errors_like_the_reference():
    t = torch.zeros(4)
    e, w = prepare_write(t, "a", 0, False)
    with pytest.raises(RuntimeError, match="is not passed to batch_write"):
        batch_write_requests([], w)
    with pytest.raises(AssertionError):
        BatchedBufferStager({(0, 4): w[0].buffer_stager, (8, 12): w[0].buffer_stager})


def test_chunking_matches_reference(golden_dir):
    gold = {c["name"]: c for c in _load(golden_dir, "chunked.json")}
    for case in _cases.chunk_cases():
        t = _cases.build_view(case["spec"])
        chunks = ChunkedTensorIOPreparer.chunk_tensor(t, chunk_sz_bytes=case["chunk_sz_bytes"])
        want = planner.chunk_tensor(list(t.shape), t.element_size(), case["chunk_sz_bytes"])
        assert [(c.offsets, c.sizes) for c in chunks] == want
        entry, wrs = ChunkedTensorIOPreparer.prepare_write("0/foo", t, chunks)
        g = gold[case["name"]]
        assert entry.dtype == g["dtype"] and entry.shape == g["shape"]
        for c, wr, gc in zip(entry.chunks, wrs, g["chunks"]):
            assert (c.offsets, c.sizes, c.tensor.location, c.tensor.shape) == (
                gc["offsets"], gc["sizes"], gc["location"], gc["shape"])
            assert c.tensor.replicated is False  # inner entries stay non-replicated (reference quirk 5)
            assert sha(run(wr.buffer_stager.stage_buffer())) == gc["sha256"]
        # and the read side fills an output tensor chunk by chunk
        out = torch.zeros_like(t.contiguous())
        rrs, _ = ChunkedTensorIOPreparer.prepare_read(entry, out)
        staged = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in wrs}
        for rr in rrs:
            run(rr.buffer_consumer.consume_buffer(_ranged(staged[rr.path], rr.byte_range)))
        assert oracle.serialize(out) == oracle.serialize(t)


def test_prepare_write_dispatch_chunks_only_above_the_limit():
    t = torch.zeros(16, 16)  # 1024 B
    with knobs.override_max_chunk_size_bytes(1024):
        e, w = prepare_write(t, "x", 3, False)
        assert e.type == "Tensor" and e.location == "3/x" and len(w) == 1
    with knobs.override_max_chunk_size_bytes(1023):
        e, w = prepare_write(t, "x", 3, True)
        assert e.type == "ChunkedTensor" and len(w) == 2 and e.replicated
        assert [c.tensor.location for c in e.chunks] == ["replicated/x_0_0", "replicated/x_8_0"]


def _sharded_entry(layout_name, full):
    dim, boxes = _cases.shard_layouts(list(full.shape))[layout_name]
    shards, wrs = [], []
    for off, sizes in boxes:
        local = _cases.box_slice(full, (off, sizes)).contiguous()
        for tensor, o, s in ShardedTensorIOPreparer.subdivide_shard(local, list(off), list(sizes), dim,
                                                                    _cases.SHARD_MAX_BYTES):
            entry, w = TensorIOPreparer.prepare_write(f"sharded/foo_{'_'.join(map(str, o))}", tensor)
            shards.append(Shard(offsets=o, sizes=s, tensor=entry))
            wrs += w
    return ShardedTensorEntry(shards=shards), wrs


class _FakeSharded:
    """Duck-typed stand-in for a ShardedTensor's local view (no process group needed)."""

    def __init__(self, boxes, dtype, global_shape):
        from types import SimpleNamespace as NS

        self._shape = list(global_shape)
        self._shards = [NS(tensor=torch.zeros(s, dtype=dtype), metadata=NS(shard_offsets=list(o), shard_sizes=list(s)))
                        for o, s in boxes]

    def local_shards(self):
        return self._shards


def test_sharded_write_and_reshard_match_reference(golden_dir, monkeypatch):
    import torchsnapshot_b200.io_preparer as iop

    gold = _load(golden_dir, "sharded.json")
    shape = _cases.SHARD_GLOBAL_SHAPE
    full = _cases.raw_tensor(_cases.SHARD_SEED, shape, _cases.DTYPES[_cases.SHARD_DTYPE])
    layouts = _cases.shard_layouts(shape)
    # let the preparer treat the duck-typed object as a sharded tensor
    monkeypatch.setattr(iop, "_is_sharded_tensor", lambda o: isinstance(o, _FakeSharded))
    monkeypatch.setattr(_FakeSharded, "metadata", lambda self: type("M", (), {"size": self._shape})(), raising=False)
    entries = {}
    stageds = {}
    for name in layouts:
        entry, wrs = _sharded_entry(name, full)
        g = gold["writes"][name]["pieces"]
        assert [(s.offsets, s.sizes, s.tensor.location) for s in entry.shards] == [
            (p["offsets"], p["sizes"], p["location"]) for p in g]
        stageds[name] = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in wrs}
        assert [sha(stageds[name][s.tensor.location]) for s in entry.shards] == [p["sha256"] for p in g]
        entries[name] = entry
    for key, g in gold["reshard"].items():
        src_name, dst_name = key.split("->")
        dst = _FakeSharded(layouts[dst_name][1], full.dtype, shape)
        rrs, _ = ShardedTensorIOPreparer.prepare_read(entries[src_name], dst)
        ptrs = [s.tensor.data_ptr() for s in dst.local_shards()]
        plan = [{"location": rr.path,
                 "regions": [{"dst_shard": ptrs.index(r.dst_tensor.data_ptr()), "overlap": [list(x) for x in r.overlap_region]}
                             for r in rr.buffer_consumer.overlapping_regions]} for rr in rrs]
        assert plan == g["plan"], key
        for rr in rrs:
            run(rr.buffer_consumer.consume_buffer(_ranged(stageds[src_name][rr.path], rr.byte_range)))
        assert [sha(oracle.serialize(s.tensor)) for s in dst.local_shards()] == g["dst_sha256"], key
    # sharded -> dense tensor (read_object of a sharded entry)
    dense = torch.zeros(shape, dtype=full.dtype)
    rrs, _ = ShardedTensorIOPreparer.prepare_read(entries["dim1x3"], dense)
    for rr in rrs:
        run(rr.buffer_consumer.consume_buffer(_ranged(stageds["dim1x3"][rr.path], rr.byte_range)))
    assert sha(oracle.serialize(dense)) == gold["dense_sha256"]


def test_overlap_matches_oracle():
    import itertools

    boxes = [([0, 0], [5, 7]), ([3, 2], [4, 9]), ([5, 0], [2, 2]), ([0, 6], [9, 1]), ([4, 6], [1, 1])]
    for (ao, asz), (bo, bsz) in itertools.product(boxes, boxes):
        got = overlap_wrt_saved(ao, asz, bo, bsz)
        if planner.boxes_overlap(bo, bsz, ao, asz):
            assert got == planner.overlap_region(ao, asz, bo, bsz)
        else:
            assert got is None


def test_subdivision_golden_counts():
    t = torch.zeros(256, 256)
    for max_bytes, dim, offsets, want in [(77, 0, [512, 0], 256), (1999, 0, [512, 0], 256), (4001, 0, [512, 0], 86),
                                          (4001, 1, [0, 512], 86), (300000, 0, [512, 0], 1)]:
        pieces = ShardedTensorIOPreparer.subdivide_shard(t, offsets, [256, 256], dim, max_bytes)
        assert len(pieces) == want
        assert [(o, s) for _t, o, s in pieces] == planner.subdivide_shard(offsets, [256, 256], 4, dim, max_bytes)
    with pytest.raises(ValueError):
        ShardedTensorIOPreparer.subdivide_shard(t, [0, 0], [256, 256], 0, 0)


def test_tiled_reads_match_reference(golden_dir):
    for g in _load(golden_dir, "tiled.json"):
        t = torch.zeros(g["shape"], dtype=_cases.DTYPES[g["dtype"]])
        entry = TensorEntry(location="x", serializer="buffer_protocol", dtype=str(t.dtype), shape=g["shape"],
                            replicated=False, byte_range=g["base"])
        rrs, _ = TensorIOPreparer.prepare_read(entry, t, buffer_size_limit_bytes=g["limit"])
        assert [list(rr.byte_range) for rr in rrs] == g["ranges"]
        assert [list(rr.buffer_consumer.tensor.shape) for rr in rrs] == g["tile_shapes"]


def test_batch_read_requests_merges_ranges_per_file():
    outs = [torch.zeros(4), torch.zeros(2), torch.zeros(3)]
    mk = lambda t, loc, br: ReadReq(path=loc, byte_range=br, buffer_consumer=prepare_read(
        TensorEntry(loc, "buffer_protocol", "torch.float32", list(t.shape), False, list(br) if br else None), t)[0][0].buffer_consumer)
    rrs = [mk(outs[0], "f", (16, 32)), mk(outs[1], "g", None), mk(outs[2], "f", (40, 52))]
    merged = batch_read_requests(rrs)
    assert len(merged) == 2
    whole = [r for r in merged if r.byte_range is None][0]
    assert whole.path == "g"
    ranged = [r for r in merged if r.byte_range is not None][0]
    assert ranged.byte_range == (16, 52) and isinstance(ranged.buffer_consumer, BatchedBufferConsumer)
    assert sorted(ranged.buffer_consumer.byte_range_to_buffer_consumer) == [(0, 16), (24, 36)]
    assert ranged.buffer_consumer.buf_sz_bytes == 36
    # agreement with the oracle's restatement
    _, want = planner.merge_reads([(r.path, r.byte_range) for r in rrs])
    assert want["f"][0] == (16, 52)
    payload = torch.arange(9, dtype=torch.float32).numpy().tobytes()
    run(ranged.buffer_consumer.consume_buffer(payload))
    assert outs[0].tolist() == [0, 1, 2, 3] and outs[2].tolist() == [6, 7, 8]


def test_partition_matches_reference(golden_dir):
    for g in _load(golden_dir, "partition.json"):
        loads = [(p, 0, s, False) for p, s in zip(g["paths"], g["sizes"])]
        got = assign_write_loads(loads, g["rank_to_size"])
        assert [[p for p, _i in r] for r in got] == g["assignment"]
        assert got and [[p for p, _ in r] for r in got] == planner.partition_write_loads(
            list(zip(g["paths"], g["sizes"])), g["rank_to_size"])


def test_chunks_are_spread_in_offset_order():
    loads = [("big", i, 100, True) for i in range(6)] + [("small", 0, 10, False)]
    got = assign_write_loads(loads, [0, 0, 0])
    assert got[0][0] == ("small", 0)  # whole paths first
    sizes = [sum(100 if p == "big" else 10 for p, _ in r) for r in got]
    assert max(sizes) - min(sizes) <= 100
    assert sorted(i for r in got for p, i in r if p == "big") == list(range(6))


def test_speed_weighted_partition_balances_finish_time():
    loads = [(f"p{i}", 0, 100, False) for i in range(40)]
    even = assign_write_loads(loads, [0, 0, 0, 0])
    assert [len(r) for r in even] == [10, 10, 10, 10]
    weighted = assign_write_loads(loads, [0, 0, 0, 0], [26.5, 38.0, 38.0, 38.0])
    counts = [len(r) for r in weighted]
    assert sum(counts) == 40 and counts[0] < counts[1] and max(counts) - min(counts) <= 4
    finish = [c * 100 / s for c, s in zip(counts, [26.5, 38.0, 38.0, 38.0])]
    assert max(finish) < 10 * 100 / 26.5  # better than the byte-balanced plan on these links
    # unusable speed vectors fall back to byte balance
    assert assign_write_loads(loads, [0, 0, 0, 0], [1.0, None, 1.0, 1.0]) == even


def test_merged_reads_split_at_large_holes():
    mk = lambda lo, hi: ReadReq(path="slab", byte_range=(lo, hi), buffer_consumer=prepare_read(
        TensorEntry("slab", "buffer_protocol", "torch.uint8", [hi - lo], False, [lo, hi]), torch.zeros(hi - lo, dtype=torch.uint8))[0][0].buffer_consumer)
    rrs = [mk(0, 100), mk(100, 300), mk(64 << 20, (64 << 20) + 50), mk(300, 1000)]
    merged = sorted(batch_read_requests(rrs), key=lambda r: r.byte_range)
    assert [r.byte_range for r in merged] == [(0, 1000), (64 << 20, (64 << 20) + 50)]
    assert sorted(merged[0].buffer_consumer.byte_range_to_buffer_consumer) == [(0, 100), (100, 300), (300, 1000)]


def test_sharded_read_fetches_only_the_needed_rows():
    """A destination that needs a few rows of a big persisted piece reads just those rows,
    and loads the same values as a whole-piece read."""
    full = _cases.raw_tensor(9, [1000, 16], torch.float32)
    entry, wrs = TensorIOPreparer.prepare_write("sharded/t_0_0", full)
    sharded = ShardedTensorEntry(shards=[Shard(offsets=[0, 0], sizes=[1000, 16], tensor=entry)])
    staged = bytes(run(wrs[0].buffer_stager.stage_buffer()))
    # a plain tensor that covers rows 0..40 of the global tensor ("only the overlapping part is loaded")
    out = torch.zeros(40, 16)
    rrs, _ = ShardedTensorIOPreparer.prepare_read(sharded, out)
    assert len(rrs) == 1 and rrs[0].byte_range == (0, 40 * 64)
    run(rrs[0].buffer_consumer.consume_buffer(staged[rrs[0].byte_range[0] : rrs[0].byte_range[1]]))
    assert torch.equal(out, full[:40])
    # a local shard in the middle: rows 300..420 of the piece, columns 4..12
    class Mid:
        def __init__(self):
            from types import SimpleNamespace as NS
            self._s = [NS(tensor=torch.zeros(120, 8), metadata=NS(shard_offsets=[300, 4], shard_sizes=[120, 8]))]
        def local_shards(self): return self._s
        def metadata(self):
            from types import SimpleNamespace as NS
            return NS(size=[1000, 16])
    import torchsnapshot_b200.io_preparer as iop
    mid = Mid()
    orig = iop._is_sharded_tensor
    iop._is_sharded_tensor = lambda o: isinstance(o, Mid)
    try:
        rrs, _ = ShardedTensorIOPreparer.prepare_read(sharded, mid)
    finally:
        iop._is_sharded_tensor = orig
    assert rrs[0].byte_range == (300 * 64, 420 * 64)
    assert rrs[0].buffer_consumer.entry.shape == [120, 16]
    run(rrs[0].buffer_consumer.consume_buffer(staged[rrs[0].byte_range[0] : rrs[0].byte_range[1]]))
    assert torch.equal(mid._s[0].tensor, full[300:420, 4:12])
    # a destination that needs (nearly) everything keeps the single whole-object read
    rrs, _ = ShardedTensorIOPreparer.prepare_read(sharded, torch.zeros(990, 16))
    assert rrs[0].byte_range is None


def test_merged_reads_stay_within_one_storage_object_size(monkeypatch):
    """Tiles of a budgeted read (read_object with memory_budget_bytes) must not be merged back into one read
    of the whole tensor: a merged read is one pinned buffer."""
    import torchsnapshot_b200.batcher as b

    monkeypatch.setattr(b, "_MAX_MERGED_READ_BYTES", 1000)
    t = torch.zeros(4096, dtype=torch.uint8)
    entry = TensorEntry("big", "buffer_protocol", "torch.uint8", [4096], False, None)
    rrs, _ = TensorIOPreparer.prepare_read(entry, t, buffer_size_limit_bytes=256)
    assert len(rrs) == 16
    merged = b.batch_read_requests(rrs)
    assert all(r.byte_range[1] - r.byte_range[0] <= 1000 for r in merged) and len(merged) >= 5
    covered = sorted(r.byte_range for r in merged)
    assert covered[0][0] == 0 and covered[-1][1] == 4096
    assert all(a[1] == c[0] for a, c in zip(covered, covered[1:]))
    payload = bytes(range(256)) * 16
    for r in merged:
        run(r.buffer_consumer.consume_buffer(payload[r.byte_range[0] : r.byte_range[1]]))
    assert bytes(t.numpy()) == payload


def _plan_signature(entries):
    """Entries with slab names replaced by the order in which the slabs first appear."""
    names, out = {}, []
    for path, e in entries.items():
        names.setdefault(e.location, len(names))
        br = tuple(e.byte_range) if e.byte_range is not None else None
        out.append((path, names[e.location] if e.location.startswith("batched/") else e.location, br, e.dtype,
                    tuple(e.shape), e.replicated, e.serializer, e.type))
    return out


@pytest.mark.parametrize("threshold", [4096, 40000, 1 << 20])
def test_bulk_planner_gives_the_per_leaf_plan(threshold):
    """batcher.plan_cuda_tensors (one vectorised pass) against prepare_write per tensor + batch_write_requests
    (the reference-shaped planner, itself held to the reference's fixtures above): same entries, same slab
    membership, same byte ranges, same stand-alone requests -- and the slabs' region tables, built by numpy,
    equal the member-by-member ones, also for strided members, 0-d tensors and empty tensors."""
    import numpy as np

    from torchsnapshot_b200.batcher import GPUSlabStager, plan_cuda_tensors
    from torchsnapshot_b200.regions import pack_regions, regions_to_array

    rs = np.random.RandomState(3)
    tensors = {}
    for i in range(400):
        kind = rs.randint(0, 10)
        n = int(rs.randint(1, 6000))
        if kind == 0:
            t = torch.zeros(n, 6, dtype=torch.float32)[:, ::2]  # strided view
        elif kind == 1:
            t = torch.tensor(float(i))  # 0-d
        elif kind == 2:
            t = torch.zeros(0, 3)  # empty
        elif kind == 3:
            t = torch.zeros(n, dtype=torch.bfloat16)
        elif kind == 4:
            t = torch.zeros(max(n // 7, 1), 7, dtype=torch.int64)
        else:
            t = torch.zeros(n, dtype=torch.float32)
        tensors[f"m/t{i}"] = t
    paths = list(tensors)
    replicated = [i % 5 == 0 for i in range(len(paths))]
    storage = [("replicated/" if r else "3/") + p for p, r in zip(paths, replicated)]
    bulk_entries, bulk_wrs = plan_cuda_tensors(paths, [tensors[p] for p in paths], storage, replicated, threshold)
    leaf_entries, leaf_wrs = {}, []
    for p, r in zip(paths, replicated):
        e, w = prepare_write(tensors[p], p, 3, r)
        leaf_entries[p] = e
        leaf_wrs += w
    _, leaf_wrs = batch_write_requests(list(leaf_entries.values()), leaf_wrs, threshold)
    assert _plan_signature(bulk_entries) == _plan_signature(leaf_entries)
    assert list(bulk_entries) == paths  # manifest order = the order given
    # stand-alone requests: same paths, in order; slabs: same sizes, in order
    solo = lambda wrs: [w.path for w in wrs if not w.path.startswith("batched/")]  # noqa: E731
    assert solo(bulk_wrs) == solo(leaf_wrs)
    slab_sizes = lambda wrs: [w.buffer_stager.slab_sz_bytes for w in wrs if w.path.startswith("batched/")]  # noqa: E731
    assert slab_sizes(bulk_wrs) == slab_sizes(leaf_wrs)
    # region tables: numpy-built == member by member
    by_slab = {}
    for p in paths:
        e = bulk_entries[p]
        if e.location.startswith("batched/"):
            by_slab.setdefault(e.location, []).append((tensors[p], e.byte_range[0]))
    base = 1 << 40
    for w in bulk_wrs:
        st = w.buffer_stager
        if not isinstance(st, GPUSlabStager):
            continue
        want = []
        for t, lo in by_slab[w.path]:
            want += pack_regions(t, base + lo)
        got = st.pack_region_array(base)
        key = lambda a: sorted(map(bytes, a))  # noqa: E731  (the strided members come last in the numpy-built table)
        assert key(got) == key(regions_to_array(want))
        assert st.pack_regions(base) and len(st.pack_regions(base)) == len(want)


def test_bulk_read_planner_gives_the_per_entry_reads():
    """batcher.plan_cuda_reads against prepare_read per entry + batch_read_requests: the same merged ranges per
    file (split at holes above 1 MiB), and region tables equal to the member-by-member ones."""
    import numpy as np

    from torchsnapshot_b200.batcher import plan_cuda_reads
    from torchsnapshot_b200.regions import regions_to_array

    rs = np.random.RandomState(4)
    items = []
    for f in range(3):
        cursor = 0
        for i in range(60):
            n = int(rs.randint(1, 5000))
            strided = i % 9 == 0
            t = torch.zeros(n, 4)[:, ::2] if strided else torch.zeros(n, 2)
            nbytes = n * 2 * 4
            if i % 20 == 19:
                cursor += 3 << 20  # a hole: members the restoring state dict does not ask for
            e = TensorEntry(location=f"batched/f{f}", serializer="buffer_protocol", dtype="torch.float32",
                            shape=[n, 2], replicated=False, byte_range=[cursor, cursor + nbytes])
            items.append((e, t))
            cursor += nbytes
    order = rs.permutation(len(items)).tolist()
    items = [items[i] for i in order]
    bulk = plan_cuda_reads(items)
    leaf = []
    for e, t in items:
        rrs, _fut = prepare_read(e, t)
        leaf += rrs
    leaf = batch_read_requests(leaf)
    key = lambda rr: (rr.path, rr.byte_range)  # noqa: E731
    assert sorted(map(key, bulk)) == sorted(map(key, leaf))
    assert len(bulk) == 12  # 3 files x 4 runs (three holes each)
    by_key = {key(rr): rr for rr in leaf}
    base = 1 << 41
    for rr in bulk:
        want = regions_to_array(by_key[key(rr)].buffer_consumer.unpack_regions(base))
        got = rr.buffer_consumer.unpack_region_array(base)
        assert sorted(map(bytes, got)) == sorted(map(bytes, want))
        assert rr.buffer_consumer.consumed_nbytes() == rr.byte_range[1] - rr.byte_range[0]
    # the size bound of a merged read cuts the same way (a budgeted read_object)
    for bound in (20000, 100000):
        leaf = []
        for e, t in items:
            rrs, _fut = prepare_read(e, t)
            leaf += rrs
        assert sorted(map(key, plan_cuda_reads(items, bound))) == sorted(map(key, batch_read_requests(leaf, bound)))


def test_take_plan_cache_hits_only_on_identical_tensors():
    """The plan of the previous take is reused when, and only when, every tensor is where it was, as it was."""
    from torchsnapshot_b200.batcher import TakePlanCache, plan_cuda_tensors

    def plan(tensors):
        paths = list(tensors)
        ts = [tensors[p] for p in paths]
        return paths, ts, plan_cuda_tensors(paths, ts, ["0/" + p for p in paths], [False] * len(paths), 4096)

    base = {f"t{i}": torch.zeros(10 + i, 3) for i in range(40)}
    base["view"] = torch.zeros(50, 8)[:, ::2]
    base["big"] = torch.zeros(5000)
    cache = TakePlanCache()
    paths, ts, (entries, wrs) = plan(base)
    cfg = ("cfg",)
    assert cache.get(cfg, paths, ts) is None
    cache.put(cfg, paths, ts, entries, wrs)
    # the same tensors through fresh Python objects (what module.state_dict() returns: detach() aliases)
    again = [t.detach() for t in ts]
    hit = cache.get(cfg, paths, again)
    assert hit is not None and hit[0] is entries and hit[1] is wrs
    slab_stagers = [w.buffer_stager for w in wrs if w.path.startswith("batched/")]
    assert all(any(t is a for a in again) for st in slab_stagers for t in st.tensors), "the stagers hold the new aliases"
    solo = [w.buffer_stager for w in wrs if not w.path.startswith("batched/")]
    assert solo and all(any(st.tensor is a for a in again) for st in solo)
    # anything different: no hit
    assert cache.get(("other",), paths, again) is None
    assert cache.get(cfg, paths[:-1], again[:-1]) is None
    moved = list(again)
    moved[3] = torch.zeros_like(moved[3])  # same shape, another address
    assert cache.get(cfg, paths, moved) is None
    reshaped = list(again)
    reshaped[5] = reshaped[5].view(-1)  # same address and bytes, another shape
    assert cache.get(cfg, paths, reshaped) is None
    restrided = list(again)
    i = paths.index("view")
    restrided[i] = torch.zeros(50, 8)[:, ::2]
    assert cache.get(cfg, paths, restrided) is None  # another base tensor
    other_stride = list(again)
    other_stride[i] = again[i].as_strided(again[i].shape, (4, 1))  # same address and shape, other strides
    assert cache.get(cfg, paths, other_stride) is None
    retyped = list(again)
    retyped[7] = retyped[7].view(torch.int32)
    assert cache.get(cfg, paths, retyped) is None
    assert cache.get(cfg, paths, again) is not None


def test_bulk_read_planner_keeps_devices_apart():
    """Members of one slab file bound for two devices never share a read (one consumer = one device's staging
    block and one launch): the device column of the table joins the sort key of plan_cuda_read_columns."""
    import numpy as np

    from torchsnapshot_b200 import tensor_table as tt
    from torchsnapshot_b200.batcher import plan_cuda_read_columns

    tensors = [torch.zeros(100, 2) for _ in range(40)]
    table = tt.scan(tensors).copy()
    table[:, tt.DEVICE] = np.arange(40) % 2  # pretend: even members on device 0, odd ones on device 1
    lo = np.arange(40, dtype=np.int64) * 800
    hi = lo + 800
    file_id = np.zeros(40, dtype=np.int64)
    reads = plan_cuda_read_columns(["batched/f"], file_id, lo, hi, tensors, table)
    assert len(reads) == 2 and sorted(rr.buffer_consumer.device for rr in reads) == [0, 1]
    for rr in reads:
        dev = rr.buffer_consumer.device
        members = [i for i in range(40) if i % 2 == dev]
        assert rr.byte_range == (int(lo[members[0]]), int(hi[members[-1]]))
        assert [t is tensors[i] for t, i in zip(rr.buffer_consumer.tensors, members)] == [True] * 20
        assert (rr.buffer_consumer._offsets == lo[members] - lo[members[0]]).all()


def test_restore_view_splits_a_manifest_into_columns_and_the_rest():
    """snapshot._RestoreView: slab members of buffer-protocol dtypes under the prefix become columns, containers
    are kept apart, everything else (whole-file tensors, pickled tensors, objects, sharded and chunked entries,
    other statefuls) is left to the per-entry planner."""
    from torchsnapshot_b200 import tensor_table as tt
    from torchsnapshot_b200.manifest import (ChunkedTensorEntry, DictEntry, ObjectEntry, PrimitiveEntry, Shard,
                                             ShardedTensorEntry)
    from torchsnapshot_b200.snapshot import _RestoreView

    def tensor(location, dtype, shape, byte_range, serializer="buffer_protocol", replicated=False):
        return TensorEntry(location=location, serializer=serializer, dtype=dtype, shape=shape, replicated=replicated,
                           byte_range=byte_range)

    manifest = {
        "m": DictEntry(keys=["a", "b", "c", "d", "e", "f", "g", "h"]),
        "m/a": tensor("batched/x", "torch.float32", [3, 4], [0, 48]),
        "m/b": tensor("batched/x", "torch.bfloat16", [], [48, 50], replicated=True),
        "m/c": tensor("0/m/c", "torch.float32", [1000, 1000], None),  # a file of its own
        "m/d": tensor("batched/y", "torch.complex64", [2], [0, 99], serializer="torch_save"),
        "m/e": ObjectEntry(location="0/m/e", serializer="torch_save", obj_type="tuple", replicated=False),
        "m/f": PrimitiveEntry.from_object(5),
        "m/g": ShardedTensorEntry(shards=[Shard(offsets=[0], sizes=[4], tensor=tensor("sharded/g_0", "torch.int64", [4], [50, 82]))]),
        "m/h": ChunkedTensorEntry(dtype="torch.float32", shape=[8], chunks=[], replicated=False),
        "m/i": tensor("batched/y", "torch.int64", [2, 2, 2, 2, 2, 2, 2, 2, 2], [100, 4196]),  # 9 dims: not a row of the table
        "other": DictEntry(keys=["a"]),
        "other/a": tensor("batched/x", "torch.float32", [1], [82, 86]),
        "mm/a": tensor("batched/x", "torch.float32", [1], [86, 90]),  # "mm" is not under "m"
    }
    view = _RestoreView(manifest, {}, "m")
    assert view.paths == ["m/a", "m/b"] and [e is manifest[p] for e, p in zip(view.entries, view.paths)] == [True, True]
    assert list(view.containers) == ["m"] and view.has_sharded
    assert [p for p, _e in view.rest] == ["m/c", "m/d", "m/e", "m/f", "m/g", "m/h", "m/i"]
    assert view.dtype_codes.tolist() == [tt.dtype_code(torch.float32), tt.dtype_code(torch.bfloat16)]
    assert view.ndims.tolist() == [2, 0] and view.shapes[0, :3].tolist() == [3, 4, 0] and not view.shapes[1].any()
    assert view.los.tolist() == [0, 48] and view.his.tolist() == [48, 50]
    assert view.locations == ["batched/x"] and view.file_ids.tolist() == [0, 0] and view.replicated.tolist() == [False, True]
    other = _RestoreView(manifest, {}, "other")
    assert other.paths == ["other/a"] and not other.has_sharded and not other.rest
    empty = _RestoreView(manifest, {}, "absent")
    assert empty.paths == [] and empty.shapes.shape == (0, tt.MAX_DIMS) and empty.los.shape == (0,)


def test_streaming_batches_count_large_files_half():
    """scheduler.streaming_batch_len: 16 slab-sized files per interleaved batch with 32 writers, up to 32 files
    of 256 MiB and more, mixtures in between; never an empty batch."""
    from torchsnapshot_b200.scheduler import streaming_batch_len

    slab, big = 128 << 20, 400 << 20
    assert streaming_batch_len([slab] * 60, 16) == 16
    assert streaming_batch_len([big] * 20, 16) == 20
    assert streaming_batch_len([big] * 40, 16) == 32
    assert streaming_batch_len([big] * 3 + [slab] * 60, 16) == 3 + 14  # 3 + 2 * 14 = 31 of 32
    assert streaming_batch_len([slab] * 5, 16) == 5
    assert streaming_batch_len([slab], 1) == 1 and streaming_batch_len([slab, slab], 1) == 1
    assert streaming_batch_len([big, slab], 1) == 1 and streaming_batch_len([big, big, big], 1) == 2
    assert streaming_batch_len([], 16) == 0
    assert streaming_batch_len([256 << 20] * 40, 8) == 16 and streaming_batch_len([(256 << 20) - 1] * 40, 8) == 8
