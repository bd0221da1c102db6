"""End-to-end take / restore / read_object of CUDA state through the B200 path, against the
reference-made fixtures and the CPU oracle.  Bit-exact comparisons."""

import asyncio
import hashlib
import io
import json
import os

import numpy as np
import pytest
import torch

import _cases
import oracle
from test_snapshot_cpu import _c1_app_state, _c1_inputs, _normalise, _payloads
from torchsnapshot_b200 import knobs
from torchsnapshot_b200.io_types import ReadIO, StoragePlugin, WriteIO
from torchsnapshot_b200.snapshot import Snapshot, StateDict

pytestmark = pytest.mark.gpu


def _ranged(buf, byte_range):
    return buf if byte_range is None else buf[byte_range[0] : byte_range[1]]


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def _bytes(t: torch.Tensor) -> bytes:
    return oracle.serialize(t.detach().cpu())


def _to_cuda(app_state):
    app_state["model"].cuda()
    optim = app_state["optim"]
    for st in optim.state.values():
        for k, v in st.items():
            if isinstance(v, torch.Tensor):
                st[k] = v.cuda()
    return app_state


def test_c1_on_gpu_matches_reference_snapshot(golden_dir, tmp_path, engine):
    """BASELINE config 0 with the state on the GPU: payload bytes and manifest equal the snapshot
    the reference wrote for the same state."""
    ref_root = os.path.join(golden_dir, "c1_simple", "snapshot")
    app_state = _to_cuda(_c1_app_state(inputs=_c1_inputs(golden_dir)))
    launches = engine.launches
    ours = str(tmp_path / "snap")
    Snapshot.take(ours, app_state)
    assert engine.launches > launches, "the CUDA path did not run"
    ref_payloads, ref_meta = _payloads(ref_root)
    our_payloads, our_meta = _payloads(ours)
    assert our_payloads == ref_payloads
    assert _normalise(our_meta) == _normalise(ref_meta)
    # restore the reference-written snapshot into CUDA state
    inputs = torch.load(os.path.join(golden_dir, "c1_simple", "inputs.pt"), weights_only=False)
    fresh = _to_cuda(_c1_app_state(seed=7))
    Snapshot(ref_root).restore(fresh)
    for name, t in fresh["model"].state_dict().items():
        assert t.is_cuda and _bytes(t) == oracle.serialize(inputs["tensors"][f"model/{name}"])
    st = fresh["optim"].state_dict()["state"]
    assert _bytes(st[2]["exp_avg_sq"]) == oracle.serialize(inputs["tensors"]["optim/state/2/exp_avg_sq"])


def _to_cuda_like_default_adam(app_state):
    """What a default (non-capturable) Adam leaves behind: parameters and moments on the GPU, `step` on the CPU."""
    app_state["model"].cuda()
    for st in app_state["optim"].state.values():
        for k, v in st.items():
            if isinstance(v, torch.Tensor) and k != "step":
                st[k] = v.cuda()
    return app_state


def test_c1_with_adam_step_counters_on_the_cpu(golden_dir, tmp_path, engine):
    """Mixed devices in one state dict (SURVEY 8d, config C2 variant): CPU members go to a CPU slab, CUDA
    members to a GPU slab (batcher.py:291-319); every object's payload still equals the reference's, and the
    snapshot restores into a state with the same device mix."""
    ref_root = os.path.join(golden_dir, "c1_simple", "snapshot")
    app_state = _to_cuda_like_default_adam(_c1_app_state(inputs=_c1_inputs(golden_dir)))
    steps = [st["step"] for st in app_state["optim"].state.values()]
    assert steps and all(isinstance(t, torch.Tensor) and not t.is_cuda for t in steps)
    launches = engine.launches
    ours = str(tmp_path / "snap")
    snap = Snapshot.take(ours, app_state)
    assert engine.launches > launches, "the CUDA path did not run"
    ref_payloads, _ref_meta = _payloads(ref_root)
    our_payloads, our_meta = _payloads(ours)
    assert our_payloads == ref_payloads
    tensors = {k: e for k, e in our_meta["manifest"].items() if e["type"] == "Tensor"}
    step_files = {e["location"] for k, e in tensors.items() if k.endswith("/step")}
    other_files = {e["location"] for k, e in tensors.items() if not k.endswith("/step")}
    assert len(step_files) == 1 and len(other_files) == 1 and step_files != other_files  # one CPU slab, one GPU slab
    fresh = _to_cuda_like_default_adam(_c1_app_state(seed=7))
    snap.restore(fresh)
    inputs = torch.load(os.path.join(golden_dir, "c1_simple", "inputs.pt"), weights_only=False)
    for name, t in fresh["model"].state_dict().items():
        assert t.is_cuda and _bytes(t) == oracle.serialize(inputs["tensors"][f"model/{name}"])
    for idx, st in fresh["optim"].state_dict()["state"].items():
        assert not st["step"].is_cuda and st["exp_avg"].is_cuda
        for k in ("step", "exp_avg", "exp_avg_sq"):
            assert _bytes(st[k]) == oracle.serialize(inputs["tensors"][f"optim/state/{idx}/{k}"]), (idx, k)


@pytest.mark.parametrize("fs_direct", ["0", "1"])
@pytest.mark.parametrize("case", ["mixed", "threshold"])
def test_gpu_take_matches_reference_fixture(golden_dir, tmp_path, case, fs_direct, monkeypatch):
    # fs_direct=1: the FS plugin moves the page-aligned body of each file with O_DIRECT; bytes are the same
    monkeypatch.setenv("TSB200_FS_DIRECT", fs_direct)
    gold = json.load(open(os.path.join(golden_dir, f"slab_{case}.json")))
    threshold, items = getattr(_cases, f"slab_case_{case}")()
    state = StateDict({p: _cases.build_view(spec, "cuda") for p, spec in items})
    with knobs.override_slab_size_threshold_bytes(threshold):
        snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
    payloads, meta = _payloads(str(tmp_path / "s"))
    slab_names = []
    for p, _spec in items:
        e = meta["manifest"][f"0/st/{p}"]
        g = gold["objects"][p]
        assert payloads[f"0/st/{p}"] == g["sha256"], p
        assert e["byte_range"] == g["byte_range"] and e["shape"] == g["shape"] and e["dtype"] == g["dtype"]
        if g["slab"] is not None:
            if e["location"] not in slab_names:
                slab_names.append(e["location"])
            assert slab_names.index(e["location"]) == g["slab"]
    for name, g in zip(slab_names, gold["slabs"]):
        blob = open(os.path.join(str(tmp_path / "s"), name), "rb").read()
        assert len(blob) == g["size"] and sha(blob) == g["sha256"]
    # restore into strided CUDA twins
    out = StateDict({p: _cases.build_view({**spec, "seed": spec["seed"] + 5}, "cuda") for p, spec in items})
    snap.restore({"st": out})
    for p, _spec in items:
        assert _bytes(out[p]) == _bytes(state[p]), p
    # read_object of one member of a slab, tiled
    obj = snap.read_object("0/st/big_f32" if case == "mixed" else "0/st/over_threshold",
                           obj_out=torch.zeros_like(state["big_f32" if case == "mixed" else "over_threshold"]),
                           memory_budget_bytes=10000)
    assert _bytes(obj) == _bytes(state["big_f32" if case == "mixed" else "over_threshold"])


class _Fake:
    def __init__(self, boxes, dtype, shape, device, full=None):
        from types import SimpleNamespace as NS

        self._shape = list(shape)
        self._shards = []
        for o, s in boxes:
            t = (_cases.box_slice(full, (o, s)).contiguous() if full is not None else torch.zeros(s, dtype=dtype)).to(device)
            self._shards.append(NS(tensor=t, metadata=NS(shard_offsets=list(o), shard_sizes=list(s))))

    def local_shards(self):
        return self._shards


def test_gpu_reshard_all_layout_pairs(golden_dir, monkeypatch):
    """Persisted pieces (host bytes) -> pinned -> H2D -> strided box kernel into CUDA shards, for every
    src x dst layout of the reference-made reshard fixture."""
    import torchsnapshot_b200.io_preparer as iop
    from test_host_planner import _sharded_entry, run

    gold = json.load(open(os.path.join(golden_dir, "sharded.json")))
    shape = _cases.SHARD_GLOBAL_SHAPE
    full = _cases.raw_tensor(_cases.SHARD_SEED, shape, _cases.DTYPES[_cases.SHARD_DTYPE])
    layouts = _cases.shard_layouts(shape)
    monkeypatch.setattr(iop, "_is_sharded_tensor", lambda o: isinstance(o, _Fake))
    monkeypatch.setattr(_Fake, "metadata", lambda self: type("M", (), {"size": self._shape})(), raising=False)
    for key, g in gold["reshard"].items():
        src_name, dst_name = key.split("->")
        # stage the source pieces from CUDA shards (pack kernel when the piece is a strided view)
        dim, boxes = layouts[src_name]
        shards, wrs = [], []
        for off, sizes in boxes:
            local = _cases.box_slice(full, (off, sizes)).contiguous().cuda()
            for tensor, o, s in iop.ShardedTensorIOPreparer.subdivide_shard(local, list(off), list(sizes), dim,
                                                                            _cases.SHARD_MAX_BYTES):
                e, w = iop.TensorIOPreparer.prepare_write(f"sharded/foo_{'_'.join(map(str, o))}", tensor)
                shards.append(iop.Shard(offsets=o, sizes=s, tensor=e))
                wrs += w
        entry = iop.ShardedTensorEntry(shards=shards)
        staged = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in wrs}
        assert [sha(staged[s.tensor.location]) for s in entry.shards] == [p["sha256"] for p in gold["writes"][src_name]["pieces"]]
        dst = _Fake(layouts[dst_name][1], full.dtype, shape, "cuda")
        rrs, _ = iop.ShardedTensorIOPreparer.prepare_read(entry, dst)
        for rr in rrs:
            assert rr.buffer_consumer.is_device_consumable()
            run(rr.buffer_consumer.consume_buffer(_ranged(staged[rr.path], rr.byte_range)))
        assert [sha(_bytes(s.tensor)) for s in dst.local_shards()] == g["dst_sha256"], key


def test_waves_and_small_arenas(tmp_path, monkeypatch):
    """A device staging bound smaller than the payload forces several pack waves; a pinned bound of a few
    slabs forces the D2H pipeline to wait for storage writes.  Result must be unchanged."""
    from torchsnapshot_b200 import engine as engmod

    engmod.shutdown_engines()
    monkeypatch.setenv("TSB200_DEVICE_STAGING_BYTES", str(96 << 20))
    monkeypatch.setenv("TSB200_PINNED_BYTES", str(24 << 20))
    try:
        g = torch.Generator(device="cuda").manual_seed(0)
        state = StateDict({f"t{i}": torch.randint(0, 255, (int(n),), dtype=torch.uint8, device="cuda", generator=g)
                           for i, n in enumerate(np.random.RandomState(0).randint(1, 3 << 20, size=60))})
        state["big"] = torch.randint(0, 255, (20 << 20,), dtype=torch.uint8, device="cuda", generator=g)
        with knobs.override_slab_size_threshold_bytes(8 << 20):
            snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
            out = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
            snap.restore({"st": out})
        for k in state:
            assert torch.equal(out[k], state[k]), k
    finally:
        engmod.shutdown_engines()


def test_no_device_staging_packs_into_and_unpacks_from_pinned_memory(tmp_path, monkeypatch):
    """Requests larger than the device staging bound (or a trainer that has filled the HBM since the bound was
    taken): the pack kernel stores straight into the pinned block and the unpack kernel reads straight out of
    it -- no staging copy in HBM, same bytes.  The reference falls back to a CPU slab here (batcher.py:145-149)."""
    from torchsnapshot_b200 import engine as engmod
    from torchsnapshot_b200 import scheduler

    engmod.shutdown_engines()
    monkeypatch.setenv("TSB200_DEVICE_STAGING_BYTES", str(16 << 20))
    try:
        g = torch.Generator(device="cuda").manual_seed(5)
        base = torch.randint(0, 2**31 - 1, (3000, 4096), dtype=torch.int32, device="cuda", generator=g)
        state = StateDict(
            dense=torch.randint(0, 255, (40 << 20,), dtype=torch.uint8, device="cuda", generator=g),  # > bound
            view=base[:, ::2],  # 24 MiB, non-contiguous: strided source, host destination
            small=torch.randn(1000, device="cuda", generator=g),
        )
        with knobs.override_slab_size_threshold_bytes(1 << 20):
            snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
            assert scheduler.last_write_stats.host_packed_bytes == (40 << 20) + 3000 * 2048 * 4
            for name, t in (("dense", state["dense"]), ("view", state["view"])):
                assert (tmp_path / "s" / "0" / "st" / name).read_bytes() == _bytes(t), name
            out_base = torch.zeros_like(base)
            out = StateDict(dense=torch.zeros_like(state["dense"]), view=out_base[:, ::2], small=torch.zeros(1000, device="cuda"))
            snap.restore({"st": out})
            rs = scheduler.last_read_stats
            # the dense tensor (and, when it is planned per entry, the one-member slab of the small one): H2D
            # straight into the destination
            assert rs.direct_bytes in (40 << 20, (40 << 20) + 4000)
            assert rs.host_unpacked_bytes == 3000 * 2048 * 4  # the strided one: unpack kernel reading pinned memory
        for k in state:
            assert torch.equal(out[k], state[k]), k
        assert not out_base[:, 1::2].any(), "the unpack must not touch the elements between the view's columns"
    finally:
        engmod.shutdown_engines()


def test_repeated_takes_reuse_the_plan_and_follow_changes(tmp_path):
    """Checkpointing the same tensors again reuses the previous bulk plan (same entries, same slab layout); new
    values in the same tensors are what gets saved; a tensor that was replaced, reshaped or re-strided makes the
    planner start over.  Every snapshot restores to exactly what the state held when it was taken."""
    from torchsnapshot_b200.batcher import take_plan_cache

    g = torch.Generator(device="cuda").manual_seed(9)
    state = StateDict({f"t{i}": torch.randn(100 + 3 * i, 5, device="cuda", generator=g) for i in range(90)})
    state["view"] = torch.randn(64, 10, device="cuda", generator=g)[:, ::2]
    state["n"] = 3

    def check(snap, want):
        out = StateDict({k: (torch.zeros_like(v) if isinstance(v, torch.Tensor) else 0) for k, v in want.items()})
        snap.restore({"st": out})
        for k, v in want.items():
            assert torch.equal(out[k], v) if isinstance(v, torch.Tensor) else out[k] == v, k

    with knobs.override_slab_size_threshold_bytes(1 << 16):
        h0 = take_plan_cache.hits
        s1 = Snapshot.take(str(tmp_path / "s1"), {"st": state})
        want1 = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in state.items()}
        assert take_plan_cache.hits == h0
        for v in state.values():
            if isinstance(v, torch.Tensor):
                v.add_(1.0)  # training goes on: same tensors, new values
        state["n"] = 4
        s2 = Snapshot.take(str(tmp_path / "s2"), {"st": state})
        assert take_plan_cache.hits == h0 + 1
        want2 = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in state.items()}
        m1, m2 = s1.get_manifest(), s2.get_manifest()
        assert {k: (e.location, e.byte_range) for k, e in m1.items() if hasattr(e, "byte_range")} == {
            k: (e.location, e.byte_range) for k, e in m2.items() if hasattr(e, "byte_range")}
        check(s1, want1)
        check(s2, want2)
        # the on-disk metadata of the second snapshot is what a fresh process reads
        check(Snapshot(str(tmp_path / "s2")), want2)
        state["t7"] = torch.randn(121, 5, device="cuda", generator=g)  # replaced: another address
        s3 = Snapshot.take(str(tmp_path / "s3"), {"st": state})
        assert take_plan_cache.hits == h0 + 1
        check(s3, dict(state))
        state["t9"] = state["t9"].view(-1)  # same memory, another shape
        s4 = Snapshot.take(str(tmp_path / "s4"), {"st": state})
        assert take_plan_cache.hits == h0 + 1 and s4.get_manifest()["0/st/t9"].shape == [127 * 5]
        check(s4, dict(state))
        s5 = Snapshot.take(str(tmp_path / "s5"), {"st": state})
        assert take_plan_cache.hits == h0 + 2
        check(s5, dict(state))


def test_bulk_planned_state_with_edge_cases(tmp_path, monkeypatch):
    """A state dict large enough for the bulk planners, with the shapes that trip byte arithmetic: empty tensors,
    0-d scalars, odd byte counts, strided and transposed views, every element size, a tensor of exactly the slab
    threshold and one just below it.  Payload bytes per object equal the oracle's serialisation, the plan equals
    the per-leaf planner's (TSB200_PLAN_CACHE off and a state below the bulk minimum take that path), and the
    round trip -- into strided destinations too -- is exact."""
    from torchsnapshot_b200 import snapshot as snapmod

    threshold = 1 << 16
    g = torch.Generator(device="cuda").manual_seed(33)

    def rnd(shape, dtype):
        es = torch.empty((), dtype=dtype).element_size()
        n = 1
        for x in shape:
            n *= x
        raw = torch.randint(0, 256, (max(n * es, 1),), dtype=torch.uint8, device="cuda", generator=g)[: n * es]
        return raw.view(dtype).reshape(shape)

    state = {}
    dtypes = [torch.uint8, torch.int16, torch.bfloat16, torch.float32, torch.int64, torch.float64, torch.bool]
    for i in range(90):
        dt = dtypes[i % len(dtypes)]
        state[f"t{i}"] = (rnd([37 + 11 * i], torch.uint8) > 127) if dt is torch.bool else rnd([37 + 11 * i], dt)
    state["empty"] = torch.empty(0, 5, device="cuda")
    state["empty1d"] = torch.empty(0, dtype=torch.int64, device="cuda")
    state["scalar"] = torch.tensor(3.25, device="cuda")
    state["odd"] = rnd([3, 7, 5], torch.uint8)
    state["strided"] = rnd([64, 40], torch.float32)[:, ::2]
    state["transposed"] = rnd([48, 80], torch.int16).t()
    state["at_threshold"] = rnd([threshold], torch.uint8)  # a request of its own (>= threshold)
    state["below_threshold"] = rnd([threshold - 1], torch.uint8)
    assert len(snapmod._bulk_plannable(state, set(), 1, None, 512 << 20).paths) == len(state)

    def take(path, cached):
        monkeypatch.setenv("TSB200_PLAN_CACHE", "1" if cached else "0")
        with knobs.override_slab_size_threshold_bytes(threshold):
            return Snapshot.take(path, {"st": StateDict(state)})

    snap = take(str(tmp_path / "bulk"), True)
    payloads, meta = _payloads(str(tmp_path / "bulk"))
    for k, t in state.items():
        assert payloads[f"0/st/{k}"] == sha(_bytes(t)), k
    assert meta["manifest"]["0/st/at_threshold"]["location"] == "0/st/at_threshold"
    assert meta["manifest"]["0/st/below_threshold"]["location"].startswith("batched/")
    # the per-leaf planner on the same state: same manifest modulo slab names
    import torchsnapshot_b200.snapshot as sm

    monkeypatch.setattr(sm, "_bulk_plannable", lambda *a, **k: None)
    take(str(tmp_path / "leaf"), False)
    _p2, meta_leaf = _payloads(str(tmp_path / "leaf"))
    assert _normalise(meta) == _normalise(meta_leaf)
    monkeypatch.undo()
    # round trip, with strided destinations where the sources were strided
    out = {k: torch.zeros_like(v) for k, v in state.items()}
    out["strided"] = torch.zeros(64, 40, device="cuda")[:, ::2]
    out["transposed"] = torch.zeros(48, 80, dtype=torch.int16, device="cuda").t()
    out_sd = StateDict(out)
    snap.restore({"st": out_sd})
    for k, t in state.items():
        assert out_sd[k].shape == t.shape and _bytes(out_sd[k]) == _bytes(t), k


def test_restore_into_mixed_destinations_matches_the_per_entry_planner(tmp_path):
    """tests/_mixed_restore.py on CUDA tensors: slab members planned over columns where the destination matches,
    entry by entry where it does not, the same values either way."""
    import _mixed_restore

    _mixed_restore.run("cuda", tmp_path)


def test_overlapping_async_takes(tmp_path):
    """A second async_take that begins while the first is still copying and writing in its background thread: the
    second repeats the first one's plan (same tensors), both drain, each snapshot holds what the state held when
    its take returned."""
    g = torch.Generator(device="cuda").manual_seed(21)
    state = StateDict({f"t{i}": torch.randn(1 << 20, device="cuda", generator=g) for i in range(80)})
    want1 = {k: v.clone() for k, v in state.items()}
    p1 = Snapshot.async_take(str(tmp_path / "a"), {"st": state})
    for v in state.values():
        v.mul_(-2.0)
    want2 = {k: v.clone() for k, v in state.items()}
    p2 = Snapshot.async_take(str(tmp_path / "b"), {"st": state})
    for v in state.values():
        v.zero_()
    s2, s1 = p2.wait(), p1.wait()
    for snap, want in ((s1, want1), (s2, want2)):
        out = StateDict({k: torch.zeros_like(v) for k, v in want.items()})
        snap.restore({"st": out})
        for k in want:
            assert torch.equal(out[k], want[k]), k


def test_async_take_is_consistent_under_mutation(tmp_path):
    g = torch.Generator(device="cuda").manual_seed(1)
    state = StateDict({f"t{i}": torch.randn(1 << 18, device="cuda", generator=g) for i in range(40)})
    want = {k: v.clone() for k, v in state.items()}
    pending = Snapshot.async_take(str(tmp_path / "s"), {"st": state})
    for v in state.values():
        v.zero_()  # training continues and overwrites the parameters
    snap = pending.wait()
    out = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
    snap.restore({"st": out})
    for k in want:
        assert torch.equal(out[k], want[k]), k


class _MemPlugin(StoragePlugin):
    """A third-party style plugin: plain StoragePlugin ABI, no read_into (bytes come back in a BytesIO)."""

    def __init__(self):
        self.files = {}

    async def write(self, write_io: WriteIO) -> None:
        assert isinstance(write_io.buf, (bytes, memoryview))
        self.files[write_io.path] = bytes(write_io.buf)

    async def read(self, read_io: ReadIO) -> None:
        data = self.files[read_io.path]
        if read_io.byte_range is not None:
            data = data[read_io.byte_range[0] : read_io.byte_range[1]]
        read_io.buf = io.BytesIO(data)

    async def delete(self, path: str) -> None:
        del self.files[path]

    async def close(self) -> None:
        pass


def test_third_party_storage_plugin_abi(monkeypatch):
    import torchsnapshot_b200.snapshot as snapmod

    plugin = _MemPlugin()
    monkeypatch.setattr(snapmod, "url_to_storage_plugin", lambda path, opts=None: plugin)
    threshold, items = _cases.slab_case_mixed()
    state = StateDict({p: _cases.build_view(spec, "cuda") for p, spec in items})
    snap = Snapshot.take("mem://x", {"st": state})
    assert ".snapshot_metadata" in plugin.files
    out = StateDict({p: torch.zeros_like(t) for p, t in state.items()})
    snap.restore({"st": out})
    for p in state:
        assert _bytes(out[p]) == _bytes(state[p]), p


def test_non_batched_large_tensors_and_chunks(tmp_path):
    g = torch.Generator(device="cuda").manual_seed(2)
    big = torch.randn(1000, 999, device="cuda", generator=g)          # > slab threshold below: written on its own
    huge = torch.randn(4000, 1000, device="cuda", generator=g)         # > chunk size below: chunked
    view = torch.randn(1200, 1200, device="cuda", generator=g)[:, 100:1100]  # strided and unbatched
    state = StateDict(big=big, huge=huge, view=view, small=torch.arange(5, device="cuda"))
    with knobs.override_slab_size_threshold_bytes(1 << 20), knobs.override_max_chunk_size_bytes(6 << 20):
        snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
        m = snap.get_manifest()
        assert m["0/st/big"].location == "0/st/big" and m["0/st/huge"].type == "ChunkedTensor"
        assert open(os.path.join(str(tmp_path / "s"), "0/st/big"), "rb").read() == _bytes(big)
        assert open(os.path.join(str(tmp_path / "s"), "0/st/view"), "rb").read() == _bytes(view)
        out = StateDict(big=torch.zeros_like(big), huge=torch.zeros_like(huge), view=torch.zeros(1200, 1000, device="cuda"),
                        small=torch.zeros(5, dtype=torch.int64, device="cuda"))
        snap.restore({"st": out})
    for k in state:
        assert torch.equal(out[k], state[k]), k


def test_read_object_row_block_from_sharded_table(tmp_path, monkeypatch):
    """BASELINE config 4 style random access: read a block of rows of a row-wise sharded table into a
    CUDA tensor.  Only the needed rows of the persisted pieces are read from storage."""
    import torchsnapshot_b200.io_preparer as iop
    import torchsnapshot_b200.snapshot as snapmod
    from torchsnapshot_b200.storage_plugins.fs import FSStoragePlugin

    rows, cols = 60000, 128
    full = _cases.raw_tensor(321, [rows, cols], torch.float32)
    monkeypatch.setattr(iop, "_is_sharded_tensor", lambda o: isinstance(o, _Fake))
    monkeypatch.setattr(snapmod, "_is_sharded_tensor", lambda o: isinstance(o, _Fake))
    monkeypatch.setattr(_Fake, "metadata", lambda self: type("M", (), {"size": self._shape})(), raising=False)
    monkeypatch.setattr(_Fake, "sharding_spec", lambda self: None, raising=False)
    table = _Fake(_cases.chunk_boxes([rows, cols], 0, 4), full.dtype, [rows, cols], "cuda", full)
    with knobs.override_max_shard_size_bytes(4 << 20), knobs.override_slab_size_threshold_bytes(1 << 20):
        snap = Snapshot.take(str(tmp_path / "s"), {"emb": StateDict(t=table)})
    read_bytes = []

    class Counting(FSStoragePlugin):
        async def read_into(self, path, byte_range, out):
            read_bytes.append(len(out))
            await super().read_into(path, byte_range, out)

    monkeypatch.setattr(snapmod, "url_to_storage_plugin", lambda path, opts=None: Counting(path))
    out = torch.zeros(5000, cols, device="cuda")  # rows 0..5000 of the global table
    got = snap.read_object("0/emb/t", obj_out=out)
    assert got is out and _bytes(out) == oracle.serialize(full[:5000])
    assert sum(read_bytes) == 5000 * cols * 4, (sum(read_bytes), len(read_bytes))  # not the 4 MiB pieces
    # the whole table into a dense CUDA tensor
    dense = torch.zeros(rows, cols, device="cuda")
    snap.read_object("0/emb/t", obj_out=dense)
    assert _bytes(dense) == oracle.serialize(full)


def test_memory_budget_bounds_pinned_bytes_in_flight(tmp_path, monkeypatch):
    """TORCHSNAPSHOT_PER_RANK_MEMORY_BUDGET_BYTES (the reference's knob, scheduler.py:46-56) bounds the pinned
    host bytes in flight on take and on restore; the result is unchanged."""
    import torchsnapshot_b200.snapshot as snapmod
    from torchsnapshot_b200.engine import get_engine
    from torchsnapshot_b200.storage_plugins.fs import FSStoragePlugin

    budget = 24 << 20
    monkeypatch.setenv("TORCHSNAPSHOT_PER_RANK_MEMORY_BUDGET_BYTES", str(budget))
    eng = get_engine(0)
    peak = [0]

    class Watching(FSStoragePlugin):
        async def write(self, write_io):
            peak[0] = max(peak[0], eng.host_in_use())
            await super().write(write_io)

        async def read_into(self, path, byte_range, out):
            peak[0] = max(peak[0], eng.host_in_use())
            await super().read_into(path, byte_range, out)

    monkeypatch.setattr(snapmod, "url_to_storage_plugin", lambda path, opts=None: Watching(path))
    g = torch.Generator(device="cuda").manual_seed(3)
    state = StateDict({f"t{i}": torch.randint(0, 255, (3 << 20,), dtype=torch.uint8, device="cuda", generator=g)
                       for i in range(40)})
    with knobs.override_slab_size_threshold_bytes(8 << 20):
        snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
        take_peak, peak[0] = peak[0], 0
        out = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
        snap.restore({"st": out})
    slab = 8 << 20
    assert 0 < take_peak <= budget + slab, take_peak
    assert 0 < peak[0] <= budget + slab, peak[0]
    for k in state:
        assert torch.equal(out[k], state[k]), k


def test_batching_disabled_on_gpu(golden_dir, tmp_path):
    """TORCHSNAPSHOT_DISABLE_BATCHING: one file per tensor (reference knobs.py:52-55); the tensors still go
    through one pack launch and the payload bytes are the reference's."""
    gold = json.load(open(os.path.join(golden_dir, "slab_mixed.json")))
    _threshold, items = _cases.slab_case_mixed()
    state = StateDict({p: _cases.build_view(spec, "cuda") for p, spec in items})
    with knobs.override_is_batching_disabled(True):
        snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
        payloads, meta = _payloads(str(tmp_path / "s"))
        for p, _spec in items:
            e = meta["manifest"][f"0/st/{p}"]
            assert e["location"] == f"0/st/{p}" and e["byte_range"] is None
            assert payloads[f"0/st/{p}"] == gold["objects"][p]["sha256"], p
        out = StateDict({p: torch.zeros_like(t) for p, t in state.items()})
        snap.restore({"st": out})
    for p in state:
        assert _bytes(out[p]) == _bytes(state[p]), p


def test_overlapping_async_and_sync_takes(tmp_path):
    """An async take still draining in its background thread while the next take starts: the arenas and
    streams are process-wide and shared (SURVEY 8b threading); both snapshots must be complete and exact."""
    g = torch.Generator(device="cuda").manual_seed(4)
    a = StateDict({f"a{i}": torch.randn(1 << 20, device="cuda", generator=g) for i in range(24)})
    b = StateDict({f"b{i}": torch.randn(1 << 19, device="cuda", generator=g) for i in range(24)})
    want_a = {k: v.clone() for k, v in a.items()}
    pending = Snapshot.async_take(str(tmp_path / "a"), {"s": a})
    for v in a.values():
        v.add_(1.0)
    snap_b = Snapshot.take(str(tmp_path / "b"), {"s": b})
    pending2 = Snapshot.async_take(str(tmp_path / "a2"), {"s": a})
    snap_a = pending.wait()
    snap_a2 = pending2.wait()
    out_a = StateDict({k: torch.zeros_like(v) for k, v in a.items()})
    snap_a.restore({"s": out_a})
    for k in want_a:
        assert torch.equal(out_a[k], want_a[k]), k
    snap_a2.restore({"s": out_a})
    for k in a:
        assert torch.equal(out_a[k], a[k]), k
    out_b = StateDict({k: torch.zeros_like(v) for k, v in b.items()})
    snap_b.restore({"s": out_b})
    for k in b:
        assert torch.equal(out_b[k], b[k]), k


def test_full_size_ddp_state_take_and_restore(tmp_path_factory):
    """BASELINE config 1 at full size (2312 tensors, 9.35 GB) through Snapshot.take / restore: a checksum per
    object (sum of its bytes, device vs file) and a bit-exact round trip -- properties that do not need the
    oracle to serialise gigabytes."""
    import shutil
    import sys
    import tempfile

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    import bench

    free, _ = torch.cuda.mem_get_info()
    if free < 40 << 30:
        pytest.skip("needs 40 GB of free HBM")
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    d = tempfile.mkdtemp(prefix="tsb_full_", dir=base)
    try:
        state = bench.build_state("cuda:0")
        snap = Snapshot.take(d, {"ddp": StateDict(state)}, replicated=["**"])
        meta = json.load(open(os.path.join(d, ".snapshot_metadata")))["manifest"]
        files = {}
        checked = 0
        for name, t in state.items():
            e = meta["0/ddp/" + name.replace("/", "%2F")]
            assert e["replicated"] is True and e["shape"] == list(t.shape)
            if e["location"] not in files:
                files = {e["location"]: np.fromfile(os.path.join(d, e["location"]), dtype=np.uint8)}  # one slab at a time
            buf = files[e["location"]]
            lo, hi = e["byte_range"] if e["byte_range"] is not None else (0, len(buf))
            assert hi - lo == t.numel() * t.element_size()
            want = int(t.reshape(-1).view(torch.uint8).sum(dtype=torch.int64))
            assert int(buf[lo:hi].sum(dtype=np.int64)) == want, name
            checked += hi - lo
        assert checked == bench.payload_bytes(state)
        out = StateDict({k: torch.zeros_like(v) for k, v in state.items()})
        snap.restore({"ddp": out})
        for k, v in state.items():
            assert torch.equal(out[k].reshape(-1).view(torch.uint8), v.reshape(-1).view(torch.uint8)), k
    finally:
        shutil.rmtree(d, ignore_errors=True)


def test_port_of_the_reference_gpu_path_writes_reference_bytes(golden_dir, tmp_path):
    """oracle/ref_path.take_gpu (bench.py's `reference_gpu_path_port` baseline) restates the reference's GPU
    stager with torch calls; its files must hold the reference's bytes too."""
    from oracle import ref_path

    gold = json.load(open(os.path.join(golden_dir, "slab_mixed.json")))
    threshold, items = _cases.slab_case_mixed()
    tensors = {p: _cases.build_view(spec, "cuda") for p, spec in items}
    manifest = ref_path.take_gpu(str(tmp_path / "s"), tensors, threshold)
    for p, _spec in items:
        e = manifest[f"0/{p}"]
        buf = open(os.path.join(str(tmp_path / "s"), e["location"]), "rb").read()
        lo, hi = e["byte_range"]
        assert e["byte_range"] == gold["objects"][p]["byte_range"] and sha(buf[lo:hi]) == gold["objects"][p]["sha256"], p
