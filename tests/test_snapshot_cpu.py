"""Snapshot.take / restore / read_object / async_take on CPU tensors (host path) and drop-in
compatibility with a snapshot the reference wrote (tests/golden/c1_simple, BASELINE config 0)."""

import hashlib
import json
import os
import re
from collections import OrderedDict

import pytest
import torch

import oracle
from torchsnapshot_b200 import knobs
from torchsnapshot_b200.manifest import SnapshotMetadata
from torchsnapshot_b200.rng_state import RNGState
from torchsnapshot_b200.snapshot import Snapshot, StateDict


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def _c1_app_state(seed=1234, step=True, inputs=None):
    """simple_example.py-style app state.  With `inputs` (tests/golden/c1_simple/inputs.pt) the tensors
    are the exact ones the reference saved -- recomputing them is not bit-reproducible across CPUs
    (BLAS reduction order), and the comparison below is on bytes."""
    torch.manual_seed(seed)
    model = torch.nn.Sequential(torch.nn.Linear(128, 64), torch.nn.Linear(64, 32), torch.nn.Linear(32, 1))
    optim = torch.optim.Adam(model.parameters(), lr=1e-3)
    if step:
        model(torch.randn(16, 128)).sum().backward()
        optim.step()
    if inputs is not None:
        tensors = inputs["tensors"]
        model.load_state_dict({k[len("model/"):]: v.clone() for k, v in tensors.items() if k.startswith("model/")})
        for i, p in enumerate(model.parameters()):
            for key in list(optim.state[p].keys()):
                optim.state[p][key] = tensors[f"optim/state/{i}/{key}"].clone()
    return {"model": model, "optim": optim, "progress": StateDict(current_epoch=0)}


def _c1_inputs(golden_dir):
    return torch.load(os.path.join(golden_dir, "c1_simple", "inputs.pt"), weights_only=False)


def _payloads(root):
    """{manifest path: sha256 of payload bytes} for every tensor entry of a snapshot directory."""
    meta = json.load(open(os.path.join(root, ".snapshot_metadata")))
    out = {}
    for path, e in meta["manifest"].items():
        if e["type"] != "Tensor":
            continue
        with open(os.path.join(root, e["location"]), "rb") as f:
            buf = f.read()
        br = e["byte_range"]
        out[path] = sha(buf if br is None else buf[br[0] : br[1]])
    return out, meta


def _normalise(meta):
    """Manifest with slab uuids factored out (slab names are random, reference batcher.py:172)."""
    text = json.dumps(meta, sort_keys=True)
    names = []
    for m in re.finditer(r"batched/[0-9a-f-]{36}", text):
        if m.group(0) not in names:
            names.append(m.group(0))
    for i, n in enumerate(names):
        text = text.replace(n, f"batched/#{i}")
    return json.loads(text)


def test_take_writes_what_the_reference_wrote(golden_dir, tmp_path):
    """Same app state -> same manifest (modulo slab uuids) and the same payload bytes per object."""
    ref_root = os.path.join(golden_dir, "c1_simple", "snapshot")
    ours = str(tmp_path / "snap")
    Snapshot.take(ours, _c1_app_state(inputs=_c1_inputs(golden_dir)))
    ref_payloads, ref_meta = _payloads(ref_root)
    our_payloads, our_meta = _payloads(ours)
    assert our_payloads == ref_payloads
    assert _normalise(our_meta) == _normalise(ref_meta)
    # the object files (pickled tuples etc.) are there as well
    for path, e in ref_meta["manifest"].items():
        if e["type"] == "object":
            assert os.path.exists(os.path.join(ours, e["location"]))


def test_restore_from_a_reference_written_snapshot(golden_dir):
    ref_root = os.path.join(golden_dir, "c1_simple", "snapshot")
    inputs = torch.load(os.path.join(golden_dir, "c1_simple", "inputs.pt"), weights_only=False)
    app_state = _c1_app_state(seed=99)  # different weights, same structure
    Snapshot(ref_root).restore(app_state)
    sd = app_state["model"].state_dict()
    for name, t in sd.items():
        assert oracle.serialize(t) == oracle.serialize(inputs["tensors"][f"model/{name}"])
    osd = app_state["optim"].state_dict()
    assert oracle.serialize(osd["state"][0]["exp_avg"]) == oracle.serialize(inputs["tensors"]["optim/state/0/exp_avg"])
    assert osd["param_groups"][0]["betas"] == (0.9, 0.999) and osd["param_groups"][0]["lr"] == 1e-3
    assert app_state["progress"]["current_epoch"] == 0
    # random access
    snap = Snapshot(ref_root)
    w = snap.read_object("0/model/0.weight")
    assert oracle.serialize(w) == oracle.serialize(inputs["tensors"]["model/0.weight"])
    assert snap.read_object("0/progress/current_epoch") == 0
    out = torch.zeros(64, 128)
    assert snap.read_object("0/model/0.weight", obj_out=out, memory_budget_bytes=4096) is not None
    assert oracle.serialize(out) == oracle.serialize(inputs["tensors"]["model/0.weight"])
    with pytest.raises(RuntimeError):
        snap.read_object("0/nope")


@pytest.mark.parametrize("batching", [True, False])
def test_round_trip_with_odd_structures(tmp_path, batching):
    state = StateDict(
        {
            "a/b": torch.arange(10, dtype=torch.int16),
            "%weird": [torch.randn(3, 3), {"x": torch.randn(2).bfloat16(), 7: torch.tensor(True)}],
            "od": OrderedDict(z=1, y=2.5, s="str", b=b"bytes", t=(1, 2)),
            "c": torch.randn(4, 4, dtype=torch.complex64),
            "nc": torch.randn(6, 8)[:, ::2],
            "scalar": torch.tensor(3.5, dtype=torch.float64),
            "collide": {1: "a", "1": "b"},
        }
    )
    path = str(tmp_path / "s")
    with knobs.override_is_batching_disabled(not batching):
        snap = Snapshot.take(path, {"st": state, "rng": RNGState()})
    rng_after_take = torch.get_rng_state()
    meta = SnapshotMetadata.from_yaml(open(os.path.join(path, ".snapshot_metadata")).read())
    assert meta.manifest["0/st/od/y"].type == "float" and meta.manifest["0/st/od/y"].get_value() == 2.5
    assert meta.manifest["0/st/collide"].type == "object"
    if batching:
        assert meta.manifest["0/st/a%2Fb"].location.startswith("batched/")
    else:
        assert meta.manifest["0/st/a%2Fb"].location == "0/st/a%2Fb"
    out = StateDict(
        {
            "a/b": torch.zeros(10, dtype=torch.int16),
            "%weird": [torch.zeros(3, 3), {"x": torch.zeros(2).bfloat16(), 7: torch.tensor(False)}],
            "od": OrderedDict(),
            "c": torch.zeros(4, 4, dtype=torch.complex64),
            "nc": torch.zeros(6, 4),
            "scalar": torch.tensor(0.0, dtype=torch.float64),
            "collide": None,
        }
    )
    torch.manual_seed(1)
    snap.restore({"st": out, "rng": RNGState()})
    assert torch.equal(torch.get_rng_state(), rng_after_take)
    assert torch.equal(out["a/b"], state["a/b"]) and torch.equal(out["%weird"][0], state["%weird"][0])
    assert torch.equal(out["%weird"][1]["x"], state["%weird"][1]["x"]) and bool(out["%weird"][1][7]) is True
    assert out["od"] == state["od"] and isinstance(out["od"], OrderedDict)
    assert torch.equal(out["c"], state["c"]) and torch.equal(out["nc"], state["nc"].contiguous())
    assert out["scalar"].item() == 3.5 and out["collide"] == {1: "a", "1": "b"}


def test_large_tensor_is_chunked_and_restored(tmp_path):
    t = torch.randn(64, 33)
    with knobs.override_max_chunk_size_bytes(1000), knobs.override_slab_size_threshold_bytes(600):
        snap = Snapshot.take(str(tmp_path / "s"), {"st": StateDict(t=t)})
    e = snap.get_manifest()["0/st/t"]
    # ceil(8448 / 1000) = 9 requested chunks -> 8 rows each -> torch.chunk semantics give 8 chunks
    assert e.type == "ChunkedTensor" and len(e.chunks) == 8
    assert [c.tensor.location for c in e.chunks] == [f"0/st/t_{8 * i}_0" for i in range(8)]
    out = StateDict(t=torch.zeros(64, 33))
    snap.restore({"st": out})
    assert torch.equal(out["t"], t)


def test_async_take_commits_metadata_last_and_surfaces_errors(tmp_path, monkeypatch):
    from torchsnapshot_b200.storage_plugins import fs as fsmod

    state = StateDict(t=torch.randn(100, 100), u=torch.arange(7))
    pending = Snapshot.async_take(str(tmp_path / "ok"), {"st": state})
    state["t"].zero_()  # mutation after async_take returns must not leak into the snapshot
    snap = pending.wait()
    assert pending.done()
    out = StateDict(t=torch.zeros(100, 100), u=torch.zeros(7, dtype=torch.int64))
    snap.restore({"st": out})
    assert out["t"].abs().sum() > 0 and torch.equal(out["u"], torch.arange(7))

    class Faulty(fsmod.FSStoragePlugin):
        async def write(self, write_io):
            raise OSError("disk on fire")

    import torchsnapshot_b200.storage_plugin as sp

    monkeypatch.setattr(sp, "FSStoragePlugin", Faulty)
    pending = Snapshot.async_take(str(tmp_path / "bad"), {"st": StateDict(t=torch.randn(10))})
    with pytest.raises(RuntimeError, match="disk on fire"):  # the reference wraps it (snapshot.py:935-941)
        pending.wait()
    assert not os.path.exists(str(tmp_path / "bad" / ".snapshot_metadata"))


def test_edge_cases_the_reference_cannot_stage(tmp_path):
    """Empty tensors and bf16 tensors with an odd element count (the reference drops bytes or
    raises there: SURVEY 8a quirk 1, serialization.py:188)."""
    state = StateDict(empty=torch.zeros(0, 3), bf_odd=torch.randn(5).bfloat16(), bf_scalar=torch.tensor(1.5).bfloat16(),
                      after=torch.arange(3, dtype=torch.int32))
    snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
    m = snap.get_manifest()
    assert m["0/st/bf_odd"].byte_range[1] - m["0/st/bf_odd"].byte_range[0] == 10
    out = StateDict(empty=torch.ones(0, 3), bf_odd=torch.zeros(5).bfloat16(), bf_scalar=torch.tensor(0.0).bfloat16(),
                    after=torch.zeros(3, dtype=torch.int32))
    snap.restore({"st": out})
    assert torch.equal(out["bf_odd"], state["bf_odd"]) and out["bf_scalar"].item() == 1.5
    assert torch.equal(out["after"], state["after"]) and out["empty"].shape == (0, 3)


def test_custom_tensor_prepare_func_and_pickled_dtypes(tmp_path):
    """`_custom_tensor_prepare_func` (reference snapshot.py:175-179, tensor.py:55-71): the prepared tensor is
    what gets written, tensors with a prepare func are never batched; complex tensors go through torch.save."""
    calls = []

    def halve(path, tensor, tracing):
        calls.append((path, tracing))
        return tensor.to(torch.float16) if tensor.dtype == torch.float32 else tensor

    state = StateDict(w=torch.arange(8, dtype=torch.float32), z=torch.ones(3, dtype=torch.complex64), k=torch.arange(4))
    snap = Snapshot.take(str(tmp_path / "s"), {"st": state}, _custom_tensor_prepare_func=halve)
    m = snap.get_manifest()
    assert m["0/st/w"].dtype == "torch.float16" and m["0/st/w"].location == "0/st/w" and m["0/st/w"].byte_range is None
    assert m["0/st/z"].serializer == "torch_save" and m["0/st/k"].location == "0/st/k"
    assert ("st/w", True) in calls and ("st/w", False) in calls
    assert os.path.getsize(os.path.join(str(tmp_path / "s"), "0/st/w")) == 16
    out = StateDict(w=torch.zeros(8, dtype=torch.float16), z=torch.zeros(3, dtype=torch.complex64), k=torch.zeros(4, dtype=torch.int64))
    snap.restore({"st": out})
    assert torch.equal(out["w"], torch.arange(8, dtype=torch.float16)) and torch.equal(out["z"], state["z"])
    # dtype mismatch on load: the entry's dtype wins and a new tensor is returned through load_state_dict
    out2 = StateDict(w=torch.zeros(8, dtype=torch.float32), z=torch.zeros(3, dtype=torch.complex64), k=torch.zeros(4, dtype=torch.int64))
    snap.restore({"st": out2})
    assert out2["w"].dtype == torch.float16


def test_pickled_dtypes_round_trip(tmp_path):
    """Dtypes outside the buffer-protocol list (complex, quantized) are pickled with torch.save like in the
    reference (tensor.py:68-71) and are never batched into slabs."""
    q = torch.quantize_per_tensor(torch.randn(6, 5), scale=0.1, zero_point=3, dtype=torch.qint8)
    state = StateDict(q=q, z=torch.randn(4, dtype=torch.complex128), f=torch.randn(3))
    snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
    m = snap.get_manifest()
    assert m["0/st/q"].serializer == "torch_save" and m["0/st/q"].dtype == "torch.qint8" and m["0/st/q"].byte_range is None
    assert m["0/st/z"].serializer == "torch_save" and m["0/st/f"].location.startswith("batched/")
    out = StateDict(q=torch.quantize_per_tensor(torch.zeros(6, 5), scale=0.1, zero_point=3, dtype=torch.qint8),
                    z=torch.zeros(4, dtype=torch.complex128), f=torch.zeros(3))
    snap.restore({"st": out})
    assert torch.equal(out["q"].int_repr(), q.int_repr()) and torch.equal(out["z"], state["z"])
    got = snap.read_object("0/st/z")
    assert torch.equal(got, state["z"])
    with pytest.raises(RuntimeError, match="quantized"):
        snap.read_object("0/st/q")  # no runtime object to learn the quantization parameters from


def test_host_slab_buffers_are_pooled_and_never_shared_while_viewed():
    """CPU slabs and host reads use pooled buffers (batcher.HostSlab): a buffer goes back to the free list
    only when the last memoryview of it is gone, and is then handed out again."""
    import gc

    from torchsnapshot_b200.batcher import HostSlab

    n = (3 << 20) + 17
    a = HostSlab(n)
    view = memoryview(a)
    assert len(view) == n and not view.readonly
    ptr = a.ptr
    view[:4] = b"abcd"
    part = view[1:3]
    del a, view
    gc.collect()
    b = HostSlab(n)  # `part` still pins the first buffer: this must be another one
    assert b.ptr != ptr and bytes(part) == b"bc"
    del part
    gc.collect()
    c = HostSlab(n)
    assert c.ptr == ptr  # the released buffer is reused
    del b, c


def test_cpu_slab_fast_path_writes_the_reference_bytes(tmp_path):
    """Members of every kind a CPU slab can hold -- dense, 0-d, strided views, bf16, bool -- through the native
    gather (BatchedBufferStager) and the native scatter (BatchedBufferConsumer), against the oracle."""
    import oracle

    base = torch.arange(600, dtype=torch.float32).reshape(20, 30)
    state = StateDict(
        dense=torch.randn(1000), scalar=torch.tensor(3.5), strided=base[:, ::3], transposed=base.t(),
        bf=torch.randn(64).bfloat16(), flags=torch.tensor([True, False, True]), i64=torch.arange(7),
        inner=base[2:9, 4:11],
    )
    snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
    meta = json.load(open(os.path.join(snap.path, ".snapshot_metadata")))["manifest"]
    files = {meta[f"0/st/{k}"]["location"] for k in state.keys()}
    assert len(files) == 1  # one CPU slab
    blob = open(os.path.join(snap.path, files.pop()), "rb").read()
    for k, t in state.items():
        lo, hi = meta[f"0/st/{k}"]["byte_range"]
        assert blob[lo:hi] == oracle.serialize(t), k
    # restore: dense destinations take the native scatter, the strided ones the per-member path
    out = StateDict(
        dense=torch.zeros(1000), scalar=torch.tensor(0.0), strided=torch.zeros(20, 30)[:, ::3],
        transposed=torch.zeros(30, 20).t(), bf=torch.zeros(64).bfloat16(), flags=torch.zeros(3, dtype=torch.bool),
        i64=torch.zeros(7, dtype=torch.int64), inner=torch.zeros(20, 30)[2:9, 4:11],
    )
    snap.restore({"st": out})
    for k, t in state.items():
        assert torch.equal(out[k], t), k


def test_large_cpu_tensors_are_read_straight_into_their_memory(tmp_path, monkeypatch):
    """Tensors above the slab threshold are files of their own; a dense CPU destination of the saved dtype is
    filled by the plugin's read_into directly (no staging buffer), chunked and tiled reads included."""
    from torchsnapshot_b200 import knobs
    from torchsnapshot_b200.storage_plugins import fs as fsmod

    calls = []
    orig = fsmod.FSStoragePlugin.read_into

    async def spy(self, path, byte_range, out):
        calls.append((path, byte_range, len(out)))
        await orig(self, path, byte_range, out)

    monkeypatch.setattr(fsmod.FSStoragePlugin, "read_into", spy)
    big = torch.randn(300, 1000)
    bf = torch.randn(70000).bfloat16()
    state = StateDict(big=big, bf=bf, small=torch.arange(5.0))
    with knobs.override_slab_size_threshold_bytes(100_000), knobs.override_max_chunk_size_bytes(500_000):
        snap = Snapshot.take(str(tmp_path / "s"), {"st": state})
        out = StateDict(big=torch.zeros(300, 1000), bf=torch.zeros(70000).bfloat16(), small=torch.zeros(5))
        snap.restore({"st": out})
    assert torch.equal(out["big"], big) and torch.equal(out["bf"], bf) and torch.equal(out["small"], state["small"])
    sizes = sorted(n for _p, _r, n in calls)
    assert 140000 in sizes  # the bf16 tensor, one read of exactly its bytes, into its own memory
    assert sum(sizes) >= big.numel() * 4 + 140000  # the chunks of `big` arrived the same way
    # random access with a memory budget: tiles are read into the tensor's own memory too
    calls.clear()
    got = snap.read_object("0/st/bf", obj_out=torch.zeros(70000).bfloat16(), memory_budget_bytes=30000)
    assert torch.equal(got, bf) and len(calls) >= 5 and max(n for _p, _r, n in calls) <= 30000


def test_restore_into_mixed_destinations(tmp_path):
    """Destinations that match, are strided, have another dtype, are absent or a Tensor subclass: the saved
    values come back, with batching on and off (tests/_mixed_restore.py; the GPU suite runs it on CUDA tensors)."""
    import _mixed_restore

    _mixed_restore.run("cpu", tmp_path)
