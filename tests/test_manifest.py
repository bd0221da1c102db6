"""`.snapshot_metadata` is the reference's document: the schema-aware writer must produce, character for
character, what json.dumps(indent=2) gives (reference manifest.py:283-289), and load what it wrote."""

import json
import os

from torchsnapshot_b200.manifest import (
    ChunkedTensorEntry,
    DictEntry,
    ListEntry,
    ObjectEntry,
    OrderedDictEntry,
    PrimitiveEntry,
    Shard,
    ShardedTensorEntry,
    SnapshotMetadata,
    TensorEntry,
)


def _generic_text(md):
    doc = {"version": md.version, "world_size": md.world_size, "manifest": {k: v.to_obj() for k, v in md.manifest.items()}}
    return json.dumps(doc, sort_keys=False, indent=2)


def test_reference_written_metadata_round_trips_to_identical_text(golden_dir):
    text = open(os.path.join(golden_dir, "c1_simple", "snapshot", ".snapshot_metadata")).read()
    md = SnapshotMetadata.from_yaml(text)
    assert md.to_yaml() == text
    assert md.manifest["0/optim/param_groups/0/lr"].get_value() == 1e-3
    assert md.manifest["0/model"].type == "OrderedDict"


def test_every_entry_kind_matches_the_generic_encoder():
    t = TensorEntry('a/b"cé', "buffer_protocol", "torch.float32", [], False, None)
    manifest = {
        "0/x": t,
        "0/y": TensorEntry("l", "buffer_protocol", "torch.int8", [3, 4], True, [0, 12]),
        "0/s": ShardedTensorEntry(shards=[Shard([0, 0], [2, 2], t), Shard([2, 0], [2, 2], t)]),
        "0/e": ShardedTensorEntry(shards=[]),
        "0/c": ChunkedTensorEntry("torch.float32", [4, 2], [Shard([0, 0], [2, 2], t)], True),
        "0/d": DictEntry(keys=["a", 1]),
        "0/l": ListEntry(),
        "0/o": ObjectEntry("loc", "torch_save", "builtins.tuple", False),
        "0/p": PrimitiveEntry.from_object(2.5),
        "0/b": PrimitiveEntry.from_object(b"\x00\xff"),
        "0/od": OrderedDictEntry(keys=[]),
    }
    md = SnapshotMetadata("0.1.0", 3, manifest)
    assert md.to_yaml() == _generic_text(md)
    back = SnapshotMetadata.from_yaml(md.to_yaml())
    assert back.to_yaml() == md.to_yaml() and back.world_size == 3
    assert back.manifest["0/b"].get_value() == b"\x00\xff" and back.manifest["0/p"].get_value() == 2.5
    empty = SnapshotMetadata("0.1.0", 1, {})
    assert empty.to_yaml() == _generic_text(empty)


def test_legacy_yaml_metadata_still_loads():
    text = "version: 0.0.1\nworld_size: 1\nmanifest:\n  0/foo:\n    type: Tensor\n    location: 0/foo\n    serializer: buffer_protocol\n    dtype: torch.float32\n    shape: [2, 3]\n    replicated: false\n    byte_range: null\n"
    md = SnapshotMetadata.from_yaml(text)
    assert md.manifest["0/foo"].shape == [2, 3] and md.manifest["0/foo"].byte_range is None


# ---- the wire form of a rank's manifest (take-side exchange) ------------------------------------------------
def _wire_manifest():
    t = TensorEntry("batched/u1", "buffer_protocol", "torch.float32", [3, 4], False, [0, 48])
    return {
        "m": DictEntry(keys=["a", "b", 3]),
        "m/a": t,
        "m/b": TensorEntry("0/m/b", "buffer_protocol", "torch.bfloat16", [], True, None),
        "m/3": PrimitiveEntry.from_object(2.5),
        "m/l": ListEntry(),
        "m/o": ObjectEntry(location="0/m/o", serializer="torch_save", obj_type="tuple", replicated=False),
        "m/s": ShardedTensorEntry(shards=[Shard([0], [4], TensorEntry("sharded/s_0", "buffer_protocol", "torch.int64", [4], False, [48, 80]))]),
        "m/c": ChunkedTensorEntry(dtype="torch.float32", shape=[8], chunks=[Shard([0], [8], t)], replicated=True),
        "m/z": TensorEntry("batched/u1", "buffer_protocol", "torch.int8", [5], False, [80, 85]),
    }


def _as_objs(manifest):
    return [(p, e.to_obj()) for p, e in manifest.items()]


def test_encoded_manifest_decodes_to_the_same_entries_in_the_same_order():
    from torchsnapshot_b200.manifest import LazySnapshotMetadata, decode_manifest, encode_manifest

    manifest = _wire_manifest()
    blob = encode_manifest(manifest)
    assert isinstance(blob, bytes)
    assert _as_objs(decode_manifest(blob)) == _as_objs(manifest)
    assert _as_objs(decode_manifest(encode_manifest({}))) == []
    lazy = LazySnapshotMetadata("0.0.3", 2, [blob, manifest])
    assert "manifest" not in lazy.__dict__
    want = [("0/" + p, o) for p, o in _as_objs(manifest)] + [("1/" + p, o) for p, o in _as_objs(manifest)]
    assert _as_objs(lazy.manifest) == want and "manifest" in lazy.__dict__
    assert SnapshotMetadata.from_yaml(lazy.to_yaml()).to_yaml() == lazy.to_yaml()


def test_cached_columns_rest_and_digests():
    """The tensor entries of a cached plan travel as columns encoded once (`tensor_columns_of`); with `rest` only
    the other entries are walked; `manifest_digest` says whether the bytes a rank would send have changed."""
    import pickle

    from torchsnapshot_b200.manifest import decode_manifest, encode_manifest, manifest_digest, tensor_columns_of

    manifest = _wire_manifest()
    cached = {p: e for p, e in manifest.items() if type(e) is TensorEntry}
    rest = {p: e for p, e in manifest.items() if p not in cached}
    columns = tensor_columns_of(cached)
    assert columns[0] is cached and len(columns[2]) == 16
    plain = decode_manifest(encode_manifest(manifest))
    for blob in (encode_manifest(manifest, columns), encode_manifest(manifest, columns, rest)):
        got = decode_manifest(blob)
        assert sorted(_as_objs(got)) == sorted(_as_objs(plain))  # cached entries come last, nothing is lost
        assert list(got)[-len(cached):] == list(cached)
    head, tail = encode_manifest(manifest, columns, rest, parts=True)
    assert tail is columns[1] and pickle.dumps((head, tail), protocol=5) == encode_manifest(manifest, columns, rest)
    # a `rest` that does not add up to the manifest is refused, not silently mis-encoded
    try:
        encode_manifest(manifest, columns, {"m": manifest["m"]})
    except AssertionError:
        pass
    else:
        raise AssertionError("encode_manifest accepted a `rest` that leaves entries out")
    # digests: stable for the same content, different when a primitive, a location or the cached columns change
    d0 = manifest_digest(head, columns)
    assert len(d0) == 32 and d0 == manifest_digest(encode_manifest(manifest, columns, rest, parts=True)[0], columns)
    changed = dict(rest)
    changed["m/3"] = PrimitiveEntry.from_object(3.5)
    head2, _ = encode_manifest({**changed, **cached}, columns, changed, parts=True)
    assert manifest_digest(head2, columns) != d0
    moved = {p: (TensorEntry("batched/u2", e.serializer, e.dtype, e.shape, e.replicated, e.byte_range) if e.location == "batched/u1" else e)
             for p, e in cached.items()}
    columns2 = tensor_columns_of(moved)
    head3, _ = encode_manifest({**rest, **moved}, columns2, rest, parts=True)
    assert head3 == head and manifest_digest(head3, columns2) != d0
    head4, tail4 = encode_manifest(manifest, None, None, parts=True)
    assert tail4 is None and manifest_digest(head4, None) != d0
