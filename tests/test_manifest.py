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
