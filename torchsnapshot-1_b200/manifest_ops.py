"""Per-rank views of the global manifest (reference torchsnapshot/manifest_ops.py:24-222).

The global manifest keys every entry by "<rank>/<logical path>".  A restoring rank sees its own entries,
the replicated ones (listed once, under rank 0), and for every sharded tensor the shards of ALL ranks.
Unlike the reference nothing is deep-copied (that copy is 0.3 s at 5 k entries): entries are
shared with the metadata object and never mutated here -- a container whose key list has to change is
replaced by a modified copy.
"""

import dataclasses
from collections import defaultdict
from typing import Any, Dict, List, Tuple

from .manifest import (
    is_container_entry,
    is_dict_entry,
    is_replicated,
    Manifest,
    ShardedTensorEntry,
    SnapshotMetadata,
)


def _split_by_rank(metadata: SnapshotMetadata) -> List[Manifest]:
    per_rank: List[Manifest] = [{} for _ in range(metadata.world_size)]
    for path, entry in metadata.manifest.items():
        rank, _, logical_path = path.partition("/")
        per_rank[int(rank)][logical_path] = entry
    return per_rank


def _merge_sharded_entries(per_rank: List[Manifest]) -> Dict[str, ShardedTensorEntry]:
    shards_by_path: Dict[str, List[Any]] = defaultdict(list)
    for manifest in per_rank:
        for logical_path, entry in manifest.items():
            if isinstance(entry, ShardedTensorEntry):
                shards_by_path[logical_path] += entry.shards
    return {
        path: ShardedTensorEntry(shards=sorted(shards, key=lambda s: s.offsets))
        for path, shards in shards_by_path.items()
    }


def _with_key_change(manifest: Manifest, parent_path: str, key: str, add: bool) -> None:
    """Replace the container at `parent_path` by a copy whose key list gained / lost `key`."""
    parent = manifest.get(parent_path)
    if parent is None or not is_dict_entry(parent):
        return
    keys = list(parent.keys)
    if add:
        if key not in keys and not (key.lstrip("-").isdigit() and int(key) in keys):
            keys.append(key)
    elif key in keys:
        keys.remove(key)
    elif key.lstrip("-").isdigit() and int(key) in keys:  # dict keys may be ints (manifest_ops.py:212-216)
        keys.remove(int(key))
    else:
        raise ValueError(f"{key!r} is not a key of the container at {parent_path!r}")
    manifest[parent_path] = dataclasses.replace(parent, keys=keys)


def _remove_entry(manifest: Manifest, logical_path: str) -> None:
    """Drop an entry and its key in the parent container (manifest_ops.py:179-216)."""
    if logical_path not in manifest:
        return
    del manifest[logical_path]
    parent_path, _, key = logical_path.rpartition("/")
    if parent_path:
        _with_key_change(manifest, parent_path, key, add=False)


def get_manifest_for_rank(metadata: SnapshotMetadata, rank: int) -> Tuple[Manifest, Dict[str, ShardedTensorEntry]]:
    """(local manifest of `rank`, sharded entries merged over all ranks) -- manifest_ops.py:24-84.

    A rank that existed at save time: its own entries, plus rank 0's replicated entries, with every
    sharded entry replaced by the merged one.  A rank beyond the saved world size: rank 0's containers and
    replicated entries only (the other entries are removed together with their keys in the parents)."""
    if metadata.world_size == 1 and rank == 0:
        # one saved rank, restored by rank 0 (the common single-process case, and what a 13 k-tensor restore pays
        # for planning): one pass, no per-rank split
        local = {}
        sharded = False
        for path, entry in metadata.manifest.items():
            local[path[2:]] = entry  # strip "0/"
            sharded = sharded or isinstance(entry, ShardedTensorEntry)
        if not sharded:
            return local, {}
    per_rank = _split_by_rank(metadata)
    merged = _merge_sharded_entries(per_rank)
    if rank < metadata.world_size:
        local = dict(per_rank[rank])
        for logical_path, entry in per_rank[0].items():
            if is_replicated(entry):
                local[logical_path] = entry
        for logical_path, entry in list(local.items()):
            if isinstance(entry, ShardedTensorEntry):
                local[logical_path] = merged[logical_path]
        return local, merged
    local = dict(per_rank[0])
    for logical_path in list(local.keys()):
        entry = local[logical_path]
        if not (is_container_entry(entry) or is_replicated(entry)):
            _remove_entry(local, logical_path)
    return local, merged


def handle_sharded_tensor_elasticity(
    manifest: Manifest, merged_sd_entries: Dict[str, ShardedTensorEntry], tensor_requests: List[str]
) -> None:
    """Make the presence of sharded tensors follow what the restoring state dict asks for
    (manifest_ops.py:118-176): a requested sharded tensor this rank did not save is added (and keyed in its
    parent container); a saved one that is not requested is dropped.  Like the reference, only when every
    sharded tensor sits at the root of its stateful's state dict ("<stateful>/<name>")."""
    if not all(len(path.split("/")) == 2 for path in merged_sd_entries):
        return
    wanted = [path for path in tensor_requests if path in merged_sd_entries]
    for logical_path in wanted:
        if logical_path not in manifest:
            manifest[logical_path] = merged_sd_entries[logical_path]
            parent_path, _, key = logical_path.rpartition("/")
            _with_key_change(manifest, parent_path, key, add=True)
    wanted_set = set(wanted)
    for logical_path in list(manifest.keys()):
        if isinstance(manifest[logical_path], ShardedTensorEntry) and logical_path not in wanted_set:
            del manifest[logical_path]
