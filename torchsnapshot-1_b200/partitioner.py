"""Replicated-write partitioning: spread the write requests of replicated entries over the ranks.

Greedy size-balanced assignment like the reference (torchsnapshot/partitioner.py:42-79): logical
paths in rank-0 order, each to the rank with the fewest bytes so far, seeded with every rank's
non-replicated bytes; the chunks of a chunked tensor are separate units.  Two deliberate
differences: chunks are assigned in offset order (the reference iterates a Python set, so its
chunk->rank map changes from run to run), and every rank computes the identical assignment from
one all_gather_object instead of rank 0 computing and broadcasting it (one collective, not two).
"""

import copy
import os
from collections import defaultdict
from typing import Any, Dict, List, Optional, Sequence, Tuple

from .io_types import WriteReq
from .manifest import ChunkedTensorEntry, Entry, is_replicated


def _write_req_size(wr: WriteReq) -> int:
    stager = wr.buffer_stager
    entry = getattr(stager, "entry", None)
    if entry is not None and hasattr(entry, "dtype"):
        from .io_preparer import TensorIOPreparer

        return TensorIOPreparer.get_tensor_size_from_entry(entry)
    return stager.get_staging_cost_bytes()


def assign_write_loads(
    loads: Sequence[Tuple[str, int, int, bool]],
    rank_to_size: Sequence[int],
    rank_to_speed: Optional[Sequence[float]] = None,
) -> List[List[Tuple[str, int]]]:
    """loads: (logical_path, write_req_idx, size, subpartitionable) in rank-0 order.  Returns, per
    rank, the (logical_path, write_req_idx) pairs it writes.

    `rank_to_speed` (relative device->host throughput of every rank, see `observed_speed`) turns
    "fewest bytes" into "earliest finish": when N GPUs copy to the host at once their links are not
    equally fast (profiles/r01_multi_gpu_peaks.json: 26.5 vs ~38 GB/s at N=4), and the take ends when
    the slowest rank ends.  Without speeds this is the reference's byte balance (partitioner.py:42-79)."""
    sizes = list(rank_to_size)
    out: List[List[Tuple[str, int]]] = [[] for _ in sizes]
    if rank_to_speed is None or len(rank_to_speed) != len(sizes) or any(not s or s <= 0 for s in rank_to_speed):
        speed = [1.0] * len(sizes)
    else:
        mean = sum(rank_to_speed) / len(rank_to_speed)
        # clamp: a noisy estimate must not starve or overload a rank
        speed = [min(max(s / mean, 0.5), 2.0) for s in rank_to_speed]

    def lightest() -> int:
        return min(range(len(sizes)), key=lambda r: (sizes[r] / speed[r], r))

    # whole logical paths first, then the chunks of sub-partitionable paths (reference order)
    grouped: Dict[str, List[Tuple[int, int]]] = defaultdict(list)
    order: List[str] = []
    chunk_units: List[Tuple[str, int, int]] = []
    for path, idx, size, sub in loads:
        if sub:
            chunk_units.append((path, idx, size))
            continue
        if path not in grouped:
            order.append(path)
        grouped[path].append((idx, size))
    for path in order:
        r = lightest()
        for idx, size in grouped[path]:
            out[r].append((path, idx))
            sizes[r] += size
    for path, idx, size in chunk_units:
        r = lightest()
        out[r].append((path, idx))
        sizes[r] += size
    return out


_speed_estimate: Optional[float] = None  # this rank's device->host GB/s under load (EMA over takes)


def observed_speed() -> Optional[float]:
    """This rank's device->host throughput as measured on its previous takes (None before the first)."""
    return _speed_estimate


def record_take_speed(nbytes: int, device_ms: float) -> None:
    """Called by the scheduler when a take has drained: bytes moved / device time of pack + D2H."""
    global _speed_estimate
    if nbytes < (64 << 20) or device_ms <= 0:
        return  # too small to say anything about the link
    gbps = nbytes / device_ms / 1e6
    _speed_estimate = gbps if _speed_estimate is None else 0.5 * _speed_estimate + 0.5 * gbps


def partition_write_reqs(
    entries: Dict[str, Entry], write_reqs: Dict[str, List[WriteReq]], pg: Any
) -> Tuple[Dict[str, Entry], Dict[str, List[WriteReq]]]:
    """Keep, on this rank, only its share of the replicated write requests (reference
    partitioner.py:169-233).  Entries must be consolidated afterwards
    (`consolidate_replicated_entries`)."""
    if not set(write_reqs.keys()).issubset(entries.keys()):
        raise RuntimeError(
            "Not all entries associated with the write reqs are passed in. "
            f"Missing: {set(write_reqs.keys()) - set(entries.keys())}."
        )
    if os.environ.get("TORCH_SNAPSHOT_DISABLE_PARTITIONER") is not None:
        raise NotImplementedError("TORCH_SNAPSHOT_DISABLE_PARTITIONER is not implemented.")
    world = pg.get_world_size()
    rank = pg.get_rank()
    replicated = {k: v for k, v in entries.items() if is_replicated(v)}
    if world == 1 or not replicated:
        return entries, write_reqs
    non_replicated_size = sum(
        _write_req_size(wr) for k, wrs in write_reqs.items() if k not in replicated for wr in wrs
    )
    loads = []
    for path, entry in replicated.items():
        sub = getattr(entry, "type", None) == "ChunkedTensor"
        for idx, wr in enumerate(write_reqs.get(path, [])):
            loads.append((path, idx, _write_req_size(wr), sub))
    gathered: List[Any] = [None] * world
    pg.all_gather_object(gathered, (non_replicated_size, loads, observed_speed()))
    # rank 0's view defines the units, like the reference; a chunked entry is only split when every
    # rank chunked it identically
    base_loads = gathered[0][1]
    same = all(g[1] == base_loads for g in gathered)
    if not same:
        base_loads = [(p, i, s, False) for p, i, s, _sub in base_loads]
    speeds = [g[2] for g in gathered]
    mine = sorted(assign_write_loads(base_loads, [g[0] for g in gathered], speeds if all(speeds) else None)[rank])
    new_entries: Dict[str, Entry] = {k: v for k, v in entries.items() if k not in replicated}
    new_wrs: Dict[str, List[WriteReq]] = {k: v for k, v in write_reqs.items() if k not in replicated}
    for path, idx in mine:
        entry = replicated[path]
        if getattr(entry, "type", None) == "ChunkedTensor":
            if path not in new_entries:
                part = copy.copy(entry)
                part.chunks = []
                new_entries[path] = part
            new_entries[path].chunks.append(entry.chunks[idx])
        else:
            new_entries[path] = entry
        new_wrs.setdefault(path, []).append(write_reqs[path][idx])
    return new_entries, new_wrs


def consolidate_replicated_entries(rank_to_entries: List[Dict[str, Entry]], dedup: bool = True) -> List[Dict[str, Entry]]:
    """Merge the partial views of replicated entries back together and list them once, under rank 0
    (reference partitioner.py:236-303)."""
    chunk_groups: Dict[str, List[ChunkedTensorEntry]] = defaultdict(list)
    for entries in rank_to_entries:
        for path, entry in entries.items():
            if is_replicated(entry) and getattr(entry, "type", None) == "ChunkedTensor":
                chunk_groups[path].append(entry)
    merged_chunked = {}
    for path, group in chunk_groups.items():
        seen = {}
        for e in group:
            for c in e.chunks:
                seen[tuple(c.offsets)] = c
        merged_chunked[path] = ChunkedTensorEntry(
            dtype=group[0].dtype, shape=group[0].shape,
            chunks=[seen[k] for k in sorted(seen)], replicated=True,
        )
    replicated: Dict[str, Entry] = {}
    for entries in rank_to_entries:
        mine = [path for path, entry in entries.items() if getattr(entry, "replicated", False)]
        for path in mine:
            if path not in replicated:
                replicated[path] = merged_chunked.get(path, entries[path])
            del entries[path]
    for rank, entries in enumerate(rank_to_entries):
        if dedup and rank != 0:
            continue
        entries.update(replicated)
    return rank_to_entries


def consolidate_replicated_entries_dist(entries: Dict[str, Entry], pg: Any, dedup: bool = True) -> Dict[str, Entry]:
    """Collective form of `consolidate_replicated_entries` (reference partitioner.py:306-316): gather every
    rank's entries, consolidate, keep this rank's."""
    gathered: List[Any] = [None] * pg.get_world_size()
    pg.all_gather_object(obj_list=gathered, obj=entries)
    return consolidate_replicated_entries(rank_to_entries=gathered, dedup=dedup)[pg.get_rank()]
