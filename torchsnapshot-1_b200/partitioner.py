"""Replicated-write partitioning: spread the write requests of replicated entries over the ranks.

Greedy size-balanced assignment like the reference (torchsnapshot/partitioner.py:42-79): logical
paths in rank-0 order, each to the rank with the fewest bytes so far, seeded with every rank's
non-replicated bytes; the chunks of a chunked tensor are separate units.  Two deliberate
differences: chunks are assigned in offset order (the reference iterates a Python set, so its
chunk->rank map changes from run to run), and every rank computes the identical assignment from
one all_gather_object instead of rank 0 computing and broadcasting it (one collective, not two).
"""

import copy
import math
import os
from collections import defaultdict
from typing import Any, Dict, List, Optional, Sequence, Tuple

from .io_types import WriteReq
from .manifest import ChunkedTensorEntry, Entry, is_replicated


_ELEMENT_SIZES: Dict[str, int] = {}


def _write_req_size(wr: WriteReq) -> int:
    stager = wr.buffer_stager
    entry = getattr(stager, "entry", None)
    if entry is not None and hasattr(entry, "dtype"):
        # (this runs once per replicated tensor on the blocked path of every take)
        es = _ELEMENT_SIZES.get(entry.dtype)
        if es is None:
            from .serialization import dtype_to_element_size, string_to_dtype

            es = _ELEMENT_SIZES[entry.dtype] = dtype_to_element_size(string_to_dtype(entry.dtype))
        return es * math.prod(entry.shape)
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
    weighted = not (rank_to_speed is None or len(rank_to_speed) != len(sizes) or any(not s or s <= 0 for s in rank_to_speed))
    if not weighted:
        speed = [1.0] * len(sizes)
    else:
        mean = sum(rank_to_speed) / len(rank_to_speed)
        # clamp: a noisy estimate must not starve or overload a rank.  The range follows what the boxes show:
        # with 8 GPUs copying at once the slowest link ran at a quarter of the fastest (11.6 vs ~45 GB/s,
        # profiles/r01_peaks_n8.txt)
        speed = [min(max(s / mean, 0.25), 4.0) for s in rank_to_speed]

    # finish time of every rank so far; the lightest rank is the first minimum (ties go to the lower rank)
    finish = [sizes[r] / speed[r] for r in range(len(sizes))]
    cost = [1.0 / x for x in speed]
    # whole logical paths first, then the chunks of sub-partitionable paths (reference order)
    grouped: Dict[str, List[Tuple[int, int]]] = {}
    chunk_units: List[Tuple[str, int, int]] = []
    for path, idx, size, sub in loads:
        if sub:
            chunk_units.append((path, idx, size))
        elif path in grouped:
            grouped[path].append((idx, size))
        else:
            grouped[path] = [(idx, size)]
    units = list(grouped.items())
    if weighted:
        # largest first: in dict order the three 160 MB embedding tensors of a GPT-2-XL state come last and
        # leave the ranks up to 8 % apart; sorted, the split lands within one small tensor of the optimum.
        # (Without speeds the order stays the reference's, which the golden partition fixture pins.)
        units.sort(key=lambda kv: -sum(size for _idx, size in kv[1]))
        chunk_units.sort(key=lambda u: -u[2])
    for path, members in units:
        r = finish.index(min(finish))
        mine = out[r]
        for idx, size in members:
            mine.append((path, idx))
            finish[r] += size * cost[r]
    for path, idx, size in chunk_units:
        r = finish.index(min(finish))
        out[r].append((path, idx))
        finish[r] += size * cost[r]
    return out


# This rank's throughput on its previous takes, GB/s: the last few measurements per KIND OF SINK (the class name
# of the storage plugin the take wrote to); the estimate is their median, so that one slow take (another job on
# the shared uplink, a page-cache hiccup) does not move the next split -- an EMA did: one 126 ms take among 78 ms
# ones cost the rank an eighth of its share for the next three takes.  What bounds a rank differs with the sink
# -- its host link when the sink is faster than the link, the sink's own write path otherwise -- and so does
# the best split.  Key "" holds the probed seed.
_SPEED_SAMPLES = 5
_speed_estimates: Dict[str, List[float]] = {}
_speed_context: str = ""


def _median(samples: Sequence[float]) -> float:
    ordered = sorted(samples)
    mid = len(ordered) // 2
    return ordered[mid] if len(ordered) % 2 else 0.5 * (ordered[mid - 1] + ordered[mid])


def set_speed_context(kind: str) -> None:
    """Name the kind of sink of the take being planned (snapshot.py: the storage plugin's class name)."""
    global _speed_context
    _speed_context = kind


def speed_context() -> str:
    return _speed_context


def observed_speed() -> Optional[float]:
    """This rank's throughput as measured on its previous takes into this kind of sink; before the first,
    the probed device->host rate (None if that could not be measured either)."""
    got = _speed_estimates.get(_speed_context) or _speed_estimates.get("")
    return _median(got) if got else None


def probe_speed(device: Optional[int] = None, nbytes: int = 256 << 20) -> Optional[float]:
    """This rank's pinned device->host rate right now, from two copies of `nbytes` (CUDA events): the seed of
    the speed estimate before the first take has been measured.  Every rank of a job reaches its first
    partitioning at about the same time (right after a collective), so the probes overlap the way the copies of
    the take will.  None when there is no device or no room for the probe's buffers."""
    try:
        import torch

        if not torch.cuda.is_available():
            return None
        from ._native import STREAM_D2H
        from .engine import get_engine

        eng = get_engine(torch.cuda.current_device() if device is None else device)
        ctx = eng.ctx
        h = ctx.host_alloc(nbytes)
        try:
            d = ctx.dev_alloc(nbytes)
            try:
                ctx.memcpy_d2h(STREAM_D2H, h, d, nbytes)  # first touch of the path
                a = ctx.event_record(STREAM_D2H)
                ctx.memcpy_d2h(STREAM_D2H, h, d, nbytes)
                ctx.memcpy_d2h(STREAM_D2H, h, d, nbytes)
                b = ctx.event_record(STREAM_D2H)
                ctx.event_sync(b)
                ms = ctx.event_elapsed_ms(a, b)
                ctx.event_release(a)
                ctx.event_release(b)
            finally:
                ctx.dev_free(d)
        finally:
            ctx.host_free(h)
        return 2 * nbytes / ms / 1e6 if ms > 0 else None
    except Exception:  # noqa: BLE001 - a seed is optional
        return None


def record_take_speed(nbytes: int, ms: float, kind: Optional[str] = None) -> None:
    """Called by the scheduler when a take has drained: this rank's payload bytes / the time from its first
    pack launch to its last byte being handed to (and accepted by) the sink."""
    if nbytes < (64 << 20) or ms <= 0:
        return  # too small to say anything
    kind = _speed_context if kind is None else kind
    gbps = nbytes / ms / 1e6
    samples = _speed_estimates.setdefault(kind, [])
    samples.append(gbps)
    del samples[:-_SPEED_SAMPLES]


class RanksDisagree(RuntimeError):
    """Raised by partition_write_reqs on EVERY rank alike when the `agree_on` digests of the ranks differ."""


def _digest(parts: Any) -> str:
    """Digest of a list of strings / tuples of strings and ints that is the same in every process (`hash()` of a
    str is not)."""
    import hashlib

    return hashlib.blake2b(repr(parts).encode("utf-8"), digest_size=12).hexdigest()


class Negotiation:
    """What the ranks told each other in the partitioner's exchange: per rank (non-replicated bytes, digest of
    its write loads, measured throughput, the caller's `agree_on` string, whether it still holds the plan it made
    for these loads last time)."""

    def __init__(self, gathered: List[Any], rank: int) -> None:
        self.gathered = gathered
        self.rank = rank

    @property
    def same_loads(self) -> bool:
        return all(g[1] == self.gathered[0][1] for g in self.gathered)

    @property
    def everybody_can_repeat(self) -> bool:
        return self.same_loads and all(g[4] for g in self.gathered)

    def key(self) -> Any:
        return (self.gathered[0][1], tuple(g[0] for g in self.gathered), len(self.gathered))


# The assignment of the previous take with the same loads: {"key", "out" (per rank: [(path, idx)]), "bytes" (per
# rank), "speeds" (what the ranks reported then), "sizes", "loads", "judged_speeds"}.  Kept while a fresh split
# for the throughputs reported now would not finish noticeably earlier (previous_assignment_stands), so that a
# job checkpointing the same state again and again writes the same files from the same ranks -- and every rank
# can reuse its whole plan (snapshot.py) instead of planning a slightly different share each time.
_last_assignment: Optional[Dict[str, Any]] = None
_JUDGE_AGAIN_DRIFT = 0.03
_REBALANCE_GAIN = 1.04


def negotiate(pg: Any, non_replicated_size: int, loads_digest: str, agree_on: Optional[str],
              can_repeat: bool = False) -> Negotiation:
    """The partitioner's exchange on its own (one small all-gather); raises RanksDisagree on every rank alike
    when the `agree_on` strings differ."""
    world = pg.get_world_size()
    gathered: List[Any] = [None] * world
    held = _last_assignment is not None and _last_assignment["digest"] == loads_digest
    pg.all_gather_object(gathered, (non_replicated_size, loads_digest, observed_speed(), agree_on, bool(can_repeat and held)))
    if any(g[3] != gathered[0][3] for g in gathered):
        raise RanksDisagree("the ranks do not agree on " + repr([g[3] for g in gathered]))
    return Negotiation(gathered, pg.get_rank())


def previous_assignment_stands(neg: Negotiation) -> bool:
    """True when every rank still holds the previous plan for these loads and the previous assignment is still
    good enough for the throughputs reported now (the same answer on every rank: gathered data only).

    "Good enough": while no rank's share of the box's throughput has moved by more than 3 % since the split was
    last judged, it stands unjudged.  Otherwise the slowest rank's predicted time under the standing split is
    compared with what a fresh split for today's throughputs would give: it stands unless the fresh one is more
    than 4 % shorter (whatever the granularity of the loads allows: a split that cannot be improved is never
    given up)."""
    if not neg.everybody_can_repeat or _last_assignment is None or _last_assignment["key"] != neg.key():
        return False
    speeds = [g[2] for g in neg.gathered]
    then = _last_assignment["speeds"]
    if not all(speeds) or not all(then):
        return all(speeds) == all(then)
    judged = _last_assignment.get("judged_speeds") or then
    mean_now, mean_judged = sum(speeds) / len(speeds), sum(judged) / len(judged)
    drift = max(abs((a / mean_now) / (b / mean_judged) - 1.0) for a, b in zip(speeds, judged))
    if drift <= _JUDGE_AGAIN_DRIFT:
        return True
    nonrep = [g[0] for g in neg.gathered]
    sizes = _last_assignment["sizes"]
    standing = max((nonrep[r] + _last_assignment["bytes"][r]) / speeds[r] for r in range(len(speeds)))
    fresh_out = assign_write_loads(_last_assignment["loads"], nonrep, speeds)
    fresh = max((nonrep[r] + sum(sizes[u] for u in fresh_out[r])) / speeds[r] for r in range(len(speeds)))
    if standing > fresh * _REBALANCE_GAIN:
        return False
    _last_assignment["judged_speeds"] = speeds
    return True


def _remember_assignment(neg: Negotiation, out: List[List[Tuple[str, int]]], sizes: Dict[Tuple[str, int], int],
                         loads: Optional[Sequence[Tuple[str, int, int, bool]]] = None) -> None:
    global _last_assignment
    _last_assignment = {"key": neg.key(), "digest": neg.gathered[0][1], "out": out,
                        "bytes": [sum(sizes[u] for u in mine) for mine in out],
                        "speeds": [g[2] for g in neg.gathered], "sizes": sizes,
                        "loads": list(loads) if loads is not None else [(p, i, sz, False) for (p, i), sz in sizes.items()]}


def probe_if_unseeded(any_device_stageable: bool) -> None:
    """First take of replicated CUDA state in this process: measure the link instead of assuming equal ones (the
    estimate is refined by every take from then on, record_take_speed)."""
    if any_device_stageable and observed_speed() is None:
        seed = probe_speed()
        if seed:
            _speed_estimates[""] = [seed]


def partition_write_reqs(
    entries: Dict[str, Entry], write_reqs: Dict[str, List[WriteReq]], pg: Any, agree_on: Optional[str] = None,
    negotiated: Optional[Negotiation] = None,
) -> Tuple[Dict[str, Entry], Dict[str, List[WriteReq]]]:
    """Keep, on this rank, only its share of the replicated write requests (reference
    partitioner.py:169-233).  Entries must be consolidated afterwards
    (`consolidate_replicated_entries`).

    One small all-gather in the common case: the ranks exchange a digest of their write loads (plus their
    non-replicated bytes and measured device->host speed) and, when the digests agree -- replicated state is
    the same on every rank by definition --, each computes the assignment from its own copy of the loads; only
    ranks that disagree exchange the lists themselves, as the reference always does (partitioner.py:209-226).
    `agree_on` is any string the caller wants compared across the ranks in the same exchange (snapshot.py sends
    a digest of its replicated-path candidates and so saves a collective); if the ranks differ in it, every
    rank raises RanksDisagree before anything is assigned."""
    if not set(write_reqs.keys()).issubset(entries.keys()):
        raise RuntimeError(
            "Not all entries associated with the write reqs are passed in. "
            f"Missing: {set(write_reqs.keys()) - set(entries.keys())}."
        )
    if os.environ.get("TORCH_SNAPSHOT_DISABLE_PARTITIONER") is not None:
        raise NotImplementedError("TORCH_SNAPSHOT_DISABLE_PARTITIONER is not implemented.")
    world = pg.get_world_size()
    rank = pg.get_rank()
    if world == 1:
        return entries, write_reqs
    replicated = {k: v for k, v in entries.items() if is_replicated(v)}
    if not replicated and agree_on is None:
        # (with `agree_on` the exchange below is also the caller's agreement check: every rank takes part,
        # whether or not it has replicated entries of its own)
        return entries, write_reqs
    non_replicated_size = sum(
        _write_req_size(wr) for k, wrs in write_reqs.items() if k not in replicated for wr in wrs
    )
    loads = []
    for path, entry in replicated.items():
        sub = getattr(entry, "type", None) == "ChunkedTensor"
        for idx, wr in enumerate(write_reqs.get(path, [])):
            loads.append((path, idx, _write_req_size(wr), sub))
    if negotiated is None:
        probe_if_unseeded(observed_speed() is None and any(
            getattr(wr.buffer_stager, "is_device_stageable", lambda: False)()
            for path in replicated for wr in write_reqs.get(path, [])[:1]
        ))
        negotiated = negotiate(pg, non_replicated_size, _digest(loads), agree_on)
    gathered = negotiated.gathered
    if negotiated.same_loads:
        base_loads = loads
    else:
        # rank 0's view defines the units, like the reference; a chunked entry is only split when every
        # rank chunked it identically
        full: List[Any] = [None] * world
        pg.all_gather_object(full, loads, large=True)
        base_loads = [(p, i, s, False) for p, i, s, _sub in full[0]]
    speeds = [g[2] for g in gathered]
    out = assign_write_loads(base_loads, [g[0] for g in gathered], speeds if all(speeds) else None)
    if negotiated.same_loads:
        _remember_assignment(negotiated, out, {(p, i): sz for p, i, sz, _sub in base_loads}, base_loads)
    mine = sorted(out[rank])
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
    global last_partition_inputs
    last_partition_inputs = (non_replicated_size, negotiated.gathered[rank][1])
    return new_entries, new_wrs


# (non-replicated bytes, digest of the loads) this rank sent in its last partition_write_reqs: what it sends again,
# without rebuilding the loads, when it repeats a take (snapshot.py)
last_partition_inputs: Optional[Tuple[int, str]] = None


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
    pg.all_gather_object(obj_list=gathered, obj=entries, large=True)
    return consolidate_replicated_entries(rank_to_entries=gathered, dedup=dedup)[pg.get_rank()]
