"""Reshard-restore of sharded tensors with peer-to-peer copies inside the box.

The reference restores a ShardedTensor by letting EVERY rank read, from storage, every persisted piece
that intersects one of its local shards (torchsnapshot/io_preparers/sharded_tensor.py:196-268): when
the layout changes (row-wise saved, column-wise loaded; 8 ranks saved, 4 loading) each piece is read
and sent over PCIe up to world-size times.  Here

  1. the overlap plan is computed on the host, identically on every rank (`plan_owners`): every
     persisted piece gets one OWNER rank -- the rank that needs most of it, ties broken towards the
     least loaded rank;
  2. each owner reads its pieces once (storage -> pinned -> H2D) into its device staging arena;
  3. every rank then fills its local shards with ONE multi-region copy launch whose sources are its
     own staging blocks or, through CUDA IPC mappings, the staging blocks of its peers -- strided
     boxes read directly over NVLink / NVSwitch by the same kernel that packs and unpacks slabs.

Two kinds of unit are told apart from the gathered needs (`scatter_keys`).  Replicated state -- every rank
needs the whole unit -- is read by every destination straight out of the owner's staging block.  A reshard
-- the needs of the ranks add up to one copy of the unit -- goes through per-destination buffers: the
destinations tell the owner which strided boxes of the unit they want (`Request`, one more object
all-gather), the owner packs, as soon as the unit is resident, each destination's boxes back to back into a
buffer of that destination's own, and the destination maps only that buffer and reads it with contiguous
loads.  That is what keeps the CUDA IPC mapping cost (10-20 ms per GB of the mapped ALLOCATION) proportional
to what a rank receives instead of to what its peers hold, and turns 64-byte strided reads over NVLink into
flat ones.

Only metadata crosses the process group (three object all-gathers per call); the payload moves
over storage -> PCIe once and then over NVLink.  The result is bit-identical to the storage-only path
(tests/test_gpu_multi.py checks both against each other and against the source data).

What a restore leaves behind for the next one (`_PeerState`, per engine and process group): the exported
staging blocks go back to a pool instead of to the driver, and the peers keep their mappings of them, so from
the second restore on there is no cudaMalloc, no cudaIpcOpenMemHandle, no unmapping and no cudaFree on the
path (at N=8 those were 1.0-1.6 s of mapping summed over the peers plus 80-155 ms of cleanup per restore).
TSB200_TRIM_ARENAS=1 gives everything back at the end of every call instead.

When the path cannot or should not run, every rank learns it from the same gathered data and the caller
loads the same objects the ordinary way (`ReshardStats.fell_back`): no staging memory on some rank, a peer
allocation that cannot be exported or mapped (ranks on different hosts never get here, snapshot.py checks),
or a plan in which the storage-only path would read (almost) nothing twice -- then sharing reads buys nothing.
"""

import asyncio
import logging
import time
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch

from . import knobs
from ._native import STREAM_H2D, STREAM_PACK
from .engine import get_engine
from .io_preparer import ShardedTensorIOPreparer, overlap_wrt_saved
from .io_types import ReadReq, StoragePlugin
from .manifest import ShardedTensorEntry
from .regions import RegionTuple, box_regions, contiguous_strides, regions_to_array
from .serialization import Serializer, dtype_to_element_size, string_to_dtype

logger: logging.Logger = logging.getLogger(__name__)

Box = Tuple[List[int], List[int]]  # (global offsets, sizes)
UnitKey = Tuple[str, int, int]  # (storage location, lower byte, upper byte; -1,-1 = whole object)


@dataclass
class Piece:
    """One persisted piece of a sharded tensor."""

    path: str  # logical path of the sharded tensor
    location: str
    byte_range: Optional[Tuple[int, int]]
    offsets: List[int]
    sizes: List[int]
    dtype: str
    nbytes: int


def pieces_of(path: str, entry: ShardedTensorEntry) -> List[Piece]:
    out = []
    for shard in entry.shards:
        t = shard.tensor
        n = dtype_to_element_size(string_to_dtype(t.dtype))
        for s in shard.sizes:
            n *= s
        out.append(Piece(path, t.location, t.byte_range_tuple, list(shard.offsets), list(shard.sizes), t.dtype, n))
    return out


def _overlap_bytes(piece: Piece, box: Box, elsize: int) -> int:
    region = overlap_wrt_saved(piece.offsets, piece.sizes, box[0], box[1])
    if region is None:
        return 0
    n = elsize
    for _dim, _so, _do, length in region:
        n *= length
    return n


def plan_owners(pieces: Sequence[Piece], boxes_by_rank: Sequence[Dict[str, List[Box]]]) -> List[Optional[int]]:
    """Owner rank of every piece (None if nobody needs it), from every rank's local boxes.  Deterministic:
    every rank computes the same answer from the same inputs.  The owner is a rank with the largest
    overlap with the piece (so that as few bytes as possible cross NVLink); among (near-)equal candidates
    the one that has been given the fewest bytes so far (so that storage reads are spread over the ranks).
    Same rule, same code as the runtime plan (`plan_unit_owners`)."""
    needs: List[Dict[UnitKey, Tuple[int, int]]] = []
    keys = [(p.location, p.byte_range[0], p.byte_range[1]) if p.byte_range is not None else (p.location, -1, -1)
            for p in pieces]
    for boxes in boxes_by_rank:
        mine: Dict[UnitKey, Tuple[int, int]] = {}
        for piece, key in zip(pieces, keys):
            elsize = dtype_to_element_size(string_to_dtype(piece.dtype))
            need = sum(_overlap_bytes(piece, box, elsize) for box in boxes.get(piece.path, []))
            if need:
                mine[key] = (need, piece.nbytes)
        needs.append(mine)
    owners = plan_unit_owners(needs)
    return [owners.get(key) for key in keys]


def cross_rank_bytes(pieces: Sequence[Piece], owners: Sequence[Optional[int]],
                     boxes_by_rank: Sequence[Dict[str, List[Box]]]) -> int:
    """Payload bytes that have to cross NVLink under a plan (for reporting / roofline arithmetic)."""
    total = 0
    for piece, owner in zip(pieces, owners):
        if owner is None:
            continue
        elsize = dtype_to_element_size(string_to_dtype(piece.dtype))
        for r, boxes in enumerate(boxes_by_rank):
            if r != owner:
                total += sum(_overlap_bytes(piece, box, elsize) for box in boxes.get(piece.path, []))
    return total


def eligible(entry: ShardedTensorEntry, obj_out: Any, device: int) -> bool:
    """The P2P path handles buffer-protocol pieces restored into CUDA shards of the same dtype on
    `device`; everything else goes through the storage-only path."""
    try:
        locals_ = ShardedTensorIOPreparer.local_boxes(obj_out)
    except RuntimeError:
        return False
    if not locals_ or not entry.shards:
        return False
    dtype = string_to_dtype(entry.shards[0].tensor.dtype)
    for shard in entry.shards:
        if shard.tensor.serializer != Serializer.BUFFER_PROTOCOL.value or shard.tensor.dtype != entry.shards[0].tensor.dtype:
            return False
    return all(t.is_cuda and t.device.index == device and t.dtype == dtype for t, _o, _s in locals_)


class ReshardStats:
    def __init__(self) -> None:
        self.read_bytes = 0  # bytes this rank read from storage (it owns these units)
        self.local_bytes = 0  # bytes copied from this rank's own staging
        self.peer_bytes = 0  # bytes this rank pulled over NVLink
        self.copy_ms = 0.0  # device time of the copy kernels (CUDA events)
        self.launches = 0  # launches that fill this rank's tensors (one per call)
        self.pack_launches = 0  # owner side: units split into per-destination buffers (one per owned unit)
        self.skipped: List[str] = []  # sharded paths some rank did not offer: load them the ordinary way
        self.phase_ms: Dict[str, float] = {}  # wall-clock phases: plan, read, exchange, copy, barrier
        self.fell_back = False  # every rank agreed not to use this path: the caller loads the objects the ordinary way
        self.reason = ""
        self.amplification: Dict[str, float] = {}  # per group: bytes the storage-only path would read / bytes this path reads
        self.skipped_groups: List[str] = []  # groups every rank left to the ordinary path (sharing their reads saves nothing)
        self.new_staging_bytes = 0  # exportable staging obtained from the driver in this call (0 once the pool is warm)
        self.opens = 0  # peer allocations mapped in this call (0 once the mappings are cached)


last_stats: Optional[ReshardStats] = None

@dataclass
class Unit:
    """A byte range of one storage object that one or more ranks need on the device: a persisted piece
    of a sharded tensor, or the merged range of the replicated members of a slab file."""

    location: str
    byte_range: Optional[Tuple[int, int]]
    nbytes: int
    need: int  # payload bytes THIS rank takes out of the unit
    regions: Callable[[int], List[RegionTuple]]  # device address of the unit's first byte -> copy regions
    solo: int = -1  # bytes this rank would read from storage for the unit on the storage-only path (-1: all of it)
    group: str = ""  # units of one group take the path together or not at all ("sharded", "replicated")

    @property
    def key(self) -> UnitKey:
        br = self.byte_range
        return (self.location, br[0], br[1]) if br is not None else (self.location, -1, -1)


def plan_unit_owners(needs_by_rank: Sequence[Dict[UnitKey, Tuple[int, ...]]]) -> Dict[UnitKey, int]:
    """Owner rank per unit from every rank's {key: (need bytes, unit bytes, ...)}; same rule as plan_owners."""
    load = [0] * len(needs_by_rank)
    owners: Dict[UnitKey, int] = {}
    for key in sorted(set().union(*(d.keys() for d in needs_by_rank))):
        need = [d.get(key, (0, 0))[0] for d in needs_by_rank]
        nbytes = max(d.get(key, (0, 0))[1] for d in needs_by_rank)
        best = max(need)
        if best == 0:
            continue
        candidates = [r for r, n in enumerate(need) if n * 1000 >= best * 999]
        owner = min(candidates, key=lambda r: (load[r], r))
        load[owner] += nbytes
        owners[key] = owner
    return owners


# A request: what one rank wants out of one unit, as strided boxes relative to the unit's first byte --
# (source offset, run bytes, outer sizes, source byte strides).  The destination half of the same region
# (address and strides in the requesting rank's memory) never leaves that rank.
Request = Tuple[int, int, Tuple[int, ...], Tuple[int, ...]]
_PACK_ALIGN = 512  # start of every packed box inside a per-destination buffer


def scatter_keys(needs_by_rank: Sequence[Dict[UnitKey, Tuple[int, ...]]]) -> List[UnitKey]:
    """Units whose bytes are PARTITIONED among the ranks (the needs add up to about one copy of the unit: a
    reshard), as opposed to replicated state every rank needs whole.  Their owner packs, per destination
    rank, exactly the bytes that rank asked for into a buffer of its own; the destination then maps and
    reads only that buffer.  Deterministic from the gathered needs."""
    total: Dict[UnitKey, int] = {}
    size: Dict[UnitKey, int] = {}
    for d in needs_by_rank:
        for key, v in d.items():
            need, nbytes = v[0], v[1]
            total[key] = total.get(key, 0) + need
            size[key] = max(size.get(key, 0), nbytes)
    return sorted(k for k in total if total[k] * 4 <= size[k] * 5)


def read_amplification(needs_by_rank: Sequence[Dict[UnitKey, Tuple[Any, ...]]]) -> Dict[str, float]:
    """Per group of units: bytes the storage-only path would read (every rank reads, on its own, the part of
    every unit it overlaps: {key: (need, unit bytes, solo bytes, group)}) over the bytes the shared-read path
    reads (every needed unit once).  Replicated state restored by N ranks: N.  A row-wise table re-cut
    row-wise (8 -> 4 ranks, shifted rows): 1 -- every rank reads only rows it needs either way, and routing
    them through an owner and NVLink only adds collectives and a staging pass."""
    once: Dict[str, Dict[UnitKey, int]] = {}
    solo_total: Dict[str, int] = {}
    for d in needs_by_rank:
        for key, v in d.items():
            if v[0] == 0:
                continue
            group = v[3] if len(v) > 3 else ""
            per = once.setdefault(group, {})
            per[key] = max(per.get(key, 0), v[1])
            solo_total[group] = solo_total.get(group, 0) + (v[2] if len(v) > 2 and v[2] >= 0 else v[1])
    return {g: (solo_total[g] / sum(per.values()) if sum(per.values()) else 0.0) for g, per in once.items()}


def packed_strides(run_bytes: int, sizes: Sequence[int]) -> Tuple[int, ...]:
    """Byte strides of the outer dims when the runs of a box are laid out back to back."""
    out = [0] * len(sizes)
    acc = run_bytes
    for i in range(len(sizes) - 1, -1, -1):
        out[i] = acc
        acc *= int(sizes[i])
    return tuple(out)


def packed_layout(requests: Dict[UnitKey, List[Request]]) -> Tuple[Dict[Tuple[UnitKey, int], int], int]:
    """Where each requested box starts inside the per-destination buffer, and the buffer's size.  The owner
    and the destination both call this on the same request table, so no offsets are exchanged."""
    offsets: Dict[Tuple[UnitKey, int], int] = {}
    cursor = 0
    for key in sorted(requests):
        for i, (_src_off, run_bytes, sizes, _strides) in enumerate(requests[key]):
            offsets[(key, i)] = cursor
            n = run_bytes
            for s in sizes:
                n *= int(s)
            cursor += (n + _PACK_ALIGN - 1) // _PACK_ALIGN * _PACK_ALIGN
    return offsets, cursor


def sharded_units(items: Sequence[Tuple[str, ShardedTensorEntry, Any]]) -> List[Unit]:
    """Units (one per persisted piece this rank overlaps) for sharded tensors restored into CUDA shards."""
    units: List[Unit] = []
    for path, entry, obj in items:
        locals_ = ShardedTensorIOPreparer.local_boxes(obj)
        for piece in pieces_of(path, entry):
            elsize = dtype_to_element_size(string_to_dtype(piece.dtype))
            overlaps = []
            need = 0
            row_lo, row_hi = None, 0
            for tensor, off, sizes in locals_:
                region = overlap_wrt_saved(piece.offsets, piece.sizes, off, sizes)
                if region is None:
                    continue
                n = elsize
                for _d, _so, _do, length in region:
                    n *= length
                need += n
                overlaps.append((tensor, region))
                if region:
                    row_lo = region[0][1] if row_lo is None else min(row_lo, region[0][1])
                    row_hi = max(row_hi, region[0][1] + region[0][3])
            if not overlaps:
                continue
            # what the storage-only path reads of this piece: the rows any local shard overlaps, or the whole
            # piece when that saves less than an eighth (ShardedTensorIOPreparer._trim_piece_to_needed_rows)
            solo = piece.nbytes
            rows = piece.sizes[0] if piece.sizes else 0
            if rows and row_lo is not None and (row_hi - row_lo) * 8 <= rows * 7:
                solo = piece.nbytes // rows * (row_hi - row_lo)

            def make(base: int, piece: Piece = piece, overlaps: Any = overlaps, elsize: int = elsize) -> List[RegionTuple]:
                src_strides = contiguous_strides(piece.sizes, elsize)
                out: List[RegionTuple] = []
                for tensor, region in overlaps:
                    dst = tensor.detach()
                    dst_strides = [s * elsize for s in dst.stride()]
                    src_addr, dst_addr, lens = base, dst.data_ptr(), []
                    for dim, so, do, length in region:
                        src_addr += so * src_strides[dim]
                        dst_addr += do * dst_strides[dim]
                        lens.append(length)
                    out += box_regions(src_addr, src_strides, dst_addr, dst_strides, lens, elsize)
                return out

            units.append(Unit(piece.location, piece.byte_range, piece.nbytes, need, make, solo, "sharded"))
    return units


def read_req_units(read_reqs: Sequence[ReadReq]) -> List[Unit]:
    """Units for (batched) read requests whose consumers can unpack on the device (replicated state)."""
    units = []
    for rr in read_reqs:
        c = rr.buffer_consumer
        n = rr.byte_range[1] - rr.byte_range[0] if rr.byte_range is not None else c.consumed_nbytes()
        units.append(Unit(rr.path, rr.byte_range, n, n, c.unpack_regions, n, "replicated"))
    return units


_POOL_GRANULE = 8 << 20  # new pool blocks are rounded up to this, so that a slightly larger plan still fits


class _PeerState:
    """What one restore leaves for the next, per (engine, process group): exportable staging blocks that are
    not in use (`free`: base -> size; they stay allocated, and mapped by the peers that read them), and the
    peer allocations this rank has mapped (`opened`: (owner rank, IPC handle) -> base address here)."""

    def __init__(self) -> None:
        self.free: Dict[int, int] = {}
        self.opened: Dict[Tuple[int, bytes], int] = {}
        self.want_flush = False  # the pool outgrew its bound: the next call starts by giving everything back
        self._sizes: Dict[int, int] = {}  # blocks handed out: base -> size
        self._tags: Dict[int, Any] = {}  # what a block was last used for: "shared", or the destination rank it served

    def pooled_bytes(self) -> int:
        return sum(self.free.values())

    def acquire(self, ctx: Any, nbytes: int, tag: Any, stats: "ReshardStats") -> int:
        """An exportable block of at least `nbytes`: a free one that is not wastefully large -- first choice the
        one that served the same purpose last time (the peer that reads it then has it mapped already), then
        the tightest -- else a new driver allocation (rounded up, so that a slightly larger plan still fits)."""
        fit = [(self._tags.get(base) != tag, size, base) for base, size in self.free.items()
               if nbytes <= size <= max(2 * nbytes, nbytes + (64 << 20))]
        if fit:
            _other, size, base = min(fit)
            del self.free[base]
        else:
            # headroom: the units a rank owns shift a little from restore to restore (the writes of replicated
            # state are re-partitioned by measured speed), and a block that is a few MiB short is a new
            # cudaMalloc plus a new mapping on every peer
            want = max(nbytes, 1) + (max(nbytes, 1) >> 2 if _POOL_GRANULE > 512 else 0)
            size = (want + _POOL_GRANULE - 1) // _POOL_GRANULE * _POOL_GRANULE
            base = ctx.dev_alloc_exportable(size)
            stats.new_staging_bytes += size
        self._sizes[base] = size
        self._tags[base] = tag
        return base

    def release(self, base: int) -> None:
        self.free[base] = self._sizes.pop(base)


def _peer_state(eng: Any, pg: Any) -> _PeerState:
    states = eng.__dict__.setdefault("_p2p_states", {})
    key = (id(getattr(pg, "pg", pg)), pg.get_rank(), pg.get_world_size())
    st = states.get(key)
    if st is None:
        st = states[key] = _PeerState()
    return st


def _pool_bound(eng: Any) -> int:
    import os

    if "TSB200_P2P_POOL_BYTES" in os.environ:
        return int(os.environ["TSB200_P2P_POOL_BYTES"])
    return int(getattr(eng, "dev_bound", 16 << 30))


def _give_back(state: _PeerState, ctx: Any, pg: Any) -> None:
    """Collective: every rank drops its peer mappings, then (after a barrier: nobody may still have a block
    mapped when its owner frees it) returns its unused staging blocks to the driver."""
    for base in state.opened.values():
        try:
            ctx.ipc_close(base)
        except Exception:  # pragma: no cover
            pass
    state.opened.clear()
    if pg.get_world_size() > 1:
        pg.barrier()
    for base in list(state.free):
        ctx.dev_free(base)
        state._tags.pop(base, None)
    state.free.clear()
    state.want_flush = False


def release_staging(device: int, pg: Any) -> None:
    """Collective over `pg`: return the staging blocks and peer mappings that peer-to-peer restores keep
    between calls (a trainer that restores once at start-up wants that HBM back)."""
    eng = get_engine(device)
    _give_back(_peer_state(eng, pg), eng.ctx, pg)


def min_amplification() -> float:
    """Below this read amplification of the storage-only path the peer-to-peer path is not worth its
    collectives (TSB200_P2P_MIN_AMPLIFICATION; 0 forces the path whenever it is eligible)."""
    import os

    return float(os.environ.get("TSB200_P2P_MIN_AMPLIFICATION", "1.25"))


def p2p_restore_units(
    units: Sequence[Unit],
    storage: StoragePlugin,
    pg: Any,
    event_loop: asyncio.AbstractEventLoop,
    device: int,
) -> ReshardStats:
    """Collective over `pg` (every rank calls it, possibly with no units): each unit is read from storage
    by ONE rank, and every rank that needs it copies out of that rank's device staging -- its own, or a
    peer's through a CUDA IPC mapping, i.e. over NVLink.

    Returns with `fell_back` set -- on every rank alike, nothing restored -- when the path is not used (see
    the module docstring); the caller then loads the same objects through the ordinary read path.  A failed
    storage read raises on every rank."""
    global last_stats
    from .scheduler import _Allocator, _read_into_pinned, read_concurrency  # shared read machinery

    stats = ReshardStats()
    t_begin = time.monotonic()
    rank, world = pg.get_rank(), pg.get_world_size()
    mine_by_key = {u.key: u for u in units}
    eng = get_engine(device)
    ctx = eng.ctx
    state = _peer_state(eng, pg)
    keep = not knobs.trim_arenas_after_use()
    # ---- collective 1: what every rank needs of which unit, and -- for units it needs only part of -- which
    # strided boxes (source side only, relative to the unit's first byte: the destination half of a region
    # never leaves its rank).  Owners, the replicated / reshard split, and what every owner packs for whom
    # follow from this one exchange, identically on every rank. ---------------------------------------------
    my_regions: Dict[UnitKey, List[RegionTuple]] = {}
    geometry: Dict[UnitKey, List[Request]] = {}
    for key, unit in mine_by_key.items():
        if 0 < unit.need < unit.nbytes:
            regs = unit.regions(0)
            my_regions[key] = regs
            geometry[key] = [(r[0], r[2], tuple(r[3]), tuple(r[4])) for r in regs]
    hello = {
        "needs": {k: (u.need, u.nbytes, u.solo if u.solo >= 0 else u.nbytes, u.group) for k, u in mine_by_key.items()},
        "geometry": geometry,
        "flush": state.want_flush,
        "scatter": knobs.is_p2p_scatter_enabled(),
    }
    hellos: List[Any] = [None] * world
    pg.all_gather_object(hellos, hello)
    gathered = [h["needs"] for h in hellos]
    if any(h["flush"] for h in hellos):
        _give_back(state, ctx, pg)
    if not any(v[0] for d in gathered for v in d.values()):
        return stats
    last_stats = stats  # only calls that had something to do are remembered
    # groups whose reads are not worth sharing are left to the caller's ordinary path, by every rank alike
    stats.amplification = read_amplification(gathered)
    stats.skipped_groups = sorted(g for g, a in stats.amplification.items() if a < min_amplification())
    if stats.skipped_groups:
        skip = set(stats.skipped_groups)
        gathered = [{k: v for k, v in d.items() if v[3] not in skip} for d in gathered]
        mine_by_key = {k: u for k, u in mine_by_key.items() if u.group not in skip}
        stats.reason = "; ".join(
            f"{g or 'units'}: the storage-only path reads {stats.amplification[g]:.2f}x the bytes of the shared-read "
            "path, nothing to gain from sharing reads" for g in stats.skipped_groups)
    owners = plan_unit_owners(gathered)
    if not owners:
        stats.fell_back = True
        return stats
    sizes: Dict[UnitKey, int] = {}
    for d in gathered:
        for k, v in d.items():
            sizes[k] = max(sizes.get(k, 0), v[1])
    owned = sorted(k for k, o in owners.items() if o == rank)
    scatter = set(scatter_keys(gathered)) if all(h["scatter"] for h in hellos) else set()
    # what destination d asks of owner o: {d: {o: {key: [requests]}}}, from the gathered geometry
    wanted: List[Dict[int, Dict[UnitKey, List[Request]]]] = [{} for _ in range(world)]
    for d, h in enumerate(hellos):
        for key, reqs in h["geometry"].items():
            owner = owners.get(key)
            if key in scatter and owner is not None and owner != d and h["needs"][key][0]:
                wanted[d].setdefault(owner, {})[key] = [tuple(r) for r in reqs]  # type: ignore[misc]
    wanted_from = wanted[rank]
    staged: Dict[UnitKey, int] = {}
    pooled: List[int] = []  # exportable blocks taken from the pool for this call
    private: List[int] = []  # arena blocks only this rank reads
    pack_by_key: Dict[UnitKey, List[RegionTuple]] = {}  # owner side: unit -> boxes to pack once it is resident
    peer_buffers: Dict[int, int] = {}  # destination rank -> its packed buffer on this rank
    executor = ThreadPoolExecutor(max_workers=4, thread_name_prefix="tsb-reshard")
    ipc_executor = ThreadPoolExecutor(max_workers=8, thread_name_prefix="tsb-ipc")

    def unit_span(k: UnitKey) -> int:
        return (sizes[k] + 511) // 512 * 512

    def release_all() -> None:
        for base in pooled:
            state.release(base)
        pooled.clear()
        for base in private:
            try:
                ctx.dev_free(base)
            except Exception:  # pragma: no cover
                pass
        private.clear()
        if state.pooled_bytes() > _pool_bound(eng):
            state.want_flush = True

    fell_back_reason: Optional[str] = None
    try:
        # ---- staging for the units this rank owns.  Units that peers read whole (replicated state) share ONE
        # exportable block: every peer maps a single handle, and what it maps is exactly that block.  Units
        # that are packed per destination stay private, in the arena; each destination gets one exportable
        # buffer that holds only its boxes. --------------------------------------------------------------
        staging_error: Optional[str] = None
        try:
            shared_keys = [k for k in owned if k not in scatter]
            private_keys = [k for k in owned if k in scatter]
            if shared_keys:
                nbytes = sum(unit_span(k) for k in shared_keys)
                if world > 1:
                    base = state.acquire(ctx, nbytes, "shared", stats)
                    pooled.append(base)
                else:
                    base = ctx.dev_alloc(nbytes)
                    private.append(base)
                off = 0
                for k in shared_keys:
                    staged[k] = base + off
                    off += unit_span(k)
            if private_keys:
                try:
                    base = ctx.dev_alloc(sum(unit_span(k) for k in private_keys))
                    private.append(base)
                    off = 0
                    for k in private_keys:
                        staged[k] = base + off
                        off += unit_span(k)
                except Exception:  # no room for one block of that size: one arena block per unit
                    for k in private_keys:
                        staged[k] = ctx.dev_alloc(sizes[k])
                        private.append(staged[k])
            for peer in range(world):
                asked = wanted[peer].get(rank) if peer != rank else None
                if not asked:
                    continue
                offsets, total = packed_layout(asked)
                buf = state.acquire(ctx, total, peer, stats)
                pooled.append(buf)
                peer_buffers[peer] = buf
                for key, reqs in asked.items():
                    for i, (src_off, run_bytes, bsizes, sstrides) in enumerate(reqs):
                        pack_by_key.setdefault(key, []).append(
                            (staged[key] + src_off, buf + offsets[(key, i)], run_bytes, tuple(bsizes),
                             tuple(sstrides), packed_strides(run_bytes, bsizes))
                        )
            exports: Any = {
                "shared": {k: ctx.ipc_export(staged[k]) for k in shared_keys} if world > 1 else {},
                "scatter": {peer: ctx.ipc_export(buf) for peer, buf in peer_buffers.items()},
            }
        except Exception as e:  # noqa: BLE001 - told to every rank right below
            staging_error = f"rank {rank}: no device staging for the units it owns ({type(e).__name__}: {e})"
            exports = staging_error
        # ---- collective 2: addresses are known before any byte is read: publish them now (IPC handle + offset),
        # or the reason this rank cannot take part -- then nobody does -------------------------------------
        all_exports: List[Any] = [None] * world
        pg.all_gather_object(all_exports, exports)
        errors = [x for x in all_exports if isinstance(x, str)]
        if errors:
            fell_back_reason = "; ".join(errors)
            raise _FallBack()
        t_read = time.monotonic()
        stats.phase_ms["plan"] = 1e3 * (t_read - t_begin)
        regions: List[RegionTuple] = []
        ipc_failure: List[str] = []

        def open_peer(hk: Tuple[int, bytes]) -> None:
            t0 = time.monotonic()
            try:
                state.opened[hk] = ctx.ipc_open(hk[1])
                stats.opens += 1
            except Exception as e:  # noqa: BLE001 - reported to every rank after the reads
                ipc_failure.append(f"rank {rank} could not map the staging of rank {hk[0]} ({e})")
            stats.phase_ms["ipc_open"] = stats.phase_ms.get("ipc_open", 0.0) + 1e3 * (time.monotonic() - t0)

        async def build_regions() -> None:
            loop = asyncio.get_running_loop()
            # mapping a peer's staging block costs milliseconds per GiB: do it on worker threads, all peers
            # at once, while the event loop keeps the storage reads and H2D copies going.  Blocks that were
            # mapped by an earlier restore (the owner took them from its pool again) cost nothing.
            needed = {(owners[key], all_exports[owners[key]]["shared"][key][0]) for key, unit in mine_by_key.items()
                      if owners.get(key) is not None and owners[key] != rank and unit.need and key not in scatter}
            needed |= {(owner, all_exports[owner]["scatter"][rank][0]) for owner in wanted_from}
            missing = sorted(hk for hk in needed if hk not in state.opened)
            if missing:
                await asyncio.gather(*(loop.run_in_executor(ipc_executor, open_peer, hk) for hk in missing))
            if ipc_failure:
                return
            n = 0
            by_owner: Dict[int, List[RegionTuple]] = {}  # copy regions, grouped by the rank whose memory they read
            for key, unit in mine_by_key.items():
                owner = owners.get(key)
                if owner is None or unit.need == 0 or (key in scatter and owner != rank):
                    continue
                if owner == rank:
                    src_base = staged[key]
                    stats.local_bytes += unit.need
                else:
                    handle, offset = all_exports[owner]["shared"][key]
                    src_base = state.opened[(owner, handle)] + offset
                    stats.peer_bytes += unit.need
                by_owner.setdefault(owner, []).extend(unit.regions(src_base))
                n += 1
                if n % 8 == 0:
                    await asyncio.sleep(0)  # let the read pipeline dispatch
            # partitioned units owned by a peer: my boxes lie back to back in that peer's buffer for me
            for owner, asked in wanted_from.items():
                handle, offset = all_exports[owner]["scatter"][rank]
                base = state.opened[(owner, handle)] + offset
                offsets, _total = packed_layout(asked)
                for key in asked:
                    for i, r in enumerate(my_regions[key]):
                        by_owner.setdefault(owner, []).append((base + offsets[(key, i)], r[1], r[2], tuple(r[3]),
                                                               packed_strides(r[2], r[3]), tuple(r[5])))
                    stats.peer_bytes += mine_by_key[key].need
            # The launch walks its tiles in table order, so the order of the regions is the order in which this
            # rank visits its peers.  Every rank starts with its own staging and goes round the ring from there:
            # at any moment the ranks read from DIFFERENT owners instead of all draining owner 0, then owner 1, ...
            # (one owner's HBM and NVLink egress shared by all readers at once).
            for owner in sorted(by_owner, key=lambda o: (o - rank) % world):
                regions.extend(by_owner[owner])

        async def read_all() -> None:
            loop = asyncio.get_running_loop()
            alloc = _Allocator()
            sem = asyncio.Semaphore(read_concurrency())

            async def one(key: UnitKey) -> None:
                nbytes = sizes[key]
                br = None if key[1] < 0 else (key[1], key[2])
                await sem.acquire()  # the slot covers the storage read only, not the H2D that follows
                held = True
                try:
                    pbuf = await alloc.pinned(eng, nbytes)
                    try:
                        rr = ReadReq(path=key[0], buffer_consumer=None, byte_range=br)  # type: ignore[arg-type]
                        await _read_into_pinned(loop, executor, storage, rr, pbuf, nbytes)
                        sem.release()
                        held = False
                        ctx.memcpy_h2d(STREAM_H2D, staged[key], pbuf.ptr, nbytes)
                        ev = ctx.event_record(STREAM_H2D)
                        packs = pack_by_key.get(key)
                        if packs:  # resident now: split it into the per-destination buffers (device side, async)
                            ctx.stream_wait_event(STREAM_PACK, ev)
                            eng.copy_region_array(regions_to_array(packs), STREAM_PACK)
                            stats.pack_launches += 1
                        try:
                            await loop.run_in_executor(executor, ctx.event_sync, ev)
                        finally:
                            ctx.event_release(ev)
                    finally:
                        pbuf.release()
                        alloc.notify(nbytes)
                finally:
                    if held:
                        sem.release()
                stats.read_bytes += nbytes

            # the copy regions are built while the reads are in flight
            await asyncio.gather(build_regions(), *(one(k) for k in owned))

        failure: Optional[Tuple[str, str]] = None
        try:
            event_loop.run_until_complete(read_all())
            if pack_by_key:
                ctx.stream_sync(STREAM_PACK)  # the per-destination buffers are complete before anybody is told so
        except Exception as e:  # noqa: BLE001 - reported to every rank below
            failure = ("read", f"rank {rank}: {type(e).__name__}: {e}")
        if failure is None and ipc_failure:
            failure = ("ipc", "; ".join(ipc_failure))
        t_bar = time.monotonic()
        stats.phase_ms["read"] = 1e3 * (t_bar - t_read)
        # ---- collective 3: "every owner's units are resident" -- or somebody failed, and then everybody
        # learns it here instead of waiting forever for bytes that will never arrive ----------------------
        verdicts: List[Any] = [None] * world
        pg.all_gather_object(verdicts, failure)
        failed = [v for v in verdicts if v is not None]
        if any(kind == "read" for kind, _msg in failed):
            raise RuntimeError(
                "peer-to-peer restore failed while reading from storage: " + "; ".join(m for _k, m in failed)
            )
        if failed:  # peer memory could not be mapped: the bytes are on storage all the same
            fell_back_reason = "; ".join(m for _k, m in failed)
            raise _FallBack()
        t_copy = time.monotonic()
        stats.phase_ms["exchange"] = 1e3 * (t_copy - t_bar)
        # ---- ONE launch: own staging + peers' staging (over NVLink) -> this rank's tensors ---------------
        arr = regions_to_array(regions)
        if arr.shape[0]:
            eng.wait_current_stream(STREAM_PACK)
            k0, k1 = eng.copy_region_array_timed(arr, STREAM_PACK)
            ctx.stream_sync(STREAM_PACK)
            stats.copy_ms += ctx.event_elapsed_ms(k0, k1)
            stats.launches += 1
            ctx.event_release(k0)
            ctx.event_release(k1)
        stats.phase_ms["copy"] = 1e3 * (time.monotonic() - t_copy)
    except _FallBack:
        stats.fell_back = True
        stats.reason = fell_back_reason or ""
        logger.warning(f"peer-to-peer restore not used, loading from storage on every rank: {stats.reason}")
    finally:
        t_cleanup = time.monotonic()
        executor.shutdown(wait=False)
        ipc_executor.shutdown(wait=True)
        if keep and not stats.fell_back:
            # Staging goes back to the pool, mapped as it is.  A peer may still be reading it: nothing
            # overwrites a pooled block before the first collective of the next call, which no rank enters
            # before its own copy launch has finished.
            release_all()
        else:
            release_all()
            try:
                _give_back(state, ctx, pg)  # unmap peers, barrier, free: every rank that got this far is here
            except Exception:  # pragma: no cover
                pass
        stats.phase_ms["cleanup"] = 1e3 * (time.monotonic() - t_cleanup)
    return stats


class _FallBack(Exception):
    """Raised (and caught) inside p2p_restore_units once every rank knows the path is not to be used."""


def p2p_reshard_restore(
    items: Sequence[Tuple[str, ShardedTensorEntry, Any]],
    storage: StoragePlugin,
    pg: Any,
    event_loop: asyncio.AbstractEventLoop,
    device: int,
    extra_units: Sequence[Unit] = (),
) -> ReshardStats:
    """Restore sharded `items` = (logical path, merged entry, runtime object), plus any `extra_units`
    (replicated state), with one shared-read + peer-copy round.  Collective over `pg`."""
    return p2p_restore_units(list(sharded_units(items)) + list(extra_units), storage, pg, event_loop, device)
