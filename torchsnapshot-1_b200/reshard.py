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

Only metadata crosses the process group (object all-gathers and barriers); the payload moves
over storage -> PCIe once and then over NVLink.  The result is bit-identical to the storage-only path
(tests/test_gpu_multi.py checks both against each other and against the source data).
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

    @property
    def key(self) -> UnitKey:
        br = self.byte_range
        return (self.location, br[0], br[1]) if br is not None else (self.location, -1, -1)


def plan_unit_owners(needs_by_rank: Sequence[Dict[UnitKey, Tuple[int, int]]]) -> Dict[UnitKey, int]:
    """Owner rank per unit from every rank's {key: (need bytes, unit bytes)}; same rule as plan_owners."""
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


def scatter_keys(needs_by_rank: Sequence[Dict[UnitKey, Tuple[int, int]]]) -> List[UnitKey]:
    """Units whose bytes are PARTITIONED among the ranks (the needs add up to about one copy of the unit: a
    reshard), as opposed to replicated state every rank needs whole.  Their owner packs, per destination
    rank, exactly the bytes that rank asked for into a buffer of its own; the destination then maps and
    reads only that buffer.  Deterministic from the gathered needs."""
    total: Dict[UnitKey, int] = {}
    size: Dict[UnitKey, int] = {}
    for d in needs_by_rank:
        for key, (need, nbytes) in d.items():
            total[key] = total.get(key, 0) + need
            size[key] = max(size.get(key, 0), nbytes)
    return sorted(k for k in total if total[k] * 4 <= size[k] * 5)


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
            for tensor, off, sizes in locals_:
                region = overlap_wrt_saved(piece.offsets, piece.sizes, off, sizes)
                if region is None:
                    continue
                n = elsize
                for _d, _so, _do, length in region:
                    n *= length
                need += n
                overlaps.append((tensor, region))
            if not overlaps:
                continue

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

            units.append(Unit(piece.location, piece.byte_range, piece.nbytes, need, make))
    return units


def read_req_units(read_reqs: Sequence[ReadReq]) -> List[Unit]:
    """Units for (batched) read requests whose consumers can unpack on the device (replicated state)."""
    units = []
    for rr in read_reqs:
        c = rr.buffer_consumer
        n = rr.byte_range[1] - rr.byte_range[0] if rr.byte_range is not None else c.consumed_nbytes()
        units.append(Unit(rr.path, rr.byte_range, n, n, c.unpack_regions))
    return units


def p2p_restore_units(
    units: Sequence[Unit],
    storage: StoragePlugin,
    pg: Any,
    event_loop: asyncio.AbstractEventLoop,
    device: int,
) -> ReshardStats:
    """Collective over `pg` (every rank calls it, possibly with no units): each unit is read from storage
    by ONE rank, and every rank that needs it copies out of that rank's device staging -- its own, or a
    peer's through a CUDA IPC mapping, i.e. over NVLink."""
    global last_stats
    from .scheduler import _Allocator, _read_into_pinned  # shared read machinery

    stats = ReshardStats()
    t_begin = time.monotonic()
    rank, world = pg.get_rank(), pg.get_world_size()
    mine_by_key = {u.key: u for u in units}
    gathered: List[Any] = [None] * world
    pg.all_gather_object(gathered, {k: (u.need, u.nbytes) for k, u in mine_by_key.items()})
    owners = plan_unit_owners(gathered)
    if not owners:
        return stats
    last_stats = stats  # only calls that had something to do are remembered
    eng = get_engine(device)
    ctx = eng.ctx
    sizes: Dict[UnitKey, int] = {}
    for d in gathered:
        for k, (_need, nbytes) in d.items():
            sizes[k] = max(sizes.get(k, 0), nbytes)
    owned = sorted(k for k, o in owners.items() if o == rank)
    staged: Dict[UnitKey, int] = {}
    blocks: List[int] = []  # device arena blocks to free at the end
    opened: Dict[Tuple[int, bytes], int] = {}
    # ---- partitioned units (a reshard): the owner packs per destination, see scatter_keys() ----------
    scatter = set(scatter_keys(gathered)) if knobs.is_p2p_scatter_enabled() else set()
    my_regions: Dict[UnitKey, List[RegionTuple]] = {}  # regions of the scatter units I pull, source relative to the unit
    wanted_from: Dict[int, Dict[UnitKey, List[Request]]] = {}
    for key, unit in mine_by_key.items():
        if key in scatter and owners.get(key) not in (None, rank) and unit.need:
            regs = unit.regions(0)
            my_regions[key] = regs
            wanted_from.setdefault(owners[key], {})[key] = [(r[0], r[2], tuple(r[3]), tuple(r[4])) for r in regs]
    all_wanted: List[Any] = [None] * world
    if scatter:
        pg.all_gather_object(all_wanted, wanted_from)
    pack_by_key: Dict[UnitKey, List[RegionTuple]] = {}  # owner side: unit -> boxes to pack once it is resident
    peer_buffers: Dict[int, int] = {}  # destination rank -> its packed buffer on this rank
    executor = ThreadPoolExecutor(max_workers=4, thread_name_prefix="tsb-reshard")
    ipc_executor = ThreadPoolExecutor(max_workers=8, thread_name_prefix="tsb-ipc")
    try:
        # ---- staging for the units this rank owns.  Units that peers read whole (replicated state) share
        # ONE block that is a driver allocation of its own: every peer maps a single IPC handle, and what it
        # maps is exactly that block, whatever the arena's segments look like after earlier takes (mapping
        # costs 10-20 ms per GB of the allocation).  Units that are packed per destination stay private, in
        # the arena. -------------------------------------------------------------------------------------
        total_owned = sum((sizes[k] + 511) // 512 * 512 for k in owned)
        staging_error: Optional[str] = None
        if owned:
            try:
                for exported, keys in ((True, [k for k in owned if k not in scatter]),
                                       (False, [k for k in owned if k in scatter])):
                    if not keys:
                        continue
                    nbytes = sum((sizes[k] + 511) // 512 * 512 for k in keys)
                    base = ctx.dev_alloc_exportable(nbytes) if exported and world > 1 else ctx.dev_alloc(nbytes)
                    blocks.append(base)
                    off = 0
                    for k in keys:
                        staged[k] = base + off
                        off += (sizes[k] + 511) // 512 * 512
            except Exception:  # no room for one block of that size: one arena block per unit
                for b in blocks:
                    ctx.dev_free(b)
                blocks.clear()
                staged.clear()
                try:
                    for k in owned:
                        staged[k] = ctx.dev_alloc(sizes[k])
                        blocks.append(staged[k])
                except Exception as e:  # noqa: BLE001 - told to every rank right below
                    staging_error = f"rank {rank}: no device staging for {total_owned} bytes ({e})"
        # ---- per-destination buffers for the partitioned units this rank owns ---------------------
        if scatter and staging_error is None:
            try:
                for peer in range(world):
                    asked = (all_wanted[peer] or {}).get(rank) if peer != rank else None
                    if not asked:
                        continue
                    offsets, total = packed_layout(asked)
                    buf = ctx.dev_alloc_exportable(total)
                    blocks.append(buf)
                    peer_buffers[peer] = buf
                    for key, reqs in asked.items():
                        for i, (src_off, run_bytes, bsizes, sstrides) in enumerate(reqs):
                            pack_by_key.setdefault(key, []).append(
                                (staged[key] + src_off, buf + offsets[(key, i)], run_bytes, tuple(bsizes),
                                 tuple(sstrides), packed_strides(run_bytes, bsizes))
                            )
            except Exception as e:  # noqa: BLE001 - told to every rank right below
                staging_error = f"rank {rank}: no per-destination staging ({e})"
        # ---- addresses are known before any byte is read: publish them now (IPC handle + offset) ---
        exports: Any = staging_error or {
            "shared": {k: ctx.ipc_export(dptr) for k, dptr in staged.items() if k not in scatter},
            "scatter": {peer: ctx.ipc_export(buf) for peer, buf in peer_buffers.items()},
        }
        all_exports: List[Any] = [None] * world
        pg.all_gather_object(all_exports, exports)
        errors = [x for x in all_exports if isinstance(x, str)]
        if errors:  # every rank sees the same list, so every rank raises instead of waiting on a peer
            raise RuntimeError("peer-to-peer restore could not stage its reads: " + "; ".join(errors))
        t_read = time.monotonic()
        stats.phase_ms["plan"] = 1e3 * (t_read - t_begin)
        regions: List[RegionTuple] = []

        def open_peer(hk: Tuple[int, bytes]) -> None:
            t0 = time.monotonic()
            opened[hk] = ctx.ipc_open(hk[1])
            stats.phase_ms["ipc_open"] = stats.phase_ms.get("ipc_open", 0.0) + 1e3 * (time.monotonic() - t0)

        async def build_regions() -> None:
            loop = asyncio.get_running_loop()
            # mapping a peer's staging block costs milliseconds per GiB: do it on worker threads, all peers
            # at once, while the event loop keeps the storage reads and H2D copies going
            wanted = {(owners[key], all_exports[owners[key]]["shared"][key][0]) for key, unit in mine_by_key.items()
                      if owners.get(key) is not None and owners[key] != rank and unit.need and key not in scatter}
            wanted |= {(owner, all_exports[owner]["scatter"][rank][0]) for owner in wanted_from}
            if wanted:
                await asyncio.gather(*(loop.run_in_executor(ipc_executor, open_peer, hk) for hk in sorted(wanted)))
            n = 0
            for key, unit in mine_by_key.items():
                owner = owners.get(key)
                if owner is None or unit.need == 0 or key in my_regions:
                    continue
                if owner == rank:
                    src_base = staged[key]
                    stats.local_bytes += unit.need
                else:
                    handle, offset = all_exports[owner]["shared"][key]
                    src_base = opened[(owner, handle)] + offset
                    stats.peer_bytes += unit.need
                regions.extend(unit.regions(src_base))
                n += 1
                if n % 8 == 0:
                    await asyncio.sleep(0)  # let the read pipeline dispatch
            # partitioned units owned by a peer: my boxes lie back to back in that peer's buffer for me
            for owner, asked in wanted_from.items():
                handle, offset = all_exports[owner]["scatter"][rank]
                base = opened[(owner, handle)] + offset
                offsets, _total = packed_layout(asked)
                for key in asked:
                    for i, r in enumerate(my_regions[key]):
                        regions.append((base + offsets[(key, i)], r[1], r[2], tuple(r[3]),
                                        packed_strides(r[2], r[3]), tuple(r[5])))
                    stats.peer_bytes += mine_by_key[key].need

        async def read_all() -> None:
            loop = asyncio.get_running_loop()
            alloc = _Allocator()
            from .scheduler import read_concurrency

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

        failure: Optional[str] = None
        try:
            event_loop.run_until_complete(read_all())
            if pack_by_key:
                ctx.stream_sync(STREAM_PACK)  # the per-destination buffers are complete before anybody is told so
        except Exception as e:  # noqa: BLE001 - reported to every rank below
            failure = f"rank {rank}: {type(e).__name__}: {e}"
        t_bar = time.monotonic()
        stats.phase_ms["read"] = 1e3 * (t_bar - t_read)
        # "every owner's units are resident" -- or somebody failed, and then everybody raises instead of
        # waiting forever for bytes that will never arrive
        verdicts: List[Any] = [None] * world
        pg.all_gather_object(verdicts, failure)
        failed = [v for v in verdicts if v is not None]
        if failed:
            raise RuntimeError("peer-to-peer restore failed while reading from storage: " + "; ".join(failed))
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
        t_end = time.monotonic()
        stats.phase_ms["copy"] = 1e3 * (t_end - t_copy)
        pg.barrier()  # nobody may free its staging while a peer still reads it
        stats.phase_ms["barrier"] = 1e3 * (time.monotonic() - t_end)
    finally:
        t_cleanup = time.monotonic()
        executor.shutdown(wait=False)
        ipc_executor.shutdown(wait=True)
        for base_addr in opened.values():
            try:
                ctx.ipc_close(base_addr)
            except Exception:  # pragma: no cover
                pass
        if world > 1:
            # exported blocks are driver allocations that are freed for good below: every peer has to have
            # dropped its mapping first (every rank that got this far has units, so every rank is here)
            try:
                pg.barrier()
            except Exception:  # pragma: no cover
                pass
        for dptr in blocks:
            ctx.dev_free(dptr)
        stats.phase_ms["cleanup"] = 1e3 * (time.monotonic() - t_cleanup)  # unmap peers, barrier, free exported blocks
    return stats


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
