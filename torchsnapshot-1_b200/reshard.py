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

Only metadata crosses the process group (two object all-gathers and two barriers); the payload moves
over storage -> PCIe once and then over NVLink.  The result is bit-identical to the storage-only path
(tests/test_gpu_multi.py checks both against each other and against the source data).
"""

import asyncio
import logging
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

from ._native import STREAM_H2D, STREAM_PACK
from .engine import get_engine
from .io_preparer import ShardedTensorIOPreparer, overlap_wrt_saved
from .io_types import ReadReq, StoragePlugin
from .manifest import ShardedTensorEntry
from .regions import RegionTuple, box_regions, contiguous_strides, regions_to_array
from .serialization import Serializer, dtype_to_element_size, string_to_dtype

logger: logging.Logger = logging.getLogger(__name__)

Box = Tuple[List[int], List[int]]  # (global offsets, sizes)


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

    @property
    def key(self) -> Tuple[str, str, Optional[Tuple[int, int]]]:
        return (self.path, self.location, self.byte_range)


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
    """Owner rank of every piece (None if nobody needs it).  Deterministic: every rank computes the same
    answer from the same inputs.  The owner is a rank with the largest overlap with the piece (so that as
    few bytes as possible cross NVLink); among (near-)equal candidates the one that has been given the
    fewest bytes so far (so that storage reads are spread over the ranks)."""
    load = [0] * len(boxes_by_rank)
    owners: List[Optional[int]] = []
    for piece in pieces:
        elsize = dtype_to_element_size(string_to_dtype(piece.dtype))
        need = [
            sum(_overlap_bytes(piece, box, elsize) for box in boxes.get(piece.path, [])) for boxes in boxes_by_rank
        ]
        best = max(need)
        if best == 0:
            owners.append(None)
            continue
        candidates = [r for r, n in enumerate(need) if n * 1000 >= best * 999]
        owner = min(candidates, key=lambda r: (load[r], r))
        load[owner] += piece.nbytes
        owners.append(owner)
    return owners


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
        self.read_bytes = 0  # bytes this rank read from storage (it owns these pieces)
        self.local_bytes = 0  # bytes copied from this rank's own staging
        self.peer_bytes = 0  # bytes this rank pulled over NVLink
        self.copy_ms = 0.0  # device time of the copy kernels (CUDA events)
        self.launches = 0
        self.skipped: List[str] = []  # paths some rank did not offer: load them through the ordinary path


last_stats: Optional[ReshardStats] = None


def p2p_reshard_restore(
    items: Sequence[Tuple[str, ShardedTensorEntry, Any]],
    storage: StoragePlugin,
    pg: Any,
    event_loop: asyncio.AbstractEventLoop,
    device: int,
) -> ReshardStats:
    """Restore `items` = (logical path, merged sharded entry, runtime object) on every rank of `pg`.
    Collective: every rank of `pg` must call it (possibly with an empty list).  Only the paths that every
    rank passed are handled; `stats.skipped` lists the others so the caller can load them the ordinary way."""
    global last_stats
    from .scheduler import _Allocator, _read_into_pinned  # shared read machinery

    stats = ReshardStats()
    rank, world = pg.get_rank(), pg.get_world_size()

    # ---- 1. everybody's local boxes (metadata all-gather), then the plan ------------------------------
    my_boxes: Dict[str, List[Box]] = {}
    my_tensors: Dict[str, List[torch.Tensor]] = {}
    for path, _entry, obj in items:
        locals_ = ShardedTensorIOPreparer.local_boxes(obj)
        my_boxes[path] = [(list(o), list(s)) for _t, o, s in locals_]
        my_tensors[path] = [t for t, _o, _s in locals_]
    gathered: List[Any] = [None] * world
    pg.all_gather_object(gathered, my_boxes)
    common = set.intersection(*(set(g.keys()) for g in gathered))
    stats.skipped = [path for path, _e, _o in items if path not in common]
    items = [it for it in items if it[0] in common]
    pieces: List[Piece] = []
    for path, entry, _obj in items:
        pieces += pieces_of(path, entry)
    if not pieces:
        return stats
    last_stats = stats  # only calls that had something to do are remembered
    eng = get_engine(device)
    ctx = eng.ctx
    owners = plan_owners(pieces, gathered)

    # ---- 2. owners read their pieces into device staging ------------------------------------------------
    mine = [i for i, o in enumerate(owners) if o == rank]
    staged: Dict[int, int] = {}  # piece index -> device address

    async def read_all() -> None:
        loop = asyncio.get_running_loop()
        executor = ThreadPoolExecutor(max_workers=4, thread_name_prefix="tsb-reshard")
        alloc = _Allocator()
        sem = asyncio.Semaphore(32)

        async def one(i: int) -> None:
            piece = pieces[i]
            async with sem:
                dptr = await alloc.device(eng, piece.nbytes)
                staged[i] = dptr
                pbuf = await alloc.pinned(eng, piece.nbytes)
                try:
                    rr = ReadReq(path=piece.location, buffer_consumer=None, byte_range=piece.byte_range)  # type: ignore
                    await _read_into_pinned(loop, executor, storage, rr, pbuf, piece.nbytes)
                    ctx.memcpy_h2d(STREAM_H2D, dptr, pbuf.ptr, piece.nbytes)
                    ev = ctx.event_record(STREAM_H2D)
                    try:
                        await loop.run_in_executor(executor, ctx.event_sync, ev)
                    finally:
                        ctx.event_release(ev)
                finally:
                    pbuf.release()
                    alloc.notify()
            stats.read_bytes += piece.nbytes

        try:
            await asyncio.gather(*(one(i) for i in mine))
        finally:
            executor.shutdown(wait=False)

    try:
        event_loop.run_until_complete(read_all())

        # ---- 3. publish where every piece lives (IPC handle of its arena segment + offset) --------------
        exports = {i: ctx.ipc_export(dptr) for i, dptr in staged.items()}
        all_exports: List[Any] = [None] * world
        pg.all_gather_object(all_exports, exports)  # doubles as "all pieces are resident"

        # ---- 4. one launch: own staging + peer staging -> local shards -----------------------------------
        opened: Dict[Tuple[int, bytes], int] = {}
        regions: List[RegionTuple] = []
        try:
            for i, piece in enumerate(pieces):
                owner = owners[i]
                if owner is None:
                    continue
                elsize = dtype_to_element_size(string_to_dtype(piece.dtype))
                src_strides = contiguous_strides(piece.sizes, elsize)
                base: Optional[int] = None
                for tensor, box in zip(my_tensors[piece.path], my_boxes[piece.path]):
                    region = overlap_wrt_saved(piece.offsets, piece.sizes, box[0], box[1])
                    if region is None:
                        continue
                    if base is None:
                        if owner == rank:
                            base = staged[i]
                        else:
                            handle, offset = all_exports[owner][i]
                            key = (owner, handle)
                            if key not in opened:
                                opened[key] = ctx.ipc_open(handle)
                            base = opened[key] + offset
                    dst = tensor.detach()
                    dst_strides = [s * elsize for s in dst.stride()]
                    src_addr, dst_addr, lens = base, dst.data_ptr(), []
                    nbytes = elsize
                    for dim, so, do, length in region:
                        src_addr += so * src_strides[dim]
                        dst_addr += do * dst_strides[dim]
                        lens.append(length)
                        nbytes *= length
                    regions += box_regions(src_addr, src_strides, dst_addr, dst_strides, lens, elsize)
                    if owner == rank:
                        stats.local_bytes += nbytes
                    else:
                        stats.peer_bytes += nbytes
            arr = regions_to_array(regions)
            if arr.shape[0]:
                eng.wait_current_stream(STREAM_PACK)
                k0, k1 = eng.copy_region_array_timed(arr, STREAM_PACK)
                ctx.stream_sync(STREAM_PACK)
                stats.copy_ms += ctx.event_elapsed_ms(k0, k1)
                stats.launches += 1
                ctx.event_release(k0)
                ctx.event_release(k1)
            # ---- 5. nobody may free its staging while a peer still reads it ------------------------------
            pg.barrier()
        finally:
            for base_addr in opened.values():
                ctx.ipc_close(base_addr)
    finally:
        for dptr in staged.values():
            ctx.dev_free(dptr)
    return stats
