"""Execution of write / read requests: device staging, pinned host ring, overlapped storage I/O.

Entry points keep the reference's names and signatures (torchsnapshot/scheduler.py:220-461):
`execute_write_reqs` returns a `PendingIOWork` as soon as every request is *staged* (a consistent
copy exists), `execute_read_reqs` returns when every destination tensor is filled.

What is different underneath:

* take -- all device-stageable requests of a rank (GPU slabs and large CUDA tensors) are packed
  into ONE device staging block by ONE gather-pack launch per wave; "staged" therefore means
  "packed in HBM", which costs a few milliseconds, not "copied to host" (the reference blocks the
  event loop on a pageable .cpu() per slab, batcher.py:141-156).  Each file's bytes then move to a
  pinned host buffer with one cudaMemcpyAsync on a dedicated copy stream and are handed to
  StoragePlugin.write as a memoryview while the next copy is already in flight.
* restore -- each read lands in pinned memory (directly, when the plugin offers `read_into`),
  crosses PCIe in one H2D copy, and is scattered into the destination tensors by ONE unpack launch
  per read request (the reference does one pageable H2D copy per tensor in a 4-thread pool,
  tensor.py:329-338).
* requests that cannot be staged on the device (CPU tensors, pickled objects, torch_save dtypes,
  user prepare funcs) go through their own stage_buffer / consume_buffer under the same memory
  budget rule as the reference.
"""

import asyncio
import logging
import os
import socket
import time
from collections import defaultdict
from concurrent.futures import Executor, ThreadPoolExecutor
from typing import Any, Dict, List, Optional, Set, Tuple

import numpy as np
import psutil
import torch

from . import _native, knobs
from ._native import ArenaFull, STREAM_D2H, STREAM_H2D, STREAM_PACK
from .engine import Engine, PinnedBuffer, get_engine, readonly_buffer_address
from .regions import regions_to_array
from .io_types import BufferType, ReadIO, ReadReq, StoragePlugin, WriteIO, WriteReq

logger: logging.Logger = logging.getLogger(__name__)

_MAX_PER_RANK_MEMORY_BUDGET_BYTES: int = 32 * 1024 * 1024 * 1024
_AVAILABLE_MEMORY_MULTIPLIER: float = 0.6
_MAX_PER_RANK_CPU_CONCURRENCY: int = int(os.environ.get("TSB200_CPU_CONCURRENCY", "4"))
# storage operations in flight per rank.  The reference uses 16 (scheduler.py:30); 32 single-threaded
# writers measured best on the B200 box (tmpfs: 28.8 vs 26.6 GB/s end-to-end take), TSB200_IO_CONCURRENCY overrides.
_MAX_PER_RANK_IO_CONCURRENCY: int = int(os.environ.get("TSB200_IO_CONCURRENCY", "32"))
_local_world_size: int = 1


def set_local_world_size(n: int) -> None:
    """Ranks sharing this host (Snapshot tells the scheduler): storage concurrency is per host, not per rank."""
    global _local_world_size
    _local_world_size = max(int(n), 1)


def io_concurrency() -> int:
    """Storage operations in flight for this rank: the knob if set, else 2 per host core available to this
    rank, within [4, 32] (8 ranks on a 32-core host otherwise run 256 reader threads and thrash)."""
    if "TSB200_IO_CONCURRENCY" in os.environ:
        return _MAX_PER_RANK_IO_CONCURRENCY
    cores = os.cpu_count() or 8
    return max(4, min(_MAX_PER_RANK_IO_CONCURRENCY, 2 * cores // _local_world_size))


def streaming_batch_len(sizes: List[int], stream_batch: int) -> int:
    """How many of the files whose sizes are given (in order) form the next interleaved batch of a streamed take:
    `stream_batch` slab-sized files, large files counting half (see drain_wave); at least one."""
    room, n = 2 * stream_batch, 0
    for nbytes in sizes:
        cost = 1 if nbytes >= _LARGE_FILE_BYTES else 2
        if n and cost > room:
            break
        room -= cost
        n += 1
    return max(n, 1) if sizes else 0


def read_concurrency() -> int:
    """Storage READS in flight for this rank.  A read's bytes are useless to the device until the whole
    request is in pinned memory, so the read side wants few requests in flight with several pread workers
    each (the first H2D starts after one request's worth of reading, not after `io_concurrency()` requests'
    worth).  Measured on the 9.35 GB DDP state (scripts/restore_concurrency_sweep.py): 32 in flight 34.2 GB/s,
    3-16 in flight 36.5-38.2 GB/s, 2 in flight 29-37 GB/s.  TSB200_READ_CONCURRENCY overrides."""
    if "TSB200_READ_CONCURRENCY" in os.environ:
        return max(int(os.environ["TSB200_READ_CONCURRENCY"]), 1)
    return min(io_concurrency(), 6)


_LARGE_FILE_BYTES = 256 << 20  # files from here on count half in a streaming batch (drain_wave)
_TRACE_READS = os.environ.get("TSB200_TRACE_READS") == "1"
_TRACE_WRITES = os.environ.get("TSB200_TRACE_WRITES") == "1"
# streamed writes: files of a batch cross the link interleaved in chunks of this size and are written as they
# arrive (plugins with `write_streamed`); TSB200_STREAM_WRITES=0 hands every plugin whole buffers
_STREAM_WRITES = os.environ.get("TSB200_STREAM_WRITES", "1").lower() not in ("0", "false")
_STREAM_CHUNK = int(os.environ.get("TSB200_D2H_CHUNK_BYTES", str(32 << 20)))
_STAGING_ALIGN = 512  # each file's bytes start 512-B aligned inside the device staging block


def get_local_world_size(pg: Any) -> int:
    hostname = socket.gethostname()
    obj_list = [None] * pg.get_world_size()
    pg.all_gather_object(obj_list=obj_list, obj=hostname)
    return sum(1 for h in obj_list if h == hostname)


def get_process_memory_budget_bytes(pg: Any = None) -> int:
    """Same rule as the reference (scheduler.py:45-65): env override, else
    min(0.6 * available / local_world_size, 32 GiB)."""
    if "TORCHSNAPSHOT_PER_RANK_MEMORY_BUDGET_BYTES" in os.environ:
        try:
            return int(os.environ["TORCHSNAPSHOT_PER_RANK_MEMORY_BUDGET_BYTES"])
        except Exception as e:
            logger.warning(f"Failed to override memory budget: {e}.")
    available = int(psutil.virtual_memory().available * _AVAILABLE_MEMORY_MULTIPLIER)
    local_world_size = get_local_world_size(pg) if pg is not None and pg.get_world_size() > 1 else 1
    return min(available // local_world_size, _MAX_PER_RANK_MEMORY_BUDGET_BYTES)


def _is_device_stageable(stager: Any) -> bool:
    fn = getattr(stager, "is_device_stageable", None)
    return bool(fn()) if fn is not None else False


def _is_device_consumable(consumer: Any) -> bool:
    fn = getattr(consumer, "is_device_consumable", None)
    return bool(fn()) if fn is not None else False


def _stager_device(stager: Any) -> int:
    if hasattr(stager, "device"):
        return int(stager.device)
    return stager.tensor.device.index or 0


def _consumer_device(consumer: Any) -> int:
    from .batcher import _consumer_device as f

    return f(consumer)


def _align(n: int, a: int = _STAGING_ALIGN) -> int:
    return (n + a - 1) // a * a


class _Stats:
    def __init__(self) -> None:
        self.begin = time.monotonic()
        self.bytes = 0
        self.staging_done: Optional[float] = None
        self.pack_launched: Optional[float] = None
        self.finished = False  # PendingIOWork.complete() has run (successfully or not)
        self.first_pack_launched: Optional[float] = None
        self.speed_context = ""  # kind of sink, for the partitioner's throughput history
        self.d2h_done: Optional[float] = None
        self.writes_done: Optional[float] = None
        # device-timed bracket of the take: first pack launch -> last D2H complete
        self.pack_events: List[Tuple[Engine, int, int, int]] = []  # (engine, start, stop, kernel bytes)
        self.d2h_end_events: List[Tuple[Engine, int]] = []
        # events around the copy kernels only (descriptor upload / tile expansion excluded)
        self.kernel_events: List[Tuple[Engine, int, int, int]] = []
        # requests packed straight into pinned host memory (no device staging available): kernel events, bytes
        self.host_pack_events: List[Tuple[Engine, int, int, int]] = []
        self.host_packed_bytes = 0
        self.write_trace: List[Tuple[float, float, float, int]] = []  # per file (TSB200_TRACE_WRITES=1): ready, start, end, bytes

    def kernel_ms(self) -> float:
        return sum(eng.ctx.event_elapsed_ms(a, b) for eng, a, b, _n in self.kernel_events)

    def kernel_bytes(self) -> int:
        """Algorithmic HBM bytes of the pack launches: 2 x packed payload (read + write)."""
        return sum(n for _e, _a, _b, n in self.kernel_events)

    def pack_ms(self) -> float:
        """Device time of the pack launches (CUDA events on the pack stream)."""
        return sum(eng.ctx.event_elapsed_ms(a, b) for eng, a, b, _n in self.pack_events)

    def pack_kernel_bytes(self) -> int:
        return sum(n for _e, _a, _b, n in self.pack_events)

    def device_ms(self) -> float:
        """First pack launch -> last D2H complete, per device, max over devices."""
        worst = 0.0
        for eng, start, _stop, _n in self.pack_events[:1]:
            for eng2, end in self.d2h_end_events:
                if eng2 is eng:
                    worst = max(worst, eng.ctx.event_elapsed_ms(start, end))
        return worst

    def release(self) -> None:
        for eng, a, b, _n in self.pack_events:
            eng.ctx.event_release(a)
            eng.ctx.event_release(b)
        for eng, e in self.d2h_end_events:
            eng.ctx.event_release(e)
        for eng, a, b, _n in self.kernel_events + self.host_pack_events:
            eng.ctx.event_release(a)
            eng.ctx.event_release(b)
        self.pack_events = []
        self.d2h_end_events = []
        self.kernel_events = []
        self.host_pack_events = []


last_write_stats: Optional[_Stats] = None  # introspection hook for bench.py
_unreleased_stats: List[_Stats] = []  # takes that were still draining when a later take began


class _ReadStats:
    def __init__(self) -> None:
        self.begin = time.monotonic()
        self.bytes = 0
        self.device_bytes = 0  # bytes that went pinned -> H2D -> unpack kernel
        self.unpack_ms = 0.0  # device time of the unpack kernels (CUDA events)
        self.unpack_launches = 0
        self.direct_bytes = 0  # dense destinations filled by the H2D copy itself (no staging, no unpack pass)
        self.host_unpacked_bytes = 0  # unpacked straight out of pinned host memory (no device staging available)
        self.trace: List[Tuple[float, float, float, float, float, int]] = []  # per device read (TSB200_TRACE_READS=1)


last_read_stats: Optional[_ReadStats] = None


async def _wait_event(loop: asyncio.AbstractEventLoop, executor: Optional[Executor], eng: Engine, ev: int) -> None:
    await loop.run_in_executor(executor, eng.ctx.event_sync, ev)


class _Allocator:
    """Arena allocation that waits (asynchronously) for a release instead of failing, and keeps the pinned
    bytes in flight under the per-rank memory budget (the reference budgets its host staging buffers the
    same way, scheduler.py:255-329: a request larger than the budget is admitted only when nothing else is
    in flight)."""

    def __init__(self, budget_bytes: Optional[int] = None) -> None:
        self.released = asyncio.Event()
        self.outstanding = 0  # allocations handed out through this object and not yet released
        self.budget = budget_bytes
        self.pinned_inflight = 0

    def notify(self, pinned_bytes: int = 0) -> None:
        self.outstanding -= 1
        self.pinned_inflight -= pinned_bytes
        self.released.set()

    async def pinned(self, eng: Engine, nbytes: int) -> PinnedBuffer:
        while True:
            over = self.budget is not None and self.pinned_inflight > 0 and self.pinned_inflight + nbytes > self.budget
            if not over:
                try:
                    try:
                        buf = eng.pinned(nbytes, False)  # out of the segments the arena already holds
                    except ArenaFull:
                        # Pinning new memory costs ~40 ms per GiB: never on the event-loop thread, where it
                        # would hold up everything else -- the consistency point of an async_take included.
                        buf = await asyncio.get_running_loop().run_in_executor(None, eng.pinned, nbytes)
                    self.outstanding += 1
                    self.pinned_inflight += nbytes
                    return buf
                except ArenaFull:
                    pass
            await self._wait(eng, "pinned host")

    async def pinned_now(self, eng: Engine, nbytes: int) -> Optional[PinnedBuffer]:
        """A pinned block if one can be had without waiting for a release (within the budget; out of the segments
        the arena holds, or out of a new one pinned from a worker thread while the arena is below its bound),
        else None."""
        if self.budget is not None and self.pinned_inflight > 0 and self.pinned_inflight + nbytes > self.budget:
            return None
        try:
            try:
                buf = eng.pinned(nbytes, False)
            except ArenaFull:
                buf = await asyncio.get_running_loop().run_in_executor(None, eng.pinned, nbytes)
        except ArenaFull:
            return None
        self.outstanding += 1
        self.pinned_inflight += nbytes
        return buf

    async def device(self, eng: Engine, nbytes: int) -> int:
        while True:
            try:
                ptr = eng.ctx.dev_alloc(max(nbytes, 1))
                self.outstanding += 1
                return ptr
            except ArenaFull:
                await self._wait(eng, "device staging")

    async def _wait(self, eng: Engine, what: str) -> None:
        if self.outstanding == 0:
            # nothing of ours will ever be released: another take/restore holds the arena
            await asyncio.sleep(0.002)
            return
        self.released.clear()
        try:
            await asyncio.wait_for(self.released.wait(), timeout=0.05)
        except asyncio.TimeoutError:
            pass


# ------------------------------------------------------------------------------------------------
# write side
# ------------------------------------------------------------------------------------------------
class PendingIOWork:
    """Work left after staging: D2H copies of staged device blocks and storage writes
    (reference scheduler.py:178-217)."""

    def __init__(self, tasks: Set["asyncio.Future[Any]"], executor: ThreadPoolExecutor, stats: _Stats, rank: int) -> None:
        self.tasks = tasks
        self.executor = executor
        self.stats = stats
        self.rank = rank

    async def complete(self) -> None:
        try:
            while self.tasks:
                done, _ = await asyncio.wait(self.tasks, return_when=asyncio.FIRST_COMPLETED)
                for d in done:
                    self.tasks.discard(d)
                    d.result()  # surface the first failure
        except BaseException:
            for t in self.tasks:
                t.cancel()
            await asyncio.gather(*self.tasks, return_exceptions=True)
            self.tasks.clear()
            raise
        finally:
            self.executor.shutdown(wait=False)
            self.stats.finished = True  # its event handles may be released by the next take
        self.stats.writes_done = time.monotonic()
        if knobs.trim_arenas_after_use():
            for eng in {e for e, _a, _b, _n in self.stats.pack_events}:
                eng.ctx.arena_trim()
        try:  # feed the replicated-write partitioner with this rank's measured throughput into this kind of sink
            from .partitioner import record_take_speed

            if self.stats.first_pack_launched is not None:
                record_take_speed(
                    self.stats.pack_kernel_bytes() // 2 + self.stats.host_packed_bytes,
                    1e3 * (self.stats.writes_done - self.stats.first_pack_launched), self.stats.speed_context,
                )
        except Exception:  # pragma: no cover - statistics only
            pass
        elapsed = max(self.stats.writes_done - self.stats.begin, 1e-9)
        logger.info(
            f"Rank {self.rank} completed writing in {elapsed:.3f} s "
            f"({self.stats.bytes / 1024**2 / elapsed:.1f} MB/s)"
        )

    def sync_complete(self, event_loop: asyncio.AbstractEventLoop) -> None:
        event_loop.run_until_complete(self.complete())


async def execute_write_reqs(
    write_reqs: List[WriteReq],
    storage: StoragePlugin,
    memory_budget_bytes: int,
    rank: int,
) -> PendingIOWork:
    global last_write_stats
    loop = asyncio.get_running_loop()
    executor = ThreadPoolExecutor(max_workers=_MAX_PER_RANK_CPU_CONCURRENCY, thread_name_prefix="tsb-stage")
    # Event handles of earlier takes go back once those takes have drained: an async_take may still be copying in
    # its background thread (it waits on its pack event there) when the next take begins.
    for old in list(_unreleased_stats):
        if old.finished:
            _unreleased_stats.remove(old)
            if old is not last_write_stats:
                try:
                    old.release()
                except Exception:
                    pass
    if last_write_stats is not None and last_write_stats.finished:
        try:
            last_write_stats.release()
        except Exception:
            pass
    elif last_write_stats is not None:
        _unreleased_stats.append(last_write_stats)
    stats = _Stats()
    last_write_stats = stats
    try:
        from .partitioner import speed_context

        stats.speed_context = speed_context()
    except Exception:  # pragma: no cover
        pass
    tasks: Set["asyncio.Future[Any]"] = set()
    io_slots = asyncio.Semaphore(io_concurrency())
    stream_batch = int(os.environ.get("TSB200_STREAM_BATCH", 0)) or max(io_concurrency() // 2, 1)  # files interleaved on the link: half of the writers, so that the
    # next batch's first chunks find free writers while this batch's last ones are still being written
    alloc = _Allocator(memory_budget_bytes)

    device_reqs: Dict[int, List[WriteReq]] = defaultdict(list)
    host_reqs: List[WriteReq] = []
    for wr in write_reqs:
        if _is_device_stageable(wr.buffer_stager):
            device_reqs[_stager_device(wr.buffer_stager)].append(wr)
        else:
            host_reqs.append(wr)

    def spawn(coro: Any) -> "asyncio.Future[Any]":
        t = asyncio.ensure_future(coro)
        tasks.add(t)
        return t

    async def write_out(path: str, buf: BufferType, on_done: Any = None) -> None:
        try:
            t_ready = time.monotonic()
            async with io_slots:
                t_start = time.monotonic()
                await storage.write(WriteIO(path=path, buf=buf))
            stats.bytes += len(buf)
            if _TRACE_WRITES:
                stats.write_trace.append((t_ready - stats.begin, t_start - stats.begin, time.monotonic() - stats.begin, len(buf)))
        finally:
            if isinstance(buf, memoryview):
                try:
                    buf.release()
                except BufferError:
                    pass
            del buf
            if on_done is not None:
                on_done()

    async def finish_file(eng: Engine, ev: int, pbuf: PinnedBuffer, path: str) -> None:
        try:
            await _wait_event(loop, executor, eng, ev)
        finally:
            eng.ctx.event_release(ev)
        view = pbuf.view()
        nbytes = pbuf.nbytes
        del pbuf
        await write_out(path, view, lambda: alloc.notify(nbytes))

    async def stream_file(eng: Engine, feed: Tuple[Any, Any], pbuf: PinnedBuffer, path: str) -> None:
        """One file of a streamed batch: the plugin's writer thread follows the D2H copies chunk by chunk.
        `feed` = (event handles, end offsets): event i done => bytes [0, ends[i]) are in pinned memory."""
        ctx = eng.ctx
        nbytes = pbuf.nbytes
        view = pbuf.view()
        del pbuf
        events, ends = [int(e) for e in feed[0]], [int(x) for x in feed[1]]

        def wait_more(done: int) -> int:
            """Called on the plugin's writer thread: blocks (GIL released) until bytes beyond `done` are in pinned
            memory, returns how many are."""
            upto = done
            for ev, end in zip(events, ends):
                if end <= done:
                    continue
                if upto == done:
                    ctx.event_sync(ev)
                elif not ctx.event_done(ev):
                    break
                upto = end
            return upto

        # a plugin that writes through libtsb200 can follow the events natively, in one GIL-free call
        wait_more.native = (ctx, feed[0], feed[1])  # type: ignore[attr-defined]

        try:
            t_ready = time.monotonic()
            async with io_slots:
                t_start = time.monotonic()
                await storage.write_streamed(path, view, wait_more)  # type: ignore[attr-defined]
            stats.bytes += nbytes
            if _TRACE_WRITES:
                stats.write_trace.append((t_ready - stats.begin, t_start - stats.begin, time.monotonic() - stats.begin, nbytes))
        finally:
            for ev in events:
                ctx.event_release(ev)
            try:
                view.release()
            except BufferError:
                pass
            del view
            alloc.notify(nbytes)

    staged_gate = asyncio.Event()

    async def drain_wave(eng: Engine, base: int, files: List[Tuple[str, int, int]], ev_pack: int,
                         gate: Optional[asyncio.Event] = None) -> None:
        """D2H of every file of a packed wave, then release of the wave's staging block.

        With a plugin that can follow the copies (`write_streamed`, the FS plugin) the files of a batch cross the
        link INTERLEAVED, chunk by chunk, and each file's writer starts with its first chunk: a write of one
        file runs at ~3 GB/s whatever the thread count (the inode lock), so the storage rate is the number of
        files being written, and a file whose bytes arrive last otherwise finishes a whole file-write time after
        the link went idle (16 x 512 MiB on one GPU: 23.4 GB/s, last file in pinned memory at 180 ms, written
        at 365 ms; profiles/r02_big_piece_trace.txt).  Other plugins get whole buffers, one file after the other."""
        ctx = eng.ctx
        streamed = _STREAM_WRITES and getattr(storage, "write_streamed", None) is not None
        try:
            if gate is not None:
                await gate.wait()
            ctx.stream_wait_event(STREAM_D2H, ev_pack)
            i = 0
            while i < len(files):
                if not streamed:
                    path, off, nbytes = files[i]
                    pbuf = await alloc.pinned(eng, nbytes)
                    ctx.memcpy_d2h(STREAM_D2H, pbuf.ptr, base + off, nbytes)
                    ev = ctx.event_record(STREAM_D2H)
                    spawn(finish_file(eng, ev, pbuf, path))
                    del pbuf
                    i += 1
                    continue
                # a batch: half as many slab-sized files as there are writers (the next batch's first chunks then
                # find free writers), up to as many LARGE files as there are writers -- a large file keeps its one
                # writer busy for 150 ms and more, and files that start late end late (20 x 400 MiB: 26 GB/s in
                # batches of 16 + 4, 31.5 GB/s in one batch; the slab workload loses 10 % with batches of 24-32:
                # profiles/r02_stream_batch_sweep.txt) -- or as pinned memory holds right now
                batch: List[Tuple[str, int, int, PinnedBuffer]] = []
                stop = i + streaming_batch_len([f[2] for f in files[i:i + 2 * stream_batch]], stream_batch)
                while i < stop:
                    path, off, nbytes = files[i]
                    if not batch:
                        pbuf = await alloc.pinned(eng, nbytes)
                    else:
                        got = await alloc.pinned_now(eng, nbytes)
                        if got is None:
                            break
                        pbuf = got
                    batch.append((path, off, nbytes, pbuf))
                    i += 1
                # one native call issues the whole batch (hundreds of copies and event records from Python would
                # fight the writer threads for the GIL on the event-loop thread)
                try:
                    feeds = ctx.memcpy_d2h_interleaved(
                        STREAM_D2H, [b[3].ptr for b in batch], [base + b[1] for b in batch], [b[2] for b in batch],
                        _STREAM_CHUNK,
                    )
                except BaseException:
                    for _path, _off, nbytes, pbuf in batch:  # nothing was handed to a writer: give the blocks back
                        pbuf.release()
                        alloc.notify(nbytes)
                    raise
                for b, (path, _off, _n, pbuf) in enumerate(batch):
                    spawn(stream_file(eng, feeds[b], pbuf, path))
                del batch, pbuf
                await asyncio.sleep(0)
            last_ev = ctx.event_record(STREAM_D2H)
            await _wait_event(loop, executor, eng, last_ev)
            stats.d2h_end_events.append((eng, last_ev))
            stats.d2h_done = time.monotonic()
        finally:
            ctx.dev_free(base)
            alloc.notify()

    async def pack_to_host(eng: Engine, wr: WriteReq, first: bool) -> int:
        """No device staging to be had for this request (the trainer owns the HBM, or the request is larger
        than the staging bound): the pack kernel stores straight into the pinned block over PCIe -- pack and
        D2H in one pass, no copy of the object in HBM.  Measured at 52.5 GB/s per GPU against 56.9 GB/s for the
        copy engine (profiles/r02_d2h_variants_n2.txt).  The price is the consistency point: the take blocks
        until these bytes have crossed the link, not just until they are packed.  The reference retreats to a
        CPU slab and per-tensor pageable copies here (batcher.py:145-149)."""
        ctx = eng.ctx
        n = wr.buffer_stager.staged_nbytes()
        pbuf = await alloc.pinned(eng, n)
        if first:
            eng.wait_current_stream(STREAM_PACK)
        ev_start = ctx.event_record(STREAM_PACK)
        arr = regions_to_array(wr.buffer_stager.pack_regions(pbuf.ptr))
        if arr.shape[0]:
            k0, k1 = eng.copy_region_array_timed(arr, STREAM_PACK)
            stats.host_pack_events.append((eng, k0, k1, n))
        ev_pack = ctx.event_record(STREAM_PACK)
        stats.pack_events.append((eng, ev_start, ev_pack, 0))
        stats.d2h_end_events.append((eng, ctx.event_record(STREAM_PACK)))
        stats.pack_launched = time.monotonic()
        if stats.first_pack_launched is None:
            stats.first_pack_launched = stats.pack_launched
        stats.host_packed_bytes += n
        spawn(finish_file(eng, ctx.event_record(STREAM_PACK), pbuf, wr.path))
        return ev_pack

    try:
        # ---- 1. device staging: one pack launch per wave -------------------------------------------
        pack_waits: List[Tuple[Engine, int]] = []
        for device, reqs in device_reqs.items():
            eng = get_engine(device)
            ctx = eng.ctx
            capacity = max(eng.dev_bound - (64 << 20), 1 << 20)  # leave room for descriptor scratch
            i = 0
            first = True
            while i < len(reqs):
                # greedy wave: as many requests as the staging bound holds (always at least one)
                files: List[Tuple[str, int, int]] = []
                wave_reqs: List[WriteReq] = []
                total = 0
                wave_start = i
                while i < len(reqs):
                    n = reqs[i].buffer_stager.staged_nbytes()
                    if wave_reqs and total + _align(n) > capacity:
                        break
                    files.append((reqs[i].path, total, n))
                    wave_reqs.append(reqs[i])
                    total += _align(n)
                    i += 1
                try:
                    base = await alloc.device(eng, total)
                except ArenaFull:
                    raise
                except _native.TsbError:
                    # The driver has no such block (the bound was taken from the free HBM at engine creation; the
                    # trainer may have filled it since): try half the wave, down to a single request, and pack that
                    # one straight into pinned host memory.
                    if len(wave_reqs) > 1:
                        capacity = max(total // 2, 1)
                        i = wave_start
                        continue
                    ev_host = await pack_to_host(eng, wave_reqs[0], first)
                    first = False
                    pack_waits.append((eng, ev_host))
                    await asyncio.sleep(0)
                    continue
                # region table of the wave: bulk-planned slabs hand over ready-made arrays
                parts = []
                regions: List[Any] = []
                for wr, (_p, off, _n) in zip(wave_reqs, files):
                    as_array = getattr(wr.buffer_stager, "pack_region_array", None)
                    if as_array is not None:
                        if regions:
                            parts.append(regions_to_array(regions))
                            regions = []
                        parts.append(as_array(base + off))
                    else:
                        regions += wr.buffer_stager.pack_regions(base + off)
                if regions or not parts:
                    parts.append(regions_to_array(regions))
                if first:
                    eng.wait_current_stream(STREAM_PACK)
                    first = False
                ev_start = ctx.event_record(STREAM_PACK)
                arr = parts[0] if len(parts) == 1 else np.concatenate(parts)
                if arr.shape[0]:
                    k0, k1 = eng.copy_region_array_timed(arr, STREAM_PACK)
                    stats.kernel_events.append((eng, k0, k1, 2 * sum(n for _p, _o, n in files)))
                ev_pack = ctx.event_record(STREAM_PACK)
                stats.pack_events.append((eng, ev_start, ev_pack, 2 * sum(n for _p, _o, n in files)))
                pack_waits.append((eng, ev_pack))
                stats.pack_launched = time.monotonic()
                if stats.first_pack_launched is None:
                    stats.first_pack_launched = stats.pack_launched
                if i < len(reqs):
                    spawn(drain_wave(eng, base, files, ev_pack))
                    # give the drain a chance to queue its D2H copies before the next wave is planned
                    await asyncio.sleep(0)
                else:
                    # The last wave of this device: its copies are queued once the caller has been told that
                    # the snapshot is consistent (the loop runs these tasks as soon as it is driven again: right
                    # away in take, on the background thread of async_take).  Queuing a few hundred chunk copies
                    # is milliseconds of event-loop time that async_take's caller would otherwise wait for.
                    spawn(drain_wave(eng, base, files, ev_pack, staged_gate))

        # ---- 2. host staging under the memory budget (reference scheduler.py:255-329) --------------
        budget = memory_budget_bytes
        budget_freed = asyncio.Event()
        inflight_host = 0

        def refund(n: int) -> Any:
            def cb() -> None:
                nonlocal budget, inflight_host
                budget += n
                inflight_host -= 1
                budget_freed.set()

            return cb

        for wr in host_reqs:
            cost = wr.buffer_stager.get_staging_cost_bytes()
            while inflight_host > 0 and cost >= budget:
                budget_freed.clear()
                await budget_freed.wait()
            budget -= cost
            inflight_host += 1
            buf = await wr.buffer_stager.stage_buffer(executor)
            # the staging cost can exceed the buffer size (2x for torch_save)
            budget += cost - len(buf)
            spawn(write_out(wr.path, buf, refund(len(buf))))
            del buf

        # ---- 3. the snapshot is consistent once every pack launch has completed ---------------------
        for eng, ev in pack_waits:
            await _wait_event(loop, executor, eng, ev)
        stats.staging_done = time.monotonic()
        staged_gate.set()
        logger.info(f"Rank {rank} completed staging in {stats.staging_done - stats.begin:.3f} s")
    except BaseException:
        # nothing is left dangling when planning or staging fails half way: queued copies and writes are
        # cancelled, the worker threads let go (reference: the exception leaves the event loop to the caller)
        staged_gate.set()
        stats.finished = True
        for t in tasks:
            t.cancel()
        await asyncio.gather(*tasks, return_exceptions=True)
        tasks.clear()
        executor.shutdown(wait=False)
        raise
    return PendingIOWork(tasks=tasks, executor=executor, stats=stats, rank=rank)


def sync_execute_write_reqs(
    write_reqs: List[WriteReq],
    storage: StoragePlugin,
    memory_budget_bytes: int,
    rank: int,
    event_loop: asyncio.AbstractEventLoop,
) -> PendingIOWork:
    return event_loop.run_until_complete(
        execute_write_reqs(write_reqs=write_reqs, storage=storage, memory_budget_bytes=memory_budget_bytes, rank=rank)
    )


# ------------------------------------------------------------------------------------------------
# read side
# ------------------------------------------------------------------------------------------------
async def _read_into_pinned(
    loop: asyncio.AbstractEventLoop,
    executor: Optional[Executor],
    storage: StoragePlugin,
    rr: ReadReq,
    pbuf: PinnedBuffer,
    nbytes: int,
) -> None:
    read_into = getattr(storage, "read_into", None)
    if read_into is not None:
        # plugin fast path: bytes go from the file straight into pinned memory
        await read_into(rr.path, rr.byte_range, pbuf.view())
        return
    read_io = ReadIO(path=rr.path, byte_range=rr.byte_range)
    await storage.read(read_io=read_io)
    src = read_io.buf.getbuffer()
    if len(src) != nbytes:
        raise AssertionError(f"{rr.path}: expected {nbytes} bytes, the storage plugin returned {len(src)}.")
    await loop.run_in_executor(executor, _native.host_memcpy, pbuf.ptr, readonly_buffer_address(src), nbytes, 0)


def _dense_host_destination(rr: ReadReq) -> Optional[Any]:
    """The CPU tensor a read request fills when a raw byte copy of the stored range is the whole job (plain
    tensor consumer, buffer-protocol entry, dense destination of the saved dtype and size); else None."""
    from .batcher import _host_destination_of

    consumer = rr.buffer_consumer
    entry = getattr(consumer, "entry", None)
    if entry is None or not hasattr(consumer, "tensor") or not hasattr(entry, "shape"):
        return None
    try:
        nbytes = consumer.consumed_nbytes()
    except Exception:
        return None
    if rr.byte_range is not None and rr.byte_range[1] - rr.byte_range[0] != nbytes:
        return None
    dst = _host_destination_of(consumer, nbytes)
    return dst.reshape(-1).view(torch.uint8) if dst is not None else None


async def execute_read_reqs(
    read_reqs: List[ReadReq],
    storage: StoragePlugin,
    memory_budget_bytes: int,
    rank: int,
) -> None:
    loop = asyncio.get_running_loop()
    executor = ThreadPoolExecutor(max_workers=_MAX_PER_RANK_CPU_CONCURRENCY, thread_name_prefix="tsb-consume")
    io_slots = asyncio.Semaphore(read_concurrency())
    alloc = _Allocator(memory_budget_bytes)
    global last_read_stats
    rstats = _ReadStats()
    last_read_stats = rstats
    begin = time.monotonic()
    bytes_read = 0
    ordered: Set[Any] = set()

    async def device_read(rr: ReadReq) -> None:
        nonlocal bytes_read
        consumer = rr.buffer_consumer
        eng = get_engine(_consumer_device(consumer))
        ctx = eng.ctx
        if eng.device not in ordered:
            eng.wait_current_stream(STREAM_PACK)  # destination tensors may still be being written
            ordered.add(eng.device)
        nbytes = rr.byte_range[1] - rr.byte_range[0] if rr.byte_range is not None else consumer.consumed_nbytes()
        if nbytes == 0:
            return
        t_q = time.monotonic()
        # the I/O slot covers the storage read only: the device phase (H2D + unpack) of one request must
        # not keep the next read from starting (it is bounded by the pinned / device arenas instead)
        await io_slots.acquire()
        slot_held = True
        try:
            t_s = time.monotonic()
            pbuf = await alloc.pinned(eng, nbytes)
            try:
                t_p = time.monotonic()
                await _read_into_pinned(loop, executor, storage, rr, pbuf, nbytes)
                io_slots.release()
                slot_held = False
                t_r = time.monotonic()
                # (a bulk-planned slab consumer is never a single dense destination worth probing for)
                probe = consumer.unpack_regions(0) if not hasattr(consumer, "unpack_region_array") else []
                if len(probe) == 1 and len(probe[0][3]) == 0 and probe[0][2] == nbytes and probe[0][0] == 0:
                    # one dense destination (a large tensor, a chunk): the H2D copy writes the tensor itself -- no
                    # staging block, no unpack pass (mirror of the dense case of stage_device_buffer)
                    if (eng.device, "h2d") not in ordered:
                        eng.wait_current_stream(STREAM_H2D)
                        ordered.add((eng.device, "h2d"))
                    ctx.memcpy_h2d(STREAM_H2D, probe[0][1], pbuf.ptr, nbytes)
                    ev = ctx.event_record(STREAM_H2D)
                    try:
                        await _wait_event(loop, executor, eng, ev)
                    finally:
                        ctx.event_release(ev)
                    rstats.direct_bytes += nbytes
                else:
                    try:
                        dptr = await alloc.device(eng, nbytes)
                    except ArenaFull:
                        raise
                    except _native.TsbError:
                        # no device staging to be had (see pack_to_host on the write side): the unpack kernel
                        # reads the pinned block over PCIe instead of a staged copy in HBM
                        dptr = None
                    try:
                        if dptr is not None:
                            ctx.memcpy_h2d(STREAM_H2D, dptr, pbuf.ptr, nbytes)
                            ev_h2d = ctx.event_record(STREAM_H2D)
                            ctx.stream_wait_event(STREAM_PACK, ev_h2d)
                            ctx.event_release(ev_h2d)
                        else:
                            rstats.host_unpacked_bytes += nbytes
                        src_addr = dptr if dptr is not None else pbuf.ptr
                        as_array = getattr(consumer, "unpack_region_array", None)
                        arr = as_array(src_addr) if as_array is not None else regions_to_array(consumer.unpack_regions(src_addr))
                        k0 = k1 = None
                        if arr.shape[0]:
                            k0, k1 = eng.copy_region_array_timed(arr, STREAM_PACK)
                        ev = ctx.event_record(STREAM_PACK)
                        try:
                            await _wait_event(loop, executor, eng, ev)
                            if k0 is not None and dptr is not None:
                                rstats.unpack_ms += ctx.event_elapsed_ms(k0, k1)
                                rstats.unpack_launches += 1
                        finally:
                            ctx.event_release(ev)
                            if k0 is not None:
                                ctx.event_release(k0)
                                ctx.event_release(k1)
                    finally:
                        if dptr is not None:
                            ctx.dev_free(dptr)
                            alloc.notify()
            finally:
                pbuf.release()
                alloc.notify(nbytes)
        finally:
            if slot_held:
                io_slots.release()
        bytes_read += nbytes
        rstats.device_bytes += nbytes
        if _TRACE_READS:
            rstats.trace.append((t_q - begin, t_s - begin, t_p - begin, t_r - begin, time.monotonic() - begin, nbytes))

    budget = memory_budget_bytes
    budget_freed = asyncio.Event()
    inflight_host = 0

    async def host_read(rr: ReadReq, cost: int) -> None:
        nonlocal bytes_read, budget, inflight_host
        try:
            async with io_slots:
                read_into = getattr(storage, "read_into", None)
                direct = _dense_host_destination(rr) if read_into is not None else None
                if direct is not None:
                    # one dense CPU tensor of the saved dtype and size: the file range IS its memory image,
                    # so the plugin reads straight into the tensor (no intermediate buffer at all)
                    await read_into(rr.path, rr.byte_range, memoryview(direct.numpy()))
                    bytes_read += direct.numel()
                    return
                if read_into is not None and rr.byte_range is not None:
                    # plugin fast path: the range lands in a pooled host buffer, no BytesIO / getvalue() copies
                    from .batcher import HostSlab

                    slab = HostSlab(rr.byte_range[1] - rr.byte_range[0])
                    buf = memoryview(slab)
                    del slab  # the view keeps the buffer alive; it returns to the pool with the last view
                    await read_into(rr.path, rr.byte_range, buf)
                else:
                    read_io = ReadIO(path=rr.path, byte_range=rr.byte_range)
                    await storage.read(read_io=read_io)
                    buf = read_io.buf.getvalue()
            await rr.buffer_consumer.consume_buffer(buf, executor)
            bytes_read += len(buf)
        finally:
            budget += cost
            inflight_host -= 1
            budget_freed.set()

    tasks: Set["asyncio.Future[Any]"] = set()
    try:
        for rr in read_reqs:
            if _is_device_consumable(rr.buffer_consumer):
                tasks.add(asyncio.ensure_future(device_read(rr)))
                continue
            cost = rr.buffer_consumer.get_consuming_cost_bytes()
            while inflight_host > 0 and cost >= budget:
                budget_freed.clear()
                await budget_freed.wait()
            budget -= cost
            inflight_host += 1
            tasks.add(asyncio.ensure_future(host_read(rr, cost)))
        while tasks:
            done, _ = await asyncio.wait(tasks, return_when=asyncio.FIRST_COMPLETED)
            for d in done:
                tasks.discard(d)
                d.result()
    except BaseException:
        for t in tasks:
            t.cancel()
        await asyncio.gather(*tasks, return_exceptions=True)
        raise
    finally:
        executor.shutdown(wait=False)
    rstats.bytes = bytes_read
    if knobs.trim_arenas_after_use():
        from .engine import _engines

        for eng in list(_engines.values()):
            eng.ctx.arena_trim()
    elapsed = max(time.monotonic() - begin, 1e-9)
    logger.info(f"Rank {rank} finished loading. Throughput: {bytes_read / 1024**2 / elapsed:.2f}MB/s")


def sync_execute_read_reqs(
    read_reqs: List[ReadReq],
    storage: StoragePlugin,
    memory_budget_bytes: int,
    rank: int,
    event_loop: asyncio.AbstractEventLoop,
) -> None:
    event_loop.run_until_complete(
        execute_read_reqs(read_reqs=read_reqs, storage=storage, memory_budget_bytes=memory_budget_bytes, rank=rank)
    )


# ------------------------------------------------------------------------------------------------
# single-request device paths: what stage_buffer / consume_buffer do when something other than
# the schedulers above drives them (the reference's scheduler in drop-in mode, or a test that
# fulfils read requests directly from write requests)
# ------------------------------------------------------------------------------------------------
async def _retry_alloc(fn: Any, *args: Any) -> Any:
    for _ in range(30000):
        try:
            return fn(*args)
        except ArenaFull:
            await asyncio.sleep(0.002)
    raise _native.TsbError("timed out waiting for staging memory")


async def stage_device_buffer(stager: Any, device: Optional[int], executor: Optional[Executor]) -> BufferType:
    eng = get_engine(device or 0)
    ctx = eng.ctx
    loop = asyncio.get_running_loop()
    nbytes = stager.staged_nbytes()
    if nbytes == 0:
        return memoryview(bytearray(0))
    pbuf = await _retry_alloc(eng.pinned, nbytes)
    probe = stager.pack_regions(0)
    if len(probe) == 1 and len(probe[0][3]) == 0 and probe[0][2] == nbytes:
        # one dense tensor: the D2H copy reads the tensor itself, no pack pass
        eng.wait_current_stream(STREAM_D2H)
        ctx.memcpy_d2h(STREAM_D2H, pbuf.ptr, probe[0][0], nbytes)
        dptr = None
    else:
        dptr = await _retry_alloc(ctx.dev_alloc, nbytes)
        eng.wait_current_stream(STREAM_PACK)
        eng.copy_regions(stager.pack_regions(dptr), STREAM_PACK)
        ev_pack = ctx.event_record(STREAM_PACK)
        ctx.stream_wait_event(STREAM_D2H, ev_pack)
        ctx.event_release(ev_pack)
        ctx.memcpy_d2h(STREAM_D2H, pbuf.ptr, dptr, nbytes)
    ev = ctx.event_record(STREAM_D2H)
    try:
        await loop.run_in_executor(executor, ctx.event_sync, ev)
    finally:
        ctx.event_release(ev)
        if dptr is not None:
            ctx.dev_free(dptr)
    view = pbuf.view()
    del pbuf
    return view


async def consume_device_buffer(consumer: Any, buf: BufferType, device: Optional[int], executor: Optional[Executor]) -> None:
    eng = get_engine(device or 0)
    ctx = eng.ctx
    loop = asyncio.get_running_loop()
    nbytes = len(buf)
    if nbytes == 0:
        return
    pbuf = await _retry_alloc(eng.pinned, nbytes)
    try:
        await loop.run_in_executor(executor, _native.host_memcpy, pbuf.ptr, readonly_buffer_address(buf), nbytes, 0)
        dptr = await _retry_alloc(ctx.dev_alloc, nbytes)
        try:
            eng.wait_current_stream(STREAM_H2D)
            ctx.memcpy_h2d(STREAM_H2D, dptr, pbuf.ptr, nbytes)
            ev_h2d = ctx.event_record(STREAM_H2D)
            ctx.stream_wait_event(STREAM_PACK, ev_h2d)
            ctx.event_release(ev_h2d)
            eng.copy_regions(consumer.unpack_regions(dptr), STREAM_PACK)
            ev = ctx.event_record(STREAM_PACK)
            try:
                await loop.run_in_executor(executor, ctx.event_sync, ev)
            finally:
                ctx.event_release(ev)
        finally:
            ctx.dev_free(dptr)
    finally:
        pbuf.release()
