"""Slab planner: batch small tensor writes into slabs, and ranged reads of one file into one read.

Same names, signatures and error behaviour as the reference (torchsnapshot/batcher.py:48-482); the
slab byte layout is the reference's exactly (members back to back in write-request order, no
padding, a new slab when `slab + tensor >= threshold`, batcher.py:291-316).  What changes is how a
GPU slab is filled: `GPUBatchedBufferStager` contributes one region per member to a single
multi-region gather-pack launch (csrc/copy_kernel.cuh) and its bytes reach pinned host memory in
one D2H copy, instead of one D2D `copy_` per tensor plus a pageable `.cpu()` (batcher.py:152-156).
"""

import asyncio
import itertools
import os
import threading
import uuid

import numpy as np
from collections import defaultdict
from concurrent.futures import Executor
from enum import Enum
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch

from . import _native, knobs
from .io_types import BufferConsumer, BufferStager, BufferType, ReadReq, WriteReq
from .regions import RegionTuple
from .serialization import Serializer


def is_tensor_stager(buffer_stager: Any) -> bool:
    """Duck-typed TensorBufferStager check, so reference-made stagers batch too."""
    return (
        hasattr(buffer_stager, "tensor")
        and hasattr(buffer_stager, "entry")
        and hasattr(buffer_stager, "_tensor_prepare_func")
    )


def is_batchable(buffer_stager: Any) -> bool:
    return (
        is_tensor_stager(buffer_stager)
        and buffer_stager.entry.serializer == Serializer.BUFFER_PROTOCOL.value
        and buffer_stager._tensor_prepare_func is None
    )


def _check_byte_ranges_contiguous(byte_ranges: Iterator[Tuple[int, int]]) -> int:
    end = next(byte_ranges)[1]
    for lo, hi in byte_ranges:
        if lo != end:
            raise AssertionError("The byte ranges are not consecutive.")
        end = hi
    return end


class HostSlab:
    """A host buffer for one CPU slab, taken from / returned to a small free list.

    The reference allocates `bytearray(slab_size)` per slab (batcher.py:64): zero-filled, and every page is
    faulted in while it is filled -- 0.7 GB/s on one thread, against 5 GB/s for the same copy into memory
    that is already mapped.  Buffers are handed out as PEP 688 exporters: `memoryview(slab)` is what the
    scheduler passes to StoragePlugin.write, and when the last view is dropped the buffer goes back to the
    free list (at most TSB200_HOST_SLAB_POOL_BYTES, default 1 GiB, are kept)."""

    _free: Dict[int, List[Any]] = {}
    _free_bytes = 0
    _lock = threading.Lock()

    def __init__(self, nbytes: int) -> None:
        import numpy as np

        self.nbytes = nbytes
        key = max(nbytes, 1)
        arr = None
        with HostSlab._lock:
            bucket = HostSlab._free.get(key)
            if bucket:
                arr = bucket.pop()
                HostSlab._free_bytes -= key
        self._arr = arr if arr is not None else np.empty(key, dtype=np.uint8)
        self.ptr: int = self._arr.ctypes.data

    def __buffer__(self, flags: int) -> memoryview:
        return memoryview(self._arr)[: self.nbytes]

    def __del__(self) -> None:
        arr, self._arr = getattr(self, "_arr", None), None
        if arr is None:
            return
        try:
            limit = int(os.environ.get("TSB200_HOST_SLAB_POOL_BYTES", str(1 << 30)))
            with HostSlab._lock:
                if HostSlab._free_bytes + arr.nbytes <= limit:
                    HostSlab._free.setdefault(arr.nbytes, []).append(arr)
                    HostSlab._free_bytes += arr.nbytes
        except Exception:  # interpreter shutdown
            pass


def _host_tensor_of(stager: Any) -> Optional[Any]:
    """The CPU tensor behind a plain, batchable tensor stager (None for anything else)."""
    try:
        if stager.entry.serializer != Serializer.BUFFER_PROTOCOL.value or stager._tensor_prepare_func is not None:
            return None
        tensor = stager.tensor
    except AttributeError:
        return None
    if not isinstance(tensor, torch.Tensor) or tensor.is_cuda or tensor.is_quantized or tensor.layout != torch.strided:
        return None
    return tensor


class BatchedBufferStager(BufferStager):
    """CPU slab: member buffers are staged by their own stagers and copied into one host slab
    (reference batcher.py:48-98).  Only CPU tensors come here."""

    def __init__(self, byte_range_to_buffer_stager: Dict[Tuple[int, int], BufferStager]) -> None:
        self.byte_range_to_buffer_stager = byte_range_to_buffer_stager
        self.slab_sz_bytes: int = _check_byte_ranges_contiguous(iter(byte_range_to_buffer_stager.keys()))

    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        tensors = [_host_tensor_of(st) for st in self.byte_range_to_buffer_stager.values()]
        if tensors and all(t is not None for t in tensors):
            return await self._stage_host_tensors(tensors, executor)
        slab = bytearray(self.slab_sz_bytes)
        for (lo, hi), stager in self.byte_range_to_buffer_stager.items():
            buf = await stager.stage_buffer(executor=executor)
            if len(buf) != hi - lo:
                raise AssertionError(
                    "The size of the buffer generated by the buffer stager does not match with the byte "
                    f"range associated with the buffer stager. Buffer size: {len(buf)}, byte range: {(lo, hi)}."
                )
            slab[lo:hi] = buf
        return memoryview(slab)

    async def _stage_host_tensors(self, tensors: List[Any], executor: Optional[Executor]) -> BufferType:
        """Every member is a plain CPU tensor: ONE native call copies them all into a pooled host buffer, on
        several threads and without the GIL (tsb_host_gather), off the event-loop thread so that the writes of
        earlier slabs keep being dispatched.  Same bytes as the member-by-member path: the contiguous
        row-major elements of each tensor at its byte range (serialization.py:162-188)."""
        slab = HostSlab(self.slab_sz_bytes)
        keep, dst, src, nbytes = [], [], [], []
        for (lo, hi), tensor in zip(self.byte_range_to_buffer_stager.keys(), tensors):
            t = tensor.detach()
            if not t.is_contiguous():
                t = t.contiguous()
            n = t.numel() * t.element_size()
            if n != hi - lo:
                raise AssertionError(
                    "The size of the buffer generated by the buffer stager does not match with the byte "
                    f"range associated with the buffer stager. Buffer size: {n}, byte range: {(lo, hi)}."
                )
            if n:
                keep.append(t)
                dst.append(slab.ptr + lo)
                src.append(t.data_ptr())
                nbytes.append(n)
        if dst:
            if executor is None:
                _native.host_gather(dst, src, nbytes)
            else:
                await asyncio.get_running_loop().run_in_executor(executor, _native.host_gather, dst, src, nbytes)
        del keep
        return memoryview(slab)

    def get_staging_cost_bytes(self) -> int:
        return sum(s.get_staging_cost_bytes() for s in self.byte_range_to_buffer_stager.values()) + self.slab_sz_bytes


class GPUBatchedBufferStager(BatchedBufferStager):
    """GPU slab: every member is a CUDA tensor whose bytes the pack kernel gathers.

    Accepts only batchable TensorBufferStagers over CUDA tensors, like the reference class of the
    same name (batcher.py:101-139); there is no CPU-slab fallback on allocation failure (the
    reference's is batcher.py:142-149) -- the scheduler waits for staging memory instead.
    """

    def __init__(self, byte_range_to_buffer_stager: Dict[Tuple[int, int], BufferStager]) -> None:
        super().__init__(byte_range_to_buffer_stager=byte_range_to_buffer_stager)
        device = None
        for stager in byte_range_to_buffer_stager.values():
            if not is_tensor_stager(stager):
                raise AssertionError(f"GPUBatchedBufferStager only supports TensorBufferStagers (got {type(stager)}).")
            if not is_batchable(stager):
                raise AssertionError(f"GPUBatchedBufferStager only supports batchable entries (got {stager.entry}).")
            if not stager.tensor.is_cuda:
                raise AssertionError("GPUBatchedBufferStager only supports GPU tensors.")
            if device is None:
                device = stager.tensor.device.index
            elif device != stager.tensor.device.index:
                raise AssertionError("GPUBatchedBufferStager members must live on one device.")
        self.device: int = device if device is not None else 0

    # -- device side ------------------------------------------------------------------------
    def is_device_stageable(self) -> bool:
        return True

    def staged_nbytes(self) -> int:
        return self.slab_sz_bytes

    def pack_regions(self, dst_addr: int) -> List[RegionTuple]:
        from .regions import pack_regions

        out: List[RegionTuple] = []
        for (lo, _hi), stager in self.byte_range_to_buffer_stager.items():
            out += pack_regions(stager.tensor, dst_addr + lo)
        return out

    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        from .scheduler import stage_device_buffer

        return await stage_device_buffer(self, self.device, executor)

    def get_staging_cost_bytes(self) -> int:
        return self.slab_sz_bytes


_new_object = object.__new__
_BUFFER_PROTOCOL_VALUE = Serializer.BUFFER_PROTOCOL.value


def _tensor_entry_class() -> Any:
    from .manifest import TensorEntry

    return TensorEntry


class GPUSlabStager(BufferStager):
    """A GPU slab planned in bulk (`plan_cuda_tensors`): the members are held as arrays of addresses, lengths and
    slab offsets, so the region table of the pack launch is built by numpy, not member by member.  Same bytes
    in the slab as GPUBatchedBufferStager, which the reference-shaped `batch_write_requests` still builds."""

    def __init__(self, device: int, tensors: List[Any], ptrs: Any, offsets: Any, nbytes: Any, contiguous: Any,
                 size: int) -> None:
        self.device = device
        self.tensors = tensors  # keeps the members alive until the slab has been packed
        self._ptrs, self._offsets, self._nbytes, self._contiguous = ptrs, offsets, nbytes, contiguous
        self.slab_sz_bytes = size
        self._dense = contiguous & (nbytes > 0)
        # regions of the non-contiguous members, destination relative to the slab's first byte (a table of its
        # own, so that packing needs no tensor object: only addresses and strides, read here once)
        from .regions import pack_regions, regions_to_array

        regs: List[RegionTuple] = []
        for i in np.flatnonzero(~contiguous):
            regs += pack_regions(tensors[int(i)], int(offsets[i]))
        self._strided = regions_to_array(regs) if regs else None
        self._template: Any = None

    def is_device_stageable(self) -> bool:
        return True

    def staged_nbytes(self) -> int:
        return self.slab_sz_bytes

    def pack_region_array(self, dst_addr: int) -> Any:
        """The slab's pack regions as a REGION_DTYPE array (members back to back at dst_addr)."""
        if self._template is None:  # built once per plan: destinations relative to the slab's first byte
            dense = self._dense
            arr = np.zeros(int(dense.sum()), dtype=_native.REGION_DTYPE)
            arr["src"] = self._ptrs[dense]
            arr["dst"] = self._offsets[dense].astype(np.uint64)
            arr["run_bytes"] = self._nbytes[dense].astype(np.uint64)
            if self._strided is not None:
                arr = np.concatenate([arr, self._strided])
            self._template = arr
        arr = self._template.copy()
        arr["dst"] += np.uint64(dst_addr)
        return arr

    def pack_regions(self, dst_addr: int) -> List[RegionTuple]:
        out: List[RegionTuple] = []
        for rec in self.pack_region_array(dst_addr):
            nd = int(rec["ndim"])
            out.append((int(rec["src"]), int(rec["dst"]), int(rec["run_bytes"]), tuple(int(x) for x in rec["size"][:nd]),
                        tuple(int(x) for x in rec["src_stride"][:nd]), tuple(int(x) for x in rec["dst_stride"][:nd])))
        return out

    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        from .scheduler import stage_device_buffer

        return await stage_device_buffer(self, self.device, executor)

    def get_staging_cost_bytes(self) -> int:
        return self.slab_sz_bytes


def plan_cuda_tensors(
    paths: List[str], tensors: List[Any], storage_paths: List[str], replicated: List[bool],
    slab_size_threshold_bytes: Optional[int] = None, is_async_snapshot: bool = False,
    max_chunk_size_bytes: Optional[int] = None, table: Optional[Any] = None,
) -> Tuple[Dict[str, Any], List[WriteReq]]:
    """Entries and write requests of plain CUDA tensors of buffer-protocol dtypes, planned in bulk.  `table` is
    tensor_table.scan(tensors) when the caller already has it.

    The result is what TensorIOPreparer.prepare_write per tensor followed by batch_write_requests gives --
    same TensorEntry fields, same slab membership and byte ranges (members back to back in the order given, a
    new slab when `slab + tensor >= threshold`, reference batcher.py:291-316), tensors of `threshold` bytes or
    more as requests of their own -- without a stager, a write request, a relocation record and a region tuple
    per member: 13 k tensors plan in about a fifth of the time.  Callers (snapshot.py) use it only when every
    member of the GPU slabs is planned here, so that the slabs are the ones the reference would form."""
    import numpy as np

    from .io_preparer import TensorBufferStager, _DTYPE_STRINGS

    threshold = slab_size_threshold_bytes or knobs.get_slab_size_threshold_bytes()
    entry_cls = _tensor_entry_class()
    new, serializer = _new_object, _BUFFER_PROTOCOL_VALUE

    def _tensor_entry(location: str, dtype: str, shape: List[int], rep: bool, byte_range: Optional[List[int]]) -> Any:
        # TensorEntry without the dataclass constructor: this runs once per tensor of a 20 k-tensor state dict
        e = new(entry_cls)
        e.__dict__ = {"type": "Tensor", "location": location, "serializer": serializer, "dtype": dtype,
                      "shape": shape, "replicated": rep, "byte_range": byte_range}
        return e

    n = len(tensors)
    entries: Dict[str, Any] = {}
    write_reqs: List[WriteReq] = []
    if n == 0:
        return entries, write_reqs
    from . import tensor_table as tt

    if table is None:
        table = tt.scan(tensors)
    if not (table[:, tt.KIND] == tt.PLAIN).all():
        raise AssertionError("plan_cuda_tensors takes plain dense tensors only")
    nbytes = tt.nbytes_of(table)
    ptrs = tt.ptrs_of(table)
    contiguous = table[:, tt.CONTIGUOUS] != 0
    devices = table[:, tt.DEVICE]
    shapes = tt.shapes_of(table)
    code_strs = {c: _DTYPE_STRINGS[tt.dtype_of(c)] for c in np.unique(table[:, tt.DTYPE]).tolist()}
    dtype_strs = [code_strs[c] for c in table[:, tt.DTYPE].tolist()]
    small = nbytes < threshold
    # tensors of `threshold` bytes or more: a request each, in order (reference batcher.py:289-291); those above
    # the chunk size get no entry here -- the caller's per-leaf planner chunks them
    chunked = nbytes > max_chunk_size_bytes if max_chunk_size_bytes is not None else np.zeros(n, dtype=np.bool_)
    for i in np.flatnonzero(~small & ~chunked):
        i = int(i)
        e = _tensor_entry(storage_paths[i], dtype_strs[i], shapes[i], replicated[i], None)
        entries[paths[i]] = e
        write_reqs.append(WriteReq(storage_paths[i], TensorBufferStager(tensors[i], e, is_async_snapshot, None)))
    for dev in sorted(set(devices[small].tolist())):
        idx = np.flatnonzero(small & (devices == dev))
        sizes = nbytes[idx]
        cum = np.cumsum(sizes)
        start, m = 0, idx.size
        while start < m:
            before = int(cum[start - 1]) if start else 0
            # member k joins while (bytes before it in the slab) + its own stay below the threshold
            stop = int(np.searchsorted(cum, before + threshold, side="left"))
            stop = max(stop, start + 1)
            members = idx[start:stop]
            offsets = cum[start:stop] - sizes[start:stop] - before
            size = int(cum[stop - 1]) - before
            location = os.path.join("batched", str(uuid.uuid4()))
            lo_list, sz_list = offsets.tolist(), sizes[start:stop].tolist()
            for k, i in enumerate(members.tolist()):
                lo = lo_list[k]
                entries[paths[i]] = _tensor_entry(location, dtype_strs[i], shapes[i], replicated[i], [lo, lo + sz_list[k]])
            write_reqs.append(WriteReq(location, GPUSlabStager(
                int(dev), [tensors[i] for i in members.tolist()], ptrs[members], offsets, sizes[start:stop],
                contiguous[members], size)))
            start = stop
    # entries in the order the tensors were given (the manifest keeps insertion order)
    return {p: entries[p] for p in paths if p in entries}, write_reqs


class TakePlanCache:
    """The last bulk plan of a process, reused by the next take of the same tensors.

    A training loop checkpoints the same state dict over and over: same paths, same tensors at the same
    addresses with the same shapes.  Validating that (addresses, devices, dtypes, sizes, strides: one table
    from tensor_table.scan, compared with the previous one) costs a small part of planning again, and a take of
    13 k tensors is mostly planning.  On a hit the entries and slab stagers of the previous plan are used as they
    are: the manifest gets the same TensorEntry objects (they are never mutated after planning) and the slab
    files the same `batched/<uuid>` names, inside the new snapshot's own directory.  Any difference -- one
    address, one shape, a knob -- and the plan is made afresh.  TSB200_PLAN_CACHE=0 turns the cache off.

    The cached stagers refer to the tensors of the LATEST take only (a hit swaps them in), so nothing older
    than what the caller just checkpointed is kept alive."""

    def __init__(self) -> None:
        self.key: Any = None
        self.value: Any = None
        self.hits = 0
        self.misses = 0
        self._columns: Any = None  # wire form of the cached entries (manifest.tensor_columns_of), made on first use

    def columns_for(self, manifest: Dict[str, Any]) -> Any:
        """Pre-encoded columns of the cached plan's entries when `manifest` carries exactly those objects (the
        take that is being planned hit, or just filled, the cache); else None."""
        if self.value is None or not self.enabled():
            return None
        entries = self.value[0]
        if not entries:
            return None
        first = next(iter(entries))
        if manifest.get(first) is not entries[first]:
            return None
        if self._columns is None or self._columns[0] is not entries:
            from .manifest import tensor_columns_of

            self._columns = tensor_columns_of(entries)
        return self._columns

    @staticmethod
    def enabled() -> bool:
        return os.environ.get("TSB200_PLAN_CACHE", "1").lower() not in ("0", "false")

    @staticmethod
    def fingerprint(tensors: List[Any], table: Optional[Any] = None) -> Any:
        """What has to be unchanged for a plan to stand: address, device, dtype, sizes, strides of every tensor
        (the rows of tensor_table.scan)."""
        if table is None:
            from . import tensor_table

            table = tensor_table.scan(tensors)
        return table

    @staticmethod
    def _same(a: Any, b: Any) -> bool:
        return a.shape == b.shape and bool(np.array_equal(a, b))

    def get(self, config: Any, paths: List[str], tensors: List[Any], table: Optional[Any] = None
            ) -> Optional[Tuple[Dict[str, Any], List[WriteReq]]]:
        if self.key is None or not self.enabled():
            return None
        cfg, cached_paths, fp = self.key
        if cfg != config or cached_paths != paths or not self._same(fp, self.fingerprint(tensors, table)):
            self.misses += 1
            return None
        self.hits += 1
        entries, write_reqs, members = self.value
        # the stagers keep THIS take's tensors alive from here on
        for wr, idx in zip(write_reqs, members):
            stager = wr.buffer_stager
            if isinstance(stager, GPUSlabStager):
                stager.tensors = [tensors[i] for i in idx]
            else:
                stager.tensor = tensors[idx[0]]
        return entries, write_reqs

    def put(self, config: Any, paths: List[str], tensors: List[Any], entries: Dict[str, Any],
            write_reqs: List[WriteReq], table: Optional[Any] = None) -> None:
        if not self.enabled():
            return
        index = {id(t): i for i, t in enumerate(tensors)}
        members = []
        for wr in write_reqs:
            stager = wr.buffer_stager
            ts = stager.tensors if isinstance(stager, GPUSlabStager) else [stager.tensor]
            members.append([index[id(t)] for t in ts])
        self.key = (config, list(paths), self.fingerprint(tensors, table))
        self.value = (entries, write_reqs, members)
        from .manifest import register_stable_entries

        register_stable_entries(entries.values())  # the metadata writer serialises these once


take_plan_cache = TakePlanCache()


class GPUSlabConsumer(BufferConsumer):
    """Mirror of GPUSlabStager for the read side (`plan_cuda_reads`): the members of one merged read of a slab
    file, held as arrays; the region table of the unpack launch is built by numpy."""

    def __init__(self, device: int, tensors: List[Any], ptrs: Any, offsets: Any, nbytes: Any, contiguous: Any,
                 size: int) -> None:
        self.device = device
        self.tensors = tensors
        self._ptrs, self._offsets, self._nbytes, self._contiguous = ptrs, offsets, nbytes, contiguous
        self.buf_sz_bytes = size

    def is_device_consumable(self) -> bool:
        return True

    def consumed_nbytes(self) -> int:
        return self.buf_sz_bytes

    def device_index(self) -> int:
        return self.device

    def unpack_region_array(self, src_addr: int) -> Any:
        import numpy as np

        from .regions import regions_to_array, unpack_regions

        dense = self._contiguous & (self._nbytes > 0)
        arr = np.zeros(int(dense.sum()), dtype=_native.REGION_DTYPE)
        arr["src"] = self._offsets[dense].astype(np.uint64) + np.uint64(src_addr)
        arr["dst"] = self._ptrs[dense]
        arr["run_bytes"] = self._nbytes[dense].astype(np.uint64)
        strided = np.flatnonzero(~self._contiguous)
        if strided.size:
            regs: List[RegionTuple] = []
            for i in strided:
                regs += unpack_regions(src_addr + int(self._offsets[i]), self.tensors[int(i)])
            arr = np.concatenate([arr, regions_to_array(regs)])
        return arr

    def unpack_regions(self, src_addr: int) -> List[RegionTuple]:
        out: List[RegionTuple] = []
        for rec in self.unpack_region_array(src_addr):
            nd = int(rec["ndim"])
            out.append((int(rec["src"]), int(rec["dst"]), int(rec["run_bytes"]), tuple(int(x) for x in rec["size"][:nd]),
                        tuple(int(x) for x in rec["src_stride"][:nd]), tuple(int(x) for x in rec["dst_stride"][:nd])))
        return out

    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        from .scheduler import consume_device_buffer

        await consume_device_buffer(self, buf, self.device, executor)

    def get_consuming_cost_bytes(self) -> int:
        return self.buf_sz_bytes


def plan_cuda_reads(items: List[Tuple[Any, Any]], max_merged_bytes: Optional[int] = None) -> List[ReadReq]:
    """Read requests for slab members restored in place into plain CUDA tensors, planned in bulk: `items` are
    (TensorEntry with a byte range, destination tensor of the entry's dtype and shape).  See
    plan_cuda_read_columns, which does the work."""
    from . import tensor_table as tt

    n = len(items)
    if n == 0:
        return []
    ids: Dict[str, int] = {}
    file_id = np.fromiter((ids.setdefault(e.location, len(ids)) for e, _t in items), dtype=np.int64, count=n)
    lo = np.fromiter((e.byte_range[0] for e, _t in items), dtype=np.int64, count=n)
    hi = np.fromiter((e.byte_range[1] for e, _t in items), dtype=np.int64, count=n)
    tensors = [t for _e, t in items]
    return plan_cuda_read_columns(list(ids), file_id, lo, hi, tensors, tt.scan(tensors), max_merged_bytes)


def plan_cuda_read_columns(locations: List[str], file_id: Any, lo: Any, hi: Any, tensors: List[Any], table: Any,
                           max_merged_bytes: Optional[int] = None) -> List[ReadReq]:
    """Read requests for slab members restored in place into plain CUDA tensors, from columns: member i is bytes
    [lo[i], hi[i]) of file locations[file_id[i]] and goes to tensors[i], whose row of tensor_table.scan is
    table[i] (a plain CUDA tensor of the saved dtype and shape: the caller has checked).  Gives the reads
    TensorIOPreparer.prepare_read per entry followed by batch_read_requests gives -- the ranged reads of one
    file merged into one read of [min lo, max hi), split at holes above 1 MiB and at the size bound -- with one
    consumer per read instead of one per member; members bound for different devices never share a read."""
    from . import tensor_table as tt

    limit = _MAX_MERGED_READ_BYTES if max_merged_bytes is None else max(int(max_merged_bytes), 1)
    n = len(tensors)
    if n == 0:
        return []
    ptrs = tt.ptrs_of(table)
    contiguous = table[:, tt.CONTIGUOUS] != 0
    devices = table[:, tt.DEVICE]
    # one sort by (device, file, offset): the members of a file are contiguous afterwards
    order = np.lexsort((lo, file_id, devices))
    file_id, lo, hi, ptrs, contiguous, devices = file_id[order], lo[order], hi[order], ptrs[order], contiguous[order], devices[order]
    order_list = order.tolist()
    # a new read starts with a new file (or device), or after a hole above the limit (members of a slab never
    # overlap, so the end of the previous member is the end of everything before it)
    starts = np.flatnonzero(np.concatenate(([True], (file_id[1:] != file_id[:-1]) | (devices[1:] != devices[:-1])
                                            | (lo[1:] - hi[:-1] > _MAX_READ_HOLE_BYTES))))
    bounds = starts.tolist() + [n]
    out: List[ReadReq] = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        # the size bound: cut where the merged range would pass it
        cuts = [a]
        if int(hi[b - 1] - lo[a]) > limit:
            begin = int(lo[a])
            for k in range(a + 1, b):
                if int(hi[k]) - begin > limit:
                    cuts.append(k)
                    begin = int(lo[k])
        cuts.append(b)
        for c, d in zip(cuts[:-1], cuts[1:]):
            first, last = int(lo[c]), int(hi[c:d].max())
            out.append(ReadReq(
                path=locations[int(file_id[c])],
                buffer_consumer=GPUSlabConsumer(
                    int(devices[c]), [tensors[i] for i in order_list[c:d]], ptrs[c:d], lo[c:d] - first,
                    hi[c:d] - lo[c:d], contiguous[c:d], last - first),
                byte_range=(first, last),
            ))
    return out


class SlabType(Enum):
    CPU = 0
    GPU = 1


class Slab:
    def __init__(self, type: SlabType, device: Optional[int] = None) -> None:
        self.type = type
        self.device = device
        self.byte_ranges: List[Tuple[int, int]] = []
        self.buffer_stagers: List[BufferStager] = []
        self.location: str = os.path.join("batched", str(uuid.uuid4()))
        self.sz_bytes: int = 0

    def add_buffer_stager(self, byte_range: Tuple[int, int], buffer_stager: BufferStager) -> None:
        self.buffer_stagers.append(buffer_stager)
        self.byte_ranges.append(byte_range)
        self.sz_bytes += byte_range[1] - byte_range[0]

    def build(self) -> BufferStager:
        members = dict(zip(self.byte_ranges, self.buffer_stagers))
        if self.type == SlabType.CPU:
            return BatchedBufferStager(byte_range_to_buffer_stager=members)
        if self.type == SlabType.GPU:
            return GPUBatchedBufferStager(byte_range_to_buffer_stager=members)
        raise ValueError(f"Unrecognized slab type: {self.type}")


def _tensor_entries_by_location(entries: List[Any]) -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    for entry in entries:
        kind = getattr(entry, "type", None)
        if kind == "Tensor":
            out[entry.location] = entry
        elif kind == "ChunkedTensor":
            for chunk in entry.chunks:
                out[chunk.tensor.location] = chunk.tensor
        elif kind == "ShardedTensor":
            for shard in entry.shards:
                out[shard.tensor.location] = shard.tensor
    return out


def batch_write_requests(
    entries: List[Any],
    write_reqs: List[WriteReq],
    slab_size_threshold_bytes: Optional[int] = None,
) -> Tuple[List[Any], List[WriteReq]]:
    """Batch small tensor write requests into slab write requests (reference batcher.py:201-352).

    Relocates the matching TensorEntry objects (also those nested in chunked / sharded entries)
    in place: `location` becomes the slab's `batched/<uuid4>` and `byte_range` the member's range.
    """
    threshold = slab_size_threshold_bytes or knobs.get_slab_size_threshold_bytes()
    batched: List[WriteReq] = []
    cpu_slabs: List[Slab] = [Slab(SlabType.CPU)]
    gpu_slabs: Dict[int, List[Slab]] = {}
    relocation: Dict[str, Tuple[str, int, int]] = {}

    buffer_protocol = Serializer.BUFFER_PROTOCOL.value
    for wr in write_reqs:
        stager = wr.buffer_stager
        # is_batchable(), inlined: this loop runs once per tensor
        try:
            batchable = stager.entry.serializer == buffer_protocol and stager._tensor_prepare_func is None
            tensor = stager.tensor
        except AttributeError:
            batchable = False
        if not batchable:
            batched.append(wr)
            continue
        nbytes = tensor.numel() * tensor.element_size()
        if nbytes >= threshold:
            batched.append(wr)
            continue
        if tensor.is_cuda:
            dev = tensor.device.index
            slabs = gpu_slabs.get(dev)
            if slabs is None:
                slabs = gpu_slabs[dev] = [Slab(SlabType.GPU, dev)]
        else:
            slabs = cpu_slabs
        slab = slabs[-1]
        if slab.sz_bytes + nbytes >= threshold:
            slab = Slab(slab.type, slab.device)
            slabs.append(slab)
        lo = slab.sz_bytes
        slab.buffer_stagers.append(stager)
        slab.byte_ranges.append((lo, lo + nbytes))
        slab.sz_bytes = lo + nbytes
        relocation[wr.path] = (slab.location, lo, lo + nbytes)

    for slab in itertools.chain(cpu_slabs, *gpu_slabs.values()):
        if slab.buffer_stagers:
            batched.append(WriteReq(path=slab.location, buffer_stager=slab.build()))

    by_location = _tensor_entries_by_location(entries) if relocation else {}
    for location, (new_location, lo, hi) in relocation.items():
        entry = by_location.get(location)
        if entry is None:
            raise RuntimeError(f"The tensor entry with the location {location} is not passed to batch_write.")
        entry.location = new_location
        entry.byte_range = [lo, hi]
    return entries, batched


class BatchedBufferConsumer(BufferConsumer):
    """Fans one buffer (a byte range of a slab file) out to member consumers (reference
    batcher.py:355-381).  When every member can consume on the device, the whole buffer goes to
    HBM in one H2D copy and ONE scatter-unpack launch; no per-member host slicing."""

    def __init__(
        self, byte_range_to_buffer_consumer: Dict[Tuple[int, int], BufferConsumer], buf_sz_bytes: int
    ) -> None:
        self.byte_range_to_buffer_consumer = byte_range_to_buffer_consumer
        self.buf_sz_bytes = buf_sz_bytes

    # -- device side ------------------------------------------------------------------------
    def device_members(self) -> bool:
        return all(
            getattr(c, "is_device_consumable", lambda: False)() for c in self.byte_range_to_buffer_consumer.values()
        )

    def is_device_consumable(self) -> bool:
        return len(self.byte_range_to_buffer_consumer) > 0 and self.device_members()

    def consumed_nbytes(self) -> int:
        return max(hi for (_lo, hi) in self.byte_range_to_buffer_consumer.keys())

    def unpack_regions(self, src_addr: int) -> List[RegionTuple]:
        out: List[RegionTuple] = []
        for (lo, _hi), consumer in self.byte_range_to_buffer_consumer.items():
            out += consumer.unpack_regions(src_addr + lo)
        return out

    def device_index(self) -> int:
        for consumer in self.byte_range_to_buffer_consumer.values():
            return _consumer_device(consumer)
        return 0

    # -- BufferConsumer ---------------------------------------------------------------------
    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        if self.is_device_consumable():
            from .scheduler import consume_device_buffer

            await consume_device_buffer(self, buf, self.device_index(), executor)
            return
        view = memoryview(buf)
        # members that load straight into a dense CPU tensor of the saved dtype and size: one native scatter
        # for all of them instead of a frombuffer + copy_ (and a thread hop) each
        fast, slow = [], []
        for (lo, hi), consumer in self.byte_range_to_buffer_consumer.items():
            dst = _host_destination_of(consumer, hi - lo)
            (fast if dst is not None else slow).append(((lo, hi), consumer, dst))
        if fast:
            from .engine import readonly_buffer_address

            base = readonly_buffer_address(view)
            dsts = [t.data_ptr() for _r, _c, t in fast]
            srcs = [base + lo for (lo, _hi), _c, _t in fast]
            sizes = [hi - lo for (lo, hi), _c, _t in fast]
            if executor is None:
                _native.host_gather(dsts, srcs, sizes)
            else:
                await asyncio.get_running_loop().run_in_executor(executor, _native.host_gather, dsts, srcs, sizes)
        tasks = [
            asyncio.ensure_future(consumer.consume_buffer(view[lo:hi], executor=executor))
            for (lo, hi), consumer, _dst in slow
        ]
        if tasks:
            await asyncio.gather(*tasks)

    def get_consuming_cost_bytes(self) -> int:
        return self.buf_sz_bytes + sum(c.get_consuming_cost_bytes() for c in self.byte_range_to_buffer_consumer.values())


def _host_destination_of(consumer: Any, nbytes: int) -> Optional[Any]:
    """The dense CPU tensor a plain tensor consumer loads `nbytes` buffer-protocol bytes into, when a raw
    byte copy is exactly what `tensor_copy(dst, frombuffer(...))` would do (same dtype, contiguous, same
    size); None otherwise."""
    try:
        entry, tensor = consumer.entry, consumer.tensor
        if entry.serializer != Serializer.BUFFER_PROTOCOL.value or type(entry).__name__ != "TensorEntry":
            return None
    except AttributeError:
        return None
    if not isinstance(tensor, torch.Tensor) or tensor.is_cuda or tensor.is_quantized or tensor.layout != torch.strided:
        return None
    if not tensor.is_contiguous() or str(tensor.dtype) != entry.dtype or nbytes == 0:
        return None
    if tensor.numel() * tensor.element_size() != nbytes or list(tensor.shape) != list(entry.shape):
        return None
    return tensor.detach()


# a gap between two wanted members larger than this starts a new read of the same file
_MAX_READ_HOLE_BYTES = 1 << 20
# upper bound of one merged read (the default maximum chunk / shard size of the writer)
_MAX_MERGED_READ_BYTES = 512 << 20


def _consumer_device(consumer: Any) -> int:
    if hasattr(consumer, "tensor"):
        return consumer.tensor.device.index or 0
    if hasattr(consumer, "overlapping_regions"):
        return consumer.overlapping_regions[0].dst_tensor.device.index or 0
    if hasattr(consumer, "device_index"):
        return consumer.device_index()
    return 0


def batch_read_requests(read_reqs: List[ReadReq], max_merged_bytes: Optional[int] = None) -> List[ReadReq]:
    """Merge the ranged reads of each file into one read of [min lo, max hi) (reference
    batcher.py:384-474).  Whole-file reads pass through."""
    limit = _MAX_MERGED_READ_BYTES if max_merged_bytes is None else max(int(max_merged_bytes), 1)
    out: List[ReadReq] = []
    by_location: Dict[str, List[ReadReq]] = defaultdict(list)
    for rr in read_reqs:
        if rr.byte_range is None:
            out.append(rr)
        else:
            by_location[rr.path].append(rr)
    for location, rrs in by_location.items():
        # One read per run of members: the reference merges everything into [min lo, max hi) and leaves
        # "split when there are large holes" as a TODO (batcher.py:436-441); random access to two members
        # at the far ends of a slab would otherwise read the whole slab.
        rrs = sorted(rrs, key=lambda rr: rr.byte_range)
        groups: List[List[ReadReq]] = [[rrs[0]]]
        lo, hi = rrs[0].byte_range
        for rr in rrs[1:]:
            too_far = rr.byte_range[0] - hi > _MAX_READ_HOLE_BYTES
            # a merged read is one pinned buffer: keep it within the size of the largest storage object the
            # writer produces, or the tiles of a budgeted read_object would be glued back into one huge read
            too_big = max(hi, rr.byte_range[1]) - lo > limit
            if too_far or too_big:
                groups.append([rr])
                lo, hi = rr.byte_range
            else:
                groups[-1].append(rr)
                hi = max(hi, rr.byte_range[1])
        for group in groups:
            lo = min(rr.byte_range[0] for rr in group)
            hi = max(rr.byte_range[1] for rr in group)
            members = {(rr.byte_range[0] - lo, rr.byte_range[1] - lo): rr.buffer_consumer for rr in group}
            out.append(
                ReadReq(
                    path=location,
                    # the merged range, not the last member's (the reference under-counts here,
                    # batcher.py:469)
                    buffer_consumer=BatchedBufferConsumer(byte_range_to_buffer_consumer=members, buf_sz_bytes=hi - lo),
                    byte_range=(lo, hi),
                )
            )
    return out
