"""Write/read planning for tensors, chunked tensors, sharded tensors and objects.

Host half of the planner.  Class and function names, arguments and error behaviour follow the
reference (torchsnapshot/io_preparer.py:46-165 and io_preparers/{tensor,chunked_tensor,
sharded_tensor,object}.py) so the same tests drive both; what differs is what a stager or consumer
*does*: for CUDA tensors it emits strided region descriptors for the multi-region copy kernel
(`pack_regions` / `unpack_regions`) instead of calling `.to("cpu")` / `.copy_()` per tensor
(reference tensor.py:252-254, 329-338; sharded_tensor.py:296-309).

CPU tensors never touch the GPU: their bytes are already in host memory and are handed to the
storage plugin as a zero-copy memoryview, exactly like the reference.  That is not a fallback of
the GPU path -- a CUDA tensor is never routed to it.
"""

import asyncio
import io
import itertools
import logging
import math
import os
from collections import defaultdict
from concurrent.futures import Executor
from dataclasses import dataclass
from functools import reduce
from operator import mul
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import torch

from . import knobs
from .io_types import BufferConsumer, BufferStager, BufferType, Future, ReadReq, WriteReq
from .manifest import (
    ChunkedTensorEntry,
    Entry,
    ObjectEntry,
    PrimitiveEntry,
    Shard,
    ShardedTensorEntry,
    TensorEntry,
)
from .regions import RegionTuple, box_regions, contiguous_strides, pack_regions, unpack_regions
from .serialization import (
    tensor_as_memoryview,
    tensor_from_memoryview,
    _BUFFER_PROTOCOL_SET as _BUFFER_PROTOCOL_DTYPES,
    _DTYPE_TO_STRING as _DTYPE_STRINGS,
    _STRING_TO_DTYPE as _STRING_DTYPES,
    Serializer,
    SUPPORTED_QUANTIZED_DTYPES,
    dtype_to_element_size,
    dtype_to_string,
    string_to_dtype,
)

logger: logging.Logger = logging.getLogger(__name__)

TensorPrepareFunc = Callable[[torch.Tensor, bool], torch.Tensor]

# hot-loop constants (prepare_write runs once per tensor: 20k times at BASELINE config 2)
_BUFFER_PROTOCOL = Serializer.BUFFER_PROTOCOL.value
_TORCH_SAVE = Serializer.TORCH_SAVE.value

try:  # ShardedTensor is optional at import time; only sharded state needs it
    from torch.distributed._shard.sharded_tensor import ShardedTensor
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec
except Exception:  # pragma: no cover
    ShardedTensor = None  # type: ignore
    ChunkShardingSpec = None  # type: ignore


def _is_sharded_tensor(obj: Any) -> bool:
    return ShardedTensor is not None and isinstance(obj, ShardedTensor)


async def _run_in_executor(executor: Optional[Executor], func: Callable[..., Any], *args: Any) -> Any:
    if executor is not None:
        return await asyncio.get_running_loop().run_in_executor(executor, func, *args)
    return func(*args)


# ----------------------------------------------------------------------------------------------
# host-side byte views (CPU tensors only)
# ----------------------------------------------------------------------------------------------
def tensor_copy(dst: torch.Tensor, src: torch.Tensor) -> None:
    """dst <- src with the reference's quantization rule (tensor.py:383-403)."""
    if src.is_quantized and (not dst.is_quantized or dst.qscheme() != src.qscheme() or dst.dtype != src.dtype):
        src = src.dequantize()
    dst.detach().copy_(src)


# ----------------------------------------------------------------------------------------------
# dense tensors
# ----------------------------------------------------------------------------------------------
class TensorBufferStager(BufferStager):
    def __init__(
        self,
        tensor: torch.Tensor,
        entry: TensorEntry,
        is_async_snapshot: bool,
        _tensor_prepare_func: Optional[TensorPrepareFunc] = None,
    ) -> None:
        self.tensor = tensor
        self.entry = entry
        self.is_async_snapshot = is_async_snapshot
        self._tensor_prepare_func = _tensor_prepare_func

    # -- device side ------------------------------------------------------------------------
    def is_device_stageable(self) -> bool:
        """True if the payload can be produced by the pack kernel straight from HBM."""
        return (
            self.tensor.is_cuda
            and self.entry.serializer == Serializer.BUFFER_PROTOCOL.value
            and self._tensor_prepare_func is None
        )

    def staged_nbytes(self) -> int:
        return self.tensor.numel() * self.tensor.element_size()

    def pack_regions(self, dst_addr: int) -> List[RegionTuple]:
        return pack_regions(self.tensor, dst_addr)  # only addresses and strides are read

    # -- BufferStager -------------------------------------------------------------------------
    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        if self.is_device_stageable():
            from .scheduler import stage_device_buffer  # single-request device path

            return await stage_device_buffer(self, self.tensor.device.index, executor)

        tensor = self.tensor
        custom_prepared = False
        if self._tensor_prepare_func is not None:
            tensor = self._tensor_prepare_func(self.tensor, False)
            custom_prepared = tensor.untyped_storage().data_ptr() != self.tensor.untyped_storage().data_ptr()
        if tensor.is_cuda:
            # torch_save dtypes or a user prepare func: host serialisation needs a host tensor
            cpu_tensor = await _run_in_executor(executor, lambda t: t.detach().to("cpu"), tensor)
        else:
            cpu_tensor = tensor.detach()
            if self.is_async_snapshot and not custom_prepared:
                # the application may mutate the tensor before the background write finishes
                cpu_tensor = cpu_tensor.clone()
        if self.entry.serializer == Serializer.TORCH_SAVE.value:
            if cpu_tensor.numel() != cpu_tensor.untyped_storage().nbytes() // max(cpu_tensor.element_size(), 1):
                cpu_tensor = cpu_tensor.clone()  # don't pickle the whole backing storage of a view
            buf = io.BytesIO()
            torch.save(cpu_tensor, buf)
            return buf.getvalue()
        if self.entry.serializer == Serializer.BUFFER_PROTOCOL.value:
            return tensor_as_memoryview(cpu_tensor)
        raise ValueError(f"Unrecognized serializer: {self.entry.serializer}.")

    def get_staging_cost_bytes(self) -> int:
        n = TensorIOPreparer.get_tensor_size_from_entry(self.entry)
        if self.entry.serializer == Serializer.TORCH_SAVE.value:
            return n * 2
        if self.entry.serializer == Serializer.BUFFER_PROTOCOL.value:
            return n
        raise ValueError(f"Unrecognized serializer: {self.entry.serializer}.")


class TensorBufferConsumer(BufferConsumer):
    def __init__(self, tensor: torch.Tensor, entry: TensorEntry) -> None:
        self.tensor = tensor
        self.entry = entry

    # -- device side ------------------------------------------------------------------------
    def is_device_consumable(self) -> bool:
        return (
            self.tensor.is_cuda
            and self.entry.serializer == Serializer.BUFFER_PROTOCOL.value
            and string_to_dtype(self.entry.dtype) == self.tensor.dtype
        )

    def consumed_nbytes(self) -> int:
        return TensorIOPreparer.get_tensor_size_from_entry(self.entry)

    def unpack_regions(self, src_addr: int) -> List[RegionTuple]:
        return unpack_regions(src_addr, self.tensor)

    # -- BufferConsumer ---------------------------------------------------------------------
    @staticmethod
    def deserialize_tensor(buf: BufferType, entry: TensorEntry) -> torch.Tensor:
        if entry.serializer == Serializer.TORCH_SAVE.value:
            return torch.load(io.BytesIO(buf), weights_only=False)
        if entry.serializer == Serializer.BUFFER_PROTOCOL.value:
            return tensor_from_memoryview(memoryview(buf), dtype=string_to_dtype(entry.dtype), shape=entry.shape)
        raise ValueError(f"Unrecognized serializer: {entry.serializer}.")

    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        if self.is_device_consumable():
            from .scheduler import consume_device_buffer

            await consume_device_buffer(self, buf, self.tensor.device.index, executor)
            return
        loaded = self.deserialize_tensor(buf=buf, entry=self.entry)
        if executor is None or (len(buf) < _INLINE_COPY_BYTES and not self.tensor.is_cuda):
            # a thread hop costs more than copying a small host tensor (the reference hops per tensor,
            # tensor.py:329-338: 4 s for 20 k tensors)
            tensor_copy(self.tensor, loaded)
        else:
            await _run_in_executor(executor, tensor_copy, self.tensor, loaded)

    def get_consuming_cost_bytes(self) -> int:
        n = TensorIOPreparer.get_tensor_size_from_entry(self.entry)
        if self.entry.serializer == Serializer.TORCH_SAVE.value:
            return n * 2
        if self.entry.serializer == Serializer.BUFFER_PROTOCOL.value:
            return n
        raise ValueError(f"Unrecognized serializer: {self.entry.serializer}.")


_INLINE_COPY_BYTES = 256 * 1024


class TensorIOPreparer:
    @staticmethod
    def prepare_write(
        storage_path: str,
        tensor: torch.Tensor,
        is_async_snapshot: bool = False,
        _tensor_prepare_func: Optional[TensorPrepareFunc] = None,
    ) -> Tuple[TensorEntry, List[WriteReq]]:
        if _tensor_prepare_func is None:
            dtype = tensor.dtype
        else:
            proc = _tensor_prepare_func(tensor, True)
            if proc.shape != tensor.shape:
                raise RuntimeError(
                    "_tensor_prepare_func shouldn't change the tensor's shape "
                    f"(changed from {tensor.shape} to {proc.shape})."
                )
            dtype = proc.dtype
        dtype_str = _DTYPE_STRINGS.get(dtype)
        if dtype_str is None:
            dtype_str = dtype_to_string(dtype)  # raises for unsupported dtypes
        entry = TensorEntry(
            storage_path,
            _BUFFER_PROTOCOL if dtype in _BUFFER_PROTOCOL_DTYPES else _TORCH_SAVE,
            dtype_str,
            list(tensor.shape),
            False,
        )
        stager = TensorBufferStager(tensor, entry, is_async_snapshot, _tensor_prepare_func)
        return entry, [WriteReq(storage_path, stager)]

    @classmethod
    def prepare_read(
        cls,
        entry: TensorEntry,
        tensor_out: Optional[torch.Tensor] = None,
        buffer_size_limit_bytes: Optional[int] = None,
    ) -> Tuple[List[ReadReq], Future]:
        if (
            tensor_out is None
            or not isinstance(tensor_out, torch.Tensor)
            or _STRING_DTYPES.get(entry.dtype) != tensor_out.dtype
            or entry.shape != list(tensor_out.shape)
        ):
            tensor_out = cls.empty_tensor_from_entry(entry)
        if buffer_size_limit_bytes is not None and entry.serializer == Serializer.BUFFER_PROTOCOL.value:
            return cls.prepare_read_tiled(entry, tensor_out, buffer_size_limit_bytes)
        br = entry.byte_range
        return [ReadReq(entry.location, TensorBufferConsumer(tensor_out, entry), None if br is None else (br[0], br[1]))], Future(
            tensor_out
        )

    @classmethod
    def prepare_read_tiled(
        cls, entry: TensorEntry, tensor_out: torch.Tensor, buffer_size_limit_bytes: int
    ) -> Tuple[List[ReadReq], Future]:
        """Split one tensor read into ceil(nbytes / limit) ranged reads (reference tensor.py:127-179)."""
        nbytes = cls.get_tensor_size_from_entry(entry)
        num_tiles = max(math.ceil(nbytes / buffer_size_limit_bytes), 1)
        try:
            tensor_out = tensor_out.view(-1)  # finer tiling when the output is viewable as 1-d
        except RuntimeError:
            pass
        if tensor_out.ndim == 0:
            tensor_out = tensor_out.reshape(1)
        elsize = dtype_to_element_size(string_to_dtype(entry.dtype))
        base = 0 if entry.byte_range is None else entry.byte_range[0]
        read_reqs: List[ReadReq] = []
        offset = 0
        for tile in torch.chunk(tensor_out, chunks=num_tiles, dim=0):
            n = tile.numel() * elsize
            tile_entry = TensorEntry(
                location=entry.location,
                serializer=entry.serializer,
                dtype=entry.dtype,
                shape=list(tile.shape),
                replicated=entry.replicated,
            )
            read_reqs.append(
                ReadReq(
                    path=entry.location,
                    byte_range=(base + offset, base + offset + n),
                    buffer_consumer=TensorBufferConsumer(tensor=tile, entry=tile_entry),
                )
            )
            offset += n
        return read_reqs, Future(obj=tensor_out)

    @staticmethod
    def get_tensor_size_from_entry(entry: Union[TensorEntry, ChunkedTensorEntry]) -> int:
        return dtype_to_element_size(string_to_dtype(entry.dtype)) * reduce(mul, entry.shape, 1)

    @staticmethod
    def can_load_inplace(entry: Union[TensorEntry, ChunkedTensorEntry], obj: Any) -> bool:
        return (
            isinstance(obj, torch.Tensor)
            and string_to_dtype(entry.dtype) == obj.dtype
            and list(entry.shape) == list(obj.shape)
        )

    @staticmethod
    def empty_tensor_from_entry(entry: Union[TensorEntry, ChunkedTensorEntry]) -> torch.Tensor:
        dtype = string_to_dtype(entry.dtype)
        if dtype in SUPPORTED_QUANTIZED_DTYPES:
            raise RuntimeError("Allocating an empty quantized tensor is not supported yet.")
        return torch.empty(entry.shape, dtype=dtype)


# ----------------------------------------------------------------------------------------------
# chunked tensors
# ----------------------------------------------------------------------------------------------
@dataclass
class Chunk:
    offsets: List[int]
    sizes: List[int]
    dtype: str


class ChunkedTensorIOPreparer:
    @staticmethod
    def chunk_tensor(
        tensor: torch.Tensor, chunking_dim: int = 0, chunk_sz_bytes: Optional[int] = None
    ) -> List[Chunk]:
        """ceil(nbytes / chunk) pieces along `chunking_dim`, each ceil(dim / n) long except the last
        (the torch.chunk rule the reference uses, chunked_tensor.py:35-62)."""
        chunk_sz_bytes = chunk_sz_bytes or knobs.get_max_chunk_size_bytes()
        shape = list(tensor.shape) if tensor.ndim > 0 else [1]
        nbytes = tensor.numel() * tensor.element_size()
        n_chunks = max(math.ceil(nbytes / chunk_sz_bytes), 1)
        extent = shape[chunking_dim]
        step = max(math.ceil(extent / n_chunks), 1)
        chunks: List[Chunk] = []
        start = 0
        while True:
            length = min(step, extent - start)
            offsets = [0] * len(shape)
            offsets[chunking_dim] = start
            sizes = list(shape)
            sizes[chunking_dim] = length
            chunks.append(Chunk(offsets=offsets, sizes=sizes, dtype=str(tensor.dtype)))
            start += length
            if start >= extent:
                break
        return chunks

    @staticmethod
    def _get_subtensor_view(tensor: torch.Tensor, chunk: Union[Shard, Chunk]) -> torch.Tensor:
        view = tensor.view(-1) if tensor.ndim == 0 else tensor
        for d, (off, size) in enumerate(zip(chunk.offsets, chunk.sizes)):
            view = view.narrow(d, off, size)
        return view

    @classmethod
    def prepare_write(
        cls,
        storage_path: str,
        tensor: torch.Tensor,
        chunking_instruction: List[Chunk],
        is_async_snapshot: bool = False,
        _tensor_prepare_func: Optional[TensorPrepareFunc] = None,
    ) -> Tuple[ChunkedTensorEntry, List[WriteReq]]:
        write_reqs: List[WriteReq] = []
        chunks: List[Shard] = []
        for chunk in chunking_instruction:
            suffix = "_".join(str(x) for x in chunk.offsets)
            chunk_entry, chunk_wrs = TensorIOPreparer.prepare_write(
                storage_path=f"{storage_path}_{suffix}",
                tensor=cls._get_subtensor_view(tensor, chunk),
                is_async_snapshot=is_async_snapshot,
                _tensor_prepare_func=_tensor_prepare_func,
            )
            chunks.append(Shard(offsets=chunk.offsets, sizes=chunk.sizes, tensor=chunk_entry))
            write_reqs += chunk_wrs
        entry = ChunkedTensorEntry(
            dtype=dtype_to_string(tensor.dtype), shape=list(tensor.shape), chunks=chunks, replicated=False
        )
        return entry, write_reqs

    @classmethod
    def prepare_read(
        cls,
        entry: ChunkedTensorEntry,
        tensor_out: Optional[torch.Tensor] = None,
        buffer_size_limit_bytes: Optional[int] = None,
    ) -> Tuple[List[ReadReq], Future]:
        if tensor_out is None or not TensorIOPreparer.can_load_inplace(entry=entry, obj=tensor_out):
            tensor_out = TensorIOPreparer.empty_tensor_from_entry(entry)
        read_reqs: List[ReadReq] = []
        for chunk in entry.chunks:
            view = cls._get_subtensor_view(tensor_out, chunk)
            rrs, _ = TensorIOPreparer.prepare_read(chunk.tensor, view, buffer_size_limit_bytes)
            read_reqs += rrs
        return read_reqs, Future(obj=tensor_out)


# ----------------------------------------------------------------------------------------------
# sharded tensors
# ----------------------------------------------------------------------------------------------
def overlap_wrt_saved(
    saved_offsets: Sequence[int], saved_sizes: Sequence[int], cur_offsets: Sequence[int], cur_sizes: Sequence[int]
) -> Optional[List[Tuple[int, int, int, int]]]:
    """Per-dim (dim, offset in saved, offset in current, length) of the intersection of two boxes
    given in global coordinates, or None when they do not overlap (reference
    sharded_tensor.py:79-125 plus the overlap test it imports from torch)."""
    out = []
    for dim, (so, ss, co, cs) in enumerate(zip(saved_offsets, saved_sizes, cur_offsets, cur_sizes)):
        lo = max(so, co)
        hi = min(so + ss, co + cs)
        if hi <= lo:
            return None
        out.append((dim, lo - so, lo - co, hi - lo))
    return out


@dataclass
class _OverlappingRegion:
    dst_tensor: torch.Tensor
    overlap_region: List[Tuple[int, int, int, int]]  # (dim, src_offset, dst_offset, length)

    def get_views(self, src_tensor: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        src_view, dst_view = src_tensor, self.dst_tensor
        for dim, src_off, dst_off, length in self.overlap_region:
            src_view = torch.narrow(src_view, dim, src_off, length)
            dst_view = torch.narrow(dst_view, dim, dst_off, length)
        return src_view, dst_view


class ShardedTensorBufferConsumer(BufferConsumer):
    def __init__(self, overlapping_regions: List[_OverlappingRegion], entry: TensorEntry) -> None:
        self.overlapping_regions = overlapping_regions
        self.entry = entry

    def is_device_consumable(self) -> bool:
        if self.entry.serializer != Serializer.BUFFER_PROTOCOL.value:
            return False
        dtype = string_to_dtype(self.entry.dtype)
        return all(r.dst_tensor.is_cuda and r.dst_tensor.dtype == dtype for r in self.overlapping_regions)

    def consumed_nbytes(self) -> int:
        return TensorIOPreparer.get_tensor_size_from_entry(self.entry)

    def unpack_regions(self, src_addr: int) -> List[RegionTuple]:
        """One strided box per overlap: persisted piece (dense, at src_addr) -> local shard."""
        elsize = dtype_to_element_size(string_to_dtype(self.entry.dtype))
        src_strides = contiguous_strides(self.entry.shape, elsize)
        out: List[RegionTuple] = []
        for region in self.overlapping_regions:
            dst = region.dst_tensor.detach()
            dst_strides = [s * elsize for s in dst.stride()]
            src = src_addr
            daddr = dst.data_ptr()
            lens = []
            for dim, src_off, dst_off, length in region.overlap_region:
                src += src_off * src_strides[dim]
                daddr += dst_off * dst_strides[dim]
                lens.append(length)
            out += box_regions(src, src_strides, daddr, dst_strides, lens, elsize)
        return out

    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        if self.is_device_consumable():
            from .scheduler import consume_device_buffer

            device = self.overlapping_regions[0].dst_tensor.device.index
            await consume_device_buffer(self, buf, device, executor)
            return
        deserialized = TensorBufferConsumer.deserialize_tensor(buf=buf, entry=self.entry)
        for region in self.overlapping_regions:
            src_view, dst_view = region.get_views(src_tensor=deserialized)
            await _run_in_executor(executor, tensor_copy, dst_view, src_view)

    def get_consuming_cost_bytes(self) -> int:
        n = TensorIOPreparer.get_tensor_size_from_entry(self.entry)
        return n * 2 if self.entry.serializer == Serializer.TORCH_SAVE.value else n


class ShardedTensorIOPreparer:
    @staticmethod
    def subdivide_shard(
        shard: torch.Tensor, offsets: List[int], sizes: List[int], dim: int, max_shard_sz_bytes: int
    ) -> List[Tuple[torch.Tensor, List[int], List[int]]]:
        """Split a local shard along `dim` into pieces of at most max_shard_sz_bytes (at least one
        slice each); reference sharded_tensor.py:47-76."""
        if max_shard_sz_bytes <= 0:
            raise ValueError(f"max_shard_sz_bytes must be a positive integer (got {max_shard_sz_bytes}).")
        slice_bytes = reduce(mul, sizes) // sizes[dim] * shard.element_size()
        step = max(max_shard_sz_bytes // slice_bytes, 1)
        pieces = []
        for start in range(0, sizes[dim], step):
            length = min(step, sizes[dim] - start)
            sub_offsets = list(offsets)
            sub_offsets[dim] += start
            sub_sizes = list(sizes)
            sub_sizes[dim] = length
            pieces.append((torch.narrow(shard, dim, start, length), sub_offsets, sub_sizes))
        return pieces

    @staticmethod
    def _shards_get_overlap_region_wrt_saved_tensor(saved_shard: Any, current_shard: Any) -> List[Tuple[int, int, int, int]]:
        region = overlap_wrt_saved(
            saved_shard.shard_offsets, saved_shard.shard_sizes, current_shard.shard_offsets, current_shard.shard_sizes
        )
        if region is None:
            raise ValueError("shards do not overlap")
        return region

    @classmethod
    def prepare_write(
        cls,
        storage_path: str,
        obj: Any,
        is_async_snapshot: bool = False,
        _tensor_prepare_func: Optional[TensorPrepareFunc] = None,
    ) -> Tuple[ShardedTensorEntry, List[WriteReq]]:
        shards: List[Shard] = []
        write_reqs: List[WriteReq] = []
        spec = obj.sharding_spec()
        sharding_dim = spec.dim if ChunkShardingSpec is not None and isinstance(spec, ChunkShardingSpec) else 0
        for local in obj.local_shards():
            pieces = cls.subdivide_shard(
                shard=local.tensor,
                offsets=list(local.metadata.shard_offsets),
                sizes=list(local.metadata.shard_sizes),
                dim=sharding_dim,
                max_shard_sz_bytes=knobs.get_max_shard_size_bytes(),
            )
            for tensor, offsets, sizes in pieces:
                suffix = "_".join(str(i) for i in offsets)
                entry, wrs = TensorIOPreparer.prepare_write(
                    storage_path=f"{storage_path}_{suffix}",
                    tensor=tensor,
                    is_async_snapshot=is_async_snapshot,
                    _tensor_prepare_func=_tensor_prepare_func,
                )
                write_reqs += wrs
                shards.append(Shard(offsets=offsets, sizes=sizes, tensor=entry))
        return ShardedTensorEntry(shards=shards), write_reqs

    @staticmethod
    def _get_global_shape(entry: ShardedTensorEntry) -> List[int]:
        shape = [0] * len(entry.shards[0].sizes)
        for shard in entry.shards:
            for d, (off, size) in enumerate(zip(shard.offsets, shard.sizes)):
                shape[d] = max(shape[d], off + size)
        return shape

    @staticmethod
    def local_boxes(obj_out: Any) -> List[Tuple[torch.Tensor, List[int], List[int]]]:
        """(tensor, global offsets, sizes) of every local shard; a plain Tensor is one shard that
        covers everything (reference sharded_tensor.py:209-225)."""
        if _is_sharded_tensor(obj_out):
            return [
                (s.tensor, list(s.metadata.shard_offsets), list(s.metadata.shard_sizes)) for s in obj_out.local_shards()
            ]
        if isinstance(obj_out, torch.Tensor):
            return [(obj_out, [0] * obj_out.ndim, list(obj_out.shape))]
        raise RuntimeError(f"obj_out must either be a Tensor or ShardedTensor (got {type(obj_out)})")

    @staticmethod
    def _trim_piece_to_needed_rows(
        piece: TensorEntry, regions: List["_OverlappingRegion"]
    ) -> Tuple[TensorEntry, List["_OverlappingRegion"]]:
        """Read only the rows of a persisted piece that some local shard overlaps.

        The reference reads every overlapping piece whole -- up to 512 MiB for a sliver
        (sharded_tensor.py:249-267).  Rows of a dense row-major piece are contiguous in storage, so the
        needed rows [lo, hi) of dim 0 are one byte range; the piece entry is narrowed to it and the
        source offsets of the overlap regions are shifted by lo.  Pickled (torch_save) pieces and
        pieces that are needed (almost) whole are left alone."""
        if piece.serializer != Serializer.BUFFER_PROTOCOL.value or not piece.shape or not regions:
            return piece, regions
        rows = piece.shape[0]
        lo, hi = rows, 0
        for r in regions:
            _dim, src_off, _dst_off, length = r.overlap_region[0]
            lo = min(lo, src_off)
            hi = max(hi, src_off + length)
        if lo >= hi or (hi - lo) * 8 > rows * 7:  # less than 1/8 to save: one whole-object read is simpler
            return piece, regions
        row_bytes = dtype_to_element_size(string_to_dtype(piece.dtype)) * reduce(mul, piece.shape[1:], 1)
        base = 0 if piece.byte_range is None else piece.byte_range[0]
        trimmed = TensorEntry(
            location=piece.location,
            serializer=piece.serializer,
            dtype=piece.dtype,
            shape=[hi - lo] + list(piece.shape[1:]),
            replicated=piece.replicated,
            byte_range=[base + lo * row_bytes, base + hi * row_bytes],
        )
        shifted = [
            _OverlappingRegion(
                dst_tensor=r.dst_tensor,
                overlap_region=[(d, so - lo if d == 0 else so, do, ln) for d, so, do, ln in r.overlap_region],
            )
            for r in regions
        ]
        return trimmed, shifted

    @classmethod
    def prepare_read(
        cls, entry: ShardedTensorEntry, obj_out: Optional[Any] = None
    ) -> Tuple[List[ReadReq], Future]:
        if obj_out is None:
            raise RuntimeError("Reading a ShardedTensor without a runtime object is not supported.")
        global_shape = cls._get_global_shape(entry)
        out_shape = list(obj_out.metadata().size) if _is_sharded_tensor(obj_out) else list(obj_out.shape)
        if out_shape != global_shape:
            logger.warning(
                f"The shape of obj_out ({out_shape}) is different from the shape of the persisted sharded "
                f"tensor ({global_shape}). Only the overlapping part will be loaded. "
            )
        locals_ = cls.local_boxes(obj_out)
        # every persisted piece is read once and fans out to all the local boxes it intersects
        regions_by_piece: Dict[Tuple[str, Optional[Tuple[int, int]]], List[_OverlappingRegion]] = defaultdict(list)
        for (tensor, l_off, l_sizes), shard in itertools.product(locals_, entry.shards):
            region = overlap_wrt_saved(shard.offsets, shard.sizes, l_off, l_sizes)
            if region is None:
                continue
            key = (shard.tensor.location, shard.tensor.byte_range_tuple)
            regions_by_piece[key].append(_OverlappingRegion(dst_tensor=tensor, overlap_region=region))
        read_reqs: List[ReadReq] = []
        for shard in entry.shards:
            key = (shard.tensor.location, shard.tensor.byte_range_tuple)
            if key not in regions_by_piece:
                continue
            piece_entry, regions = cls._trim_piece_to_needed_rows(shard.tensor, regions_by_piece[key])
            read_reqs.append(
                ReadReq(
                    path=piece_entry.location,
                    buffer_consumer=ShardedTensorBufferConsumer(overlapping_regions=regions, entry=piece_entry),
                    byte_range=piece_entry.byte_range_tuple,
                )
            )
        return read_reqs, Future(obj=obj_out)


# ----------------------------------------------------------------------------------------------
# objects and primitives (host only; unchanged semantics, reference io_preparers/object.py)
# ----------------------------------------------------------------------------------------------
class ObjectBufferStager(BufferStager):
    def __init__(self, obj: Any) -> None:
        self.obj = obj

    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        buf = io.BytesIO()
        torch.save(self.obj, buf)
        return buf.getvalue()

    def get_staging_cost_bytes(self) -> int:
        import sys

        return sys.getsizeof(self.obj)


class ObjectBufferConsumer(BufferConsumer):
    def __init__(self, obj_out: Future) -> None:
        self.fut = obj_out
        self.consuming_cost_bytes = 0

    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        self.fut.obj = torch.load(io.BytesIO(buf), weights_only=False)

    def set_consuming_cost_bytes(self, n: int) -> None:
        self.consuming_cost_bytes = n

    def get_consuming_cost_bytes(self) -> int:
        return self.consuming_cost_bytes


class ObjectIOPreparer:
    @staticmethod
    def prepare_write(storage_path: str, obj: Any) -> Tuple[ObjectEntry, List[WriteReq]]:
        entry = ObjectEntry(
            location=storage_path,
            serializer=Serializer.TORCH_SAVE.value,
            obj_type=f"{type(obj).__module__}.{type(obj).__name__}",
            replicated=False,
        )
        return entry, [WriteReq(path=storage_path, buffer_stager=ObjectBufferStager(obj))]

    @staticmethod
    def prepare_read(entry: ObjectEntry, obj_out: Optional[Any] = None) -> Tuple[List[ReadReq], Future]:
        fut: Future = Future()
        return [ReadReq(path=entry.location, buffer_consumer=ObjectBufferConsumer(fut))], fut


class PrimitivePreparer:
    @staticmethod
    def should_inline(obj: Any) -> bool:
        return type(obj).__name__ in PrimitiveEntry.supported_types

    @staticmethod
    def prepare_write(obj: Any) -> PrimitiveEntry:
        return PrimitiveEntry.from_object(obj)

    @staticmethod
    def prepare_read(entry: PrimitiveEntry) -> Tuple[List[ReadReq], Future]:
        return [], Future(obj=entry.get_value())


# ----------------------------------------------------------------------------------------------
# dispatch (reference io_preparer.py:46-165)
# ----------------------------------------------------------------------------------------------
def get_storage_path(obj: Any, logical_path: str, rank: int, replicated: bool) -> str:
    if _is_sharded_tensor(obj):
        return "sharded/" + logical_path
    if replicated:
        return "replicated/" + logical_path
    return f"{rank}/{logical_path}"


def prepare_write(
    obj: Any,
    logical_path: str,
    rank: int,
    replicated: bool,
    is_async_snapshot: bool = False,
    _tensor_prepare_func: Optional[TensorPrepareFunc] = None,
    max_chunk_size_bytes: Optional[int] = None,
) -> Tuple[Entry, List[WriteReq]]:
    """Entry + write requests of one leaf (reference io_preparer.py:73-128).  `max_chunk_size_bytes`
    lets a caller that plans thousands of leaves read the environment knob once."""
    if type(obj) is torch.Tensor or type(obj) is torch.nn.Parameter:
        # the overwhelmingly common leaf: skip the primitive / sharded dispatch
        if max_chunk_size_bytes is None:
            max_chunk_size_bytes = knobs.get_max_chunk_size_bytes()
        storage_path = ("replicated/" if replicated else f"{rank}/") + logical_path
        if obj.numel() * obj.element_size() > max_chunk_size_bytes:
            entry, wrs = ChunkedTensorIOPreparer.prepare_write(
                storage_path=storage_path,
                tensor=obj,
                chunking_instruction=ChunkedTensorIOPreparer.chunk_tensor(obj, chunk_sz_bytes=max_chunk_size_bytes),
                is_async_snapshot=is_async_snapshot,
                _tensor_prepare_func=_tensor_prepare_func,
            )
        else:
            entry, wrs = TensorIOPreparer.prepare_write(storage_path, obj, is_async_snapshot, _tensor_prepare_func)
        entry.replicated = replicated
        return entry, wrs
    if PrimitivePreparer.should_inline(obj):
        entry = PrimitivePreparer.prepare_write(obj)
        entry.replicated = replicated
        return entry, []
    storage_path = get_storage_path(obj, logical_path, rank, replicated)
    if _is_sharded_tensor(obj):
        return ShardedTensorIOPreparer.prepare_write(
            storage_path=storage_path, obj=obj, is_async_snapshot=is_async_snapshot,
            _tensor_prepare_func=_tensor_prepare_func,
        )
    if isinstance(obj, torch.Tensor):
        if max_chunk_size_bytes is None:
            max_chunk_size_bytes = knobs.get_max_chunk_size_bytes()
        if obj.numel() * obj.element_size() > max_chunk_size_bytes:
            entry, wrs = ChunkedTensorIOPreparer.prepare_write(
                storage_path=storage_path,
                tensor=obj,
                chunking_instruction=ChunkedTensorIOPreparer.chunk_tensor(obj, chunk_sz_bytes=max_chunk_size_bytes),
                is_async_snapshot=is_async_snapshot,
                _tensor_prepare_func=_tensor_prepare_func,
            )
        else:
            entry, wrs = TensorIOPreparer.prepare_write(
                storage_path=storage_path, tensor=obj, is_async_snapshot=is_async_snapshot,
                _tensor_prepare_func=_tensor_prepare_func,
            )
    else:
        entry, wrs = ObjectIOPreparer.prepare_write(storage_path, obj)
    entry.replicated = replicated
    return entry, wrs


def prepare_read(
    entry: Entry, obj_out: Optional[Any] = None, buffer_size_limit_bytes: Optional[int] = None
) -> Tuple[List[ReadReq], Future]:
    kind = getattr(entry, "type", None)
    if kind == "ShardedTensor":
        if obj_out is None:
            raise RuntimeError("Reading a ShardedTensor without a runtime object is not supported.")
        return ShardedTensorIOPreparer.prepare_read(entry, obj_out)
    if kind == "ChunkedTensor":
        return ChunkedTensorIOPreparer.prepare_read(entry, obj_out, buffer_size_limit_bytes=buffer_size_limit_bytes)
    if kind == "Tensor":
        return TensorIOPreparer.prepare_read(entry, obj_out, buffer_size_limit_bytes=buffer_size_limit_bytes)
    if kind == "object":
        return ObjectIOPreparer.prepare_read(entry, obj_out)
    if kind in PrimitiveEntry.supported_types:
        return PrimitivePreparer.prepare_read(entry)
    raise Exception(f"Unsupported entry type: {entry} ({kind}).")
