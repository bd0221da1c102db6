"""Snapshot.take / async_take / restore / read_object over the B200 hot path.

The public surface and the on-disk result are the reference's (torchsnapshot/snapshot.py:66-947):
same arguments, same `.snapshot_metadata` document, same file layout (`<rank>/...`,
`replicated/...`, `sharded/...`, `batched/<uuid>`), same commit rule (payload first, metadata last,
written by rank 0 after every rank has finished).  This module is deliberately thin -- it is the
caller of the hot path, kept only so the path can be driven and measured end to end without the
reference installed; the planning and byte movement live in io_preparer / batcher / scheduler.

Fewer metadata collectives than the reference (3 object all-gathers per take instead of ~12,
SURVEY 2.3): path + replication globs + keys + hostname travel in one, the replicated-path
verification and the write partitioning in a second, the manifest in a third.
"""

import asyncio
import contextlib
import fnmatch
import functools
import gc
import itertools
import logging
import os
import socket
import sys
import threading
import time
from collections import defaultdict
from concurrent.futures import ThreadPoolExecutor
from typing import Any, Callable, Dict, List, NamedTuple, Optional, Set, Tuple

import numpy as np

import torch

from . import knobs
from .batcher import batch_read_requests, batch_write_requests
from .flatten import _encode, flatten, inflate
from .io_preparer import _is_sharded_tensor, prepare_read, prepare_write
from .io_types import ReadIO, ReadReq, StoragePlugin, WriteIO, WriteReq
from .manifest import (
    manifest_digest,
    Entry,
    Manifest,
    PrimitiveEntry,
    ShardedTensorEntry,
    SnapshotMetadata,
    TensorEntry,
    LazySnapshotMetadata,
    decode_manifest,
    encode_manifest,
    is_container_entry,
    is_replicated,
)
from .partitioner import RanksDisagree, _digest as _paths_digest, consolidate_replicated_entries, partition_write_reqs
from .pg_wrapper import PGWrapper
from .scheduler import (
    PendingIOWork,
    _MAX_PER_RANK_MEMORY_BUDGET_BYTES,
    get_process_memory_budget_bytes,
    set_local_world_size,
    sync_execute_read_reqs,
    sync_execute_write_reqs,
)
from .manifest_ops import get_manifest_for_rank, handle_sharded_tensor_elasticity
from .state_dict import StateDict  # noqa: F401  (re-exported: `from ...snapshot import Snapshot, StateDict`)
from .stateful import AppState
from .storage_plugin import url_to_storage_plugin

logger: logging.Logger = logging.getLogger(__name__)

SNAPSHOT_METADATA_FNAME = ".snapshot_metadata"
VERSION = "0.1.0"  # the metadata schema version the reference writes (torchsnapshot/version.py:17)

# wall-clock milliseconds of the phases of the last take on this rank, up to the point where take /
# async_take hands over to the storage writes (what async_take's caller is blocked for): introspection for
# bench.py and scripts/, a handful of clock reads per take
last_take_phases: Dict[str, float] = {}

# (what every rank sent in the manifest exchange of the last take, the metadata built from it): a job that
# checkpoints the same state again sends the same bytes, and the document is the same
_last_gathered_manifests: Optional[Tuple[Any, Any]] = None



def _validate_app_state(app_state: AppState) -> None:
    for key, value in app_state.items():
        if not (hasattr(value, "state_dict") and hasattr(value, "load_state_dict")):
            raise TypeError(f"Expected Stateful in app_state for key {key}, got {type(value)}.")


def _infer_replicated(replicated: List[str], app_state: AppState) -> List[str]:
    out = list(replicated)
    if "**" in out:
        return out
    from torch.nn.parallel import DistributedDataParallel as DDP

    for key, val in app_state.items():
        if isinstance(val, DDP):
            ignored = set(getattr(val, "parameters_to_ignore", []) or [])
            if not ignored:
                out.append(os.path.join(key, "**"))
                continue
            for name, _ in itertools.chain(val.named_parameters(), val.named_buffers()):
                if name not in ignored:
                    out.append(os.path.join(key, name))
    return out


class _Bulk(NamedTuple):
    """The leaves batcher.plan_cuda_tensors plans: paths, tensors, and their rows of tensor_table.scan."""
    paths: List[str]
    tensors: List[Any]
    table: Any


def _bulk_plannable(
    flattened: Dict[str, Any], replicated_paths: Set[str], world: int, prepare_func: Optional[Any], max_chunk: int,
    table: Optional[Any] = None,
) -> Optional[_Bulk]:
    """The leaves batcher.plan_cuda_tensors plans (plain CUDA tensors of buffer-protocol dtypes that are not
    chunked), or None when the per-leaf planner has to do everything: bulk planning must cover EVERY member of
    the GPU slabs, or the slabs would not be the ones the reference forms (batcher.py:291-316).  So none of it
    when batching is off, a prepare function is set, chunks can be smaller than a slab, or some would-be slab
    member cannot be planned here: a sharded tensor (its pieces join slabs), a tensor subclass, a tensor the
    table does not describe, replicated state of a multi-rank job (the partitioner assigns it request by
    request).  `table` is tensor_table.scan(list(flattened.values())) when the caller already has it."""
    if prepare_func is not None or knobs.is_batching_disabled():
        return None
    threshold = knobs.get_slab_size_threshold_bytes()
    if max_chunk < threshold:
        return None
    from . import tensor_table as tt

    values = list(flattened.values())
    if table is None:
        table = tt.scan(values)
    kind = table[:, tt.KIND]
    if (kind >= tt.UNDESCRIBED).any():
        return None
    for i in np.flatnonzero(kind == tt.NOT_A_TENSOR).tolist():
        if _is_sharded_tensor(values[i]):
            return None
    # (tensors above the chunk size are handed back by plan_cuda_tensors: their chunks are at least a slab long
    # here, so they never join a slab)
    chosen = (kind == tt.PLAIN) & (table[:, tt.DEVICE] >= 0) & _buffer_protocol_mask()[table[:, tt.DTYPE]]
    count = int(chosen.sum())
    if count < 64:  # a handful of tensors plans just as fast the ordinary way
        return None
    paths = list(flattened.keys())
    if count == len(values):
        tensors = values
    else:
        idx = np.flatnonzero(chosen)
        table = table[idx]
        idx = idx.tolist()
        paths, tensors = [paths[i] for i in idx], [values[i] for i in idx]
    if world > 1 and replicated_paths and not replicated_paths.isdisjoint(paths):
        return None
    return _Bulk(paths, tensors, table)


_buffer_protocol_codes: Any = None


def _buffer_protocol_mask() -> Any:
    """Look-up table over tensor_table dtype codes: True for the dtypes serialised through the buffer protocol."""
    global _buffer_protocol_codes
    if _buffer_protocol_codes is None:
        from . import tensor_table as tt
        from .serialization import _BUFFER_PROTOCOL_SET

        _buffer_protocol_codes = tt.dtype_mask(_BUFFER_PROTOCOL_SET)
    return _buffer_protocol_codes


@contextlib.contextmanager
def _gc_paused():
    """Planning creates a few objects per tensor; a generational GC pass in the middle of it was the
    largest source of jitter in async_take's blocked time (25 ms vs 60-80 ms at 2312 tensors)."""
    was_enabled = gc.isenabled()
    gc.disable()
    try:
        yield
    finally:
        if was_enabled:
            gc.enable()


def _memory_budget(local_world_size: int) -> int:
    if "TORCHSNAPSHOT_PER_RANK_MEMORY_BUDGET_BYTES" in os.environ:
        return get_process_memory_budget_bytes(None)
    import psutil

    return min(int(psutil.virtual_memory().available * 0.6) // max(local_world_size, 1), _MAX_PER_RANK_MEMORY_BUDGET_BYTES)


class _RepeatTake:
    """The plan a rank of a multi-rank job made for its share of a replicated state dict, kept for the next take of
    the same tensors: entries (after partitioning and slab batching) and write requests.

    Eligible state dicts hold only containers, inline primitives and plain tensors of buffer-protocol dtypes that
    are not chunked (a DDP model plus tensor-valued optimizer state).  The next take may repeat the plan
    when (1) this rank's tensors are where and what they were (same paths, addresses, shapes, strides, dtypes:
    `TakePlanCache.fingerprint`), (2) every other rank says the same of its own, and (3) the previous split of the
    writes still stands: a fresh split for the throughputs reported now would not end noticeably earlier (partitioner.py) -- (2) and (3)
    are decided from the gathered exchange, identically on every rank.  Anything else plans afresh.
    TSB200_PLAN_CACHE=0 turns it off."""

    last: Optional["_RepeatTake"] = None
    repeats = 0  # takes that repeated the previous plan (introspection for tests and bench.py)

    def __init__(self, config: Any, paths: List[str], fingerprint: Any) -> None:
        self.config, self.paths, self.fingerprint = config, paths, fingerprint
        self.nonrep = 0
        self.digest = ""
        self.object_entries: Dict[str, Entry] = {}
        self.write_reqs: List[WriteReq] = []
        self.swaps: List[Tuple[Any, Any]] = []  # (stager, logical path | list of logical paths)
        self.primitive_paths: List[str] = []
        self._columns: Any = None

    @staticmethod
    def candidate(flattened: Dict[str, Any], rank: int, world: int, globs: Set[str], prepare_func: Optional[Any],
                  max_chunk: int, table: Optional[Any] = None) -> Optional["_RepeatTake"]:
        """`table` is tensor_table.scan(list(flattened.values())) when the caller already has it."""
        from . import tensor_table as tt
        from .batcher import TakePlanCache
        from .io_preparer import PrimitivePreparer

        if prepare_func is not None or not TakePlanCache.enabled():
            return None
        values = list(flattened.values())
        if table is None:
            table = tt.scan(values)
        kind = table[:, tt.KIND]
        plain = kind == tt.PLAIN
        if (kind >= tt.UNDESCRIBED).any() or not plain.any():
            return None
        if not _buffer_protocol_mask()[table[plain, tt.DTYPE]].all():
            return None
        all_paths = list(flattened.keys())
        primitives: List[str] = []
        for i in np.flatnonzero(~plain).tolist():
            if not PrimitivePreparer.should_inline(values[i]):
                return None
            primitives.append(all_paths[i])
        if primitives:
            paths = [all_paths[i] for i in np.flatnonzero(plain).tolist()]
            table = table[plain]
        else:
            paths = all_paths
        config = (rank, world, tuple(sorted(globs)), max_chunk, knobs.get_slab_size_threshold_bytes(),
                  knobs.is_batching_disabled(), tuple(primitives))
        rt = _RepeatTake(config, paths, table)
        rt.primitive_paths = primitives
        return rt

    def columns_for(self, manifest: Dict[str, Any]) -> Any:
        """Wire form of this plan's entries (manifest.tensor_columns_of), when `manifest` carries those objects."""
        if not self.object_entries:
            return None
        first = next(iter(self.object_entries))
        if manifest.get(first) is not self.object_entries[first]:
            return None
        if self._columns is None:
            from .manifest import tensor_columns_of

            self._columns = tensor_columns_of(self.object_entries)
        return self._columns

    def matches(self, other: "_RepeatTake") -> bool:
        from .batcher import TakePlanCache

        return (self.config == other.config and self.paths == other.paths
                and TakePlanCache._same(self.fingerprint, other.fingerprint))

    @staticmethod
    def remember(rt: "_RepeatTake", flattened: Dict[str, Any], object_entries: Dict[str, Entry],
                 write_reqs: List[WriteReq]) -> None:
        from . import partitioner
        from .batcher import GPUSlabStager

        if partitioner.last_partition_inputs is None:
            return
        by_id = {id(flattened[p]): p for p in rt.paths}
        swaps: List[Tuple[Any, Any]] = []
        try:
            for wr in write_reqs:
                stager = wr.buffer_stager
                if isinstance(stager, GPUSlabStager):
                    swaps.append((stager, [by_id[id(t)] for t in stager.tensors]))
                elif hasattr(stager, "byte_range_to_buffer_stager"):
                    for member in stager.byte_range_to_buffer_stager.values():
                        swaps.append((member, by_id[id(member.tensor)]))
                else:
                    swaps.append((stager, by_id[id(stager.tensor)]))
        except (KeyError, AttributeError):
            return  # a stager that does not hold one of the flattened tensors as it is (a chunk view, ...): no repeat
        if any(getattr(e, "type", None) != "Tensor" for e in object_entries.values()):
            return
        rt.nonrep, rt.digest = partitioner.last_partition_inputs
        rt.object_entries, rt.write_reqs, rt.swaps = object_entries, write_reqs, swaps
        _RepeatTake.last = rt

    def reuse(self, flattened: Dict[str, Any], is_async: bool, replicated_paths: Set[str]
              ) -> Tuple[Dict[str, Entry], List[WriteReq], Dict[str, Entry]]:
        """The previous plan over this take's tensor objects (the stagers keep the caller's current aliases alive,
        not last take's), plus this take's inline primitives."""
        for stager, where in self.swaps:
            if isinstance(where, list):
                stager.tensors = [flattened[p] for p in where]
            else:
                stager.tensor = flattened[where]
                stager.is_async_snapshot = is_async
        primitives: Dict[str, Entry] = {}
        for p in self.primitive_paths:
            entry = PrimitiveEntry.from_object(flattened[p])
            entry.replicated = p in replicated_paths
            primitives[p] = entry
        _RepeatTake.repeats += 1
        return self.object_entries, self.write_reqs, primitives


def clear_plan_caches() -> None:
    """Drop the plans kept for the next take (`batcher.TakePlanCache`, `_RepeatTake`, the last metadata exchange).
    They refer to the tensors of the LAST take only, but that is still a reference: a caller that frees a model
    and wants its memory back before the next checkpoint calls this (or sets TSB200_PLAN_CACHE=0)."""
    global _last_gathered_manifests
    from .batcher import take_plan_cache
    from .manifest import register_stable_entries

    take_plan_cache.key = take_plan_cache.value = None
    take_plan_cache._columns = None
    register_stable_entries([])
    _RepeatTake.last = None
    _last_gathered_manifests = None


class _RestoreView:
    """One rank's entries under one stateful's prefix, split for `Snapshot._load_stateful`: slab members that
    batcher.plan_cuda_read_columns can plan, as columns (what a destination has to match: dtype code, rank,
    sizes; where the bytes are: file, byte range), container entries, and the rest, which is planned entry by
    entry.  Made once per Snapshot object, rank and prefix; the entries are the manifest's own objects."""

    def __init__(self, manifest: Manifest, merged_sd: Dict[str, ShardedTensorEntry], prefix: str) -> None:
        from . import tensor_table as tt
        from .serialization import _STRING_TO_DTYPE, Serializer

        buffer_protocol = Serializer.BUFFER_PROTOCOL.value
        self.manifest, self.merged_sd = manifest, merged_sd
        self.containers: Manifest = {}
        self.rest: List[Tuple[str, Entry]] = []
        self.paths: List[str] = []
        self.entries: List[Entry] = []
        self.has_sharded = False
        self.locations: List[str] = []
        file_of: Dict[str, int] = {}
        codes: List[int] = []
        shapes: List[List[int]] = []
        ranges: List[Any] = []
        files: List[int] = []
        replicated: List[bool] = []
        prefix_slash = prefix + "/"
        code_of: Dict[str, int] = {}  # dtype string -> table code (-1: not a dtype the table names)
        max_dims = tt.MAX_DIMS
        paths, entries, rest, containers = self.paths, self.entries, self.rest, self.containers
        for logical_path, entry in manifest.items():
            if not logical_path.startswith(prefix_slash) and logical_path != prefix:
                continue
            if type(entry) is TensorEntry:
                span = entry.byte_range
                if span is not None and entry.serializer == buffer_protocol:
                    dtype = entry.dtype
                    code = code_of.get(dtype)
                    if code is None:
                        code = code_of[dtype] = tt.dtype_code(_STRING_TO_DTYPE.get(dtype))
                    shape = entry.shape
                    if code >= 0 and len(shape) <= max_dims:
                        paths.append(logical_path)
                        entries.append(entry)
                        codes.append(code)
                        shapes.append(shape)
                        ranges.append(span)
                        location = entry.location
                        file_id = file_of.get(location)
                        if file_id is None:
                            file_id = file_of[location] = len(file_of)
                        files.append(file_id)
                        replicated.append(entry.replicated)
                        continue
                rest.append((logical_path, entry))
            elif is_container_entry(entry):
                containers[logical_path] = entry
            else:
                if isinstance(entry, ShardedTensorEntry):
                    self.has_sharded = True
                rest.append((logical_path, entry))
        n = len(self.paths)
        self.index = set(self.paths)
        self.locations = list(file_of)
        self.dtype_codes = np.array(codes, dtype=np.int64)
        self.ndims = np.fromiter(map(len, shapes), dtype=np.int64, count=n)
        self.shapes = np.zeros((n, tt.MAX_DIMS), dtype=np.int64)
        flat = np.fromiter(itertools.chain.from_iterable(shapes), dtype=np.int64, count=int(self.ndims.sum()))
        first = np.cumsum(self.ndims) - self.ndims
        self.shapes[np.repeat(np.arange(n), self.ndims), np.arange(flat.size) - np.repeat(first, self.ndims)] = flat
        span = np.fromiter(itertools.chain.from_iterable(ranges), dtype=np.int64, count=2 * n).reshape(n, 2)
        self.los, self.his = span[:, 0].copy(), span[:, 1].copy()
        self.file_ids = np.array(files, dtype=np.int64)
        self.replicated = np.array(replicated, dtype=np.bool_)  # (None counts as not replicated)


class _Ready:
    """A future that is already resolved: the object restored in place."""

    __slots__ = ("obj",)

    def __init__(self, obj: Any) -> None:
        self.obj = obj


class Snapshot:
    def __init__(self, path: str, pg: Optional[Any] = None, storage_options: Optional[Dict[str, Any]] = None) -> None:
        self.path = path
        self.pg = pg
        self._storage_options = storage_options
        self._metadata: Optional[SnapshotMetadata] = None

    # ------------------------------------------------------------------------------------------
    # take
    # ------------------------------------------------------------------------------------------
    @classmethod
    def take(
        cls,
        path: str,
        app_state: AppState,
        pg: Optional[Any] = None,
        replicated: Optional[List[str]] = None,
        storage_options: Optional[Dict[str, Any]] = None,
        _custom_tensor_prepare_func: Optional[Callable[[str, torch.Tensor, bool], torch.Tensor]] = None,
    ) -> "Snapshot":
        _validate_app_state(app_state)
        event_loop = asyncio.new_event_loop()
        pg_wrapper = PGWrapper(pg)
        with _gc_paused():
            path, pending, metadata, storage = cls._take_impl(
                path, app_state, replicated or [], pg_wrapper, storage_options, event_loop, False,
                _custom_tensor_prepare_func,
            )
        # serialise the manifest while the D2H copies and storage writes drain
        yaml_future = None
        if pg_wrapper.get_rank() == 0:
            yaml_pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="tsb-meta")
            yaml_future = yaml_pool.submit(metadata.to_yaml)
        pending.sync_complete(event_loop)
        # commit: every rank's payload is durable before rank 0 publishes the metadata
        pg_wrapper.barrier()
        if yaml_future is not None:
            storage.sync_write(
                WriteIO(path=SNAPSHOT_METADATA_FNAME, buf=yaml_future.result().encode("utf-8")), event_loop=event_loop
            )
            yaml_pool.shutdown(wait=False)
        storage.sync_close(event_loop=event_loop)
        event_loop.close()
        snapshot = cls(path=path, pg=pg, storage_options=storage_options)
        snapshot._metadata = metadata
        return snapshot

    @classmethod
    def async_take(
        cls,
        path: str,
        app_state: AppState,
        pg: Optional[Any] = None,
        replicated: Optional[List[str]] = None,
        storage_options: Optional[Dict[str, Any]] = None,
        _custom_tensor_prepare_func: Optional[Callable[[str, torch.Tensor, bool], torch.Tensor]] = None,
    ) -> "PendingSnapshot":
        """Returns once a consistent copy of the state exists (packed in HBM or staged on the host);
        D2H copies and storage I/O continue in a background thread."""
        _validate_app_state(app_state)
        event_loop = asyncio.new_event_loop()
        pg_wrapper = PGWrapper(pg)
        with _gc_paused():
            path, pending, metadata, storage = cls._take_impl(
                path, app_state, replicated or [], pg_wrapper, storage_options, event_loop, True,
                _custom_tensor_prepare_func,
            )
        return PendingSnapshot(path, pending, pg_wrapper, metadata, storage, event_loop, storage_options)

    @staticmethod
    def _coalesce_replicated(global_replicated: List[List[str]]) -> Set[str]:
        """Only globs every rank named are replicated (reference snapshot.py:809-812)."""
        return set.intersection(*map(set, global_replicated))

    @staticmethod
    def _infer_replicated(replicated: List[str], app_state: AppState) -> List[str]:
        return _infer_replicated(replicated, app_state)

    @classmethod
    def _take_impl(
        cls,
        path: str,
        app_state: AppState,
        replicated: List[str],
        pg: PGWrapper,
        storage_options: Optional[Dict[str, Any]],
        event_loop: asyncio.AbstractEventLoop,
        is_async: bool,
        prepare_func: Optional[Callable[[str, torch.Tensor, bool], torch.Tensor]],
    ) -> Tuple[str, PendingIOWork, SnapshotMetadata, StoragePlugin]:
        rank, world = pg.get_rank(), pg.get_world_size()
        app_state = dict(app_state)
        phases: Dict[str, float] = {}
        t_mark = time.monotonic()

        def mark(name: str) -> None:
            nonlocal t_mark
            now = time.monotonic()
            phases[name] = phases.get(name, 0.0) + 1e3 * (now - t_mark)
            t_mark = now

        # ---- collective 1: path, replication globs, keys, host ---------------------------------
        hello: List[Any] = [None] * world
        pg.all_gather_object(hello, (path, _infer_replicated(replicated, app_state), list(app_state.keys()),
                                     socket.gethostname()))
        if hello[0][0] != path:
            logger.warning(f"Rank {rank} specified a path ({path}) different from rank 0 ({hello[0][0]}).")
        path = hello[0][0]
        globs = cls._coalesce_replicated([h[1] for h in hello])
        global_keys = sorted(set(itertools.chain.from_iterable(h[2] for h in hello)))
        local_world = sum(1 for h in hello if h[3] == hello[rank][3])
        set_local_world_size(local_world)
        storage = url_to_storage_plugin(path, storage_options)
        from .partitioner import set_speed_context

        set_speed_context(type(storage).__name__)  # throughput history is kept per kind of sink
        mark("hello_gather")

        # ---- state dicts -> flat leaves (RNG state first, restored afterwards) -----------------
        manifest: Manifest = {}
        flattened: Dict[str, Any] = {}
        rng_items = [(k, v) for k, v in app_state.items() if type(v).__name__ == "RNGState"]
        if len(rng_items) > 1:
            raise RuntimeError(f"Multiple RNGState objects in app state: {[k for k, _ in rng_items]}")
        rng_sd = None
        if rng_items:
            rng_key, rng_obj = rng_items[0]
            del app_state[rng_key]
            rng_sd = rng_obj.state_dict()
            m, f = flatten(rng_sd, prefix=rng_key)
            manifest.update(m)
            flattened.update(f)
        for key in global_keys:
            if key in app_state:
                m, f = flatten(app_state[key].state_dict(), prefix=key)
                manifest.update(m)
                flattened.update(f)
            if world > 1:
                pg.barrier()  # state_dict() may run collectives; keep ranks in step
        if rng_items:
            rng_items[0][1].load_state_dict(rng_sd)
        mark("state_dict_flatten")

        # ---- plan --------------------------------------------------------------------------------
        if "**" in globs:  # everything is replicated (benchmarks/ddp): no need to match 20 k paths one by one
            candidates = [p for p, v in flattened.items() if not _is_sharded_tensor(v)]
        elif globs:
            candidates = [
                p for p, v in flattened.items()
                if not _is_sharded_tensor(v) and any(fnmatch.fnmatch(p, g) for g in globs)
            ]
        else:
            candidates = []
        max_chunk = knobs.get_max_chunk_size_bytes()
        # address, device, dtype, sizes and strides of every leaf, read once (what the bulk planner and both plan
        # caches look at)
        from . import tensor_table

        leaf_table = tensor_table.scan(list(flattened.values()))

        bulk_write_reqs: List[WriteReq] = []

        def plan(replicated_paths: Set[str]) -> Tuple[Dict[str, Entry], Dict[str, List[WriteReq]], Dict[str, Entry]]:
            objects: Dict[str, Entry] = {}
            wrs_by_path: Dict[str, List[WriteReq]] = {}
            primitives: Dict[str, Entry] = {}
            # plain CUDA tensors are planned in one vectorised pass when they are ALL of the GPU slabs' members
            bulk_write_reqs.clear()
            found = _bulk_plannable(flattened, replicated_paths, world, prepare_func, max_chunk, leaf_table)
            bulk_entries: Dict[str, Entry] = {}
            if found is not None:
                from .batcher import plan_cuda_tensors, take_plan_cache

                bulk, tensors, table = found
                flags = [p in replicated_paths for p in bulk] if replicated_paths else None
                # (sync / async is not part of the key: it only matters to CPU tensors, which are not planned here)
                config = (rank, world, max_chunk, knobs.get_slab_size_threshold_bytes(),
                          tuple(flags) if flags is not None else None)
                hit = take_plan_cache.get(config, bulk, tensors, table)
                if hit is not None:
                    bulk_entries, wrs = hit
                else:
                    own = f"{rank}/"
                    if flags is not None:
                        storage_paths = [("replicated/" if r else own) + p for p, r in zip(bulk, flags)]
                    else:
                        flags = [False] * len(bulk)
                        storage_paths = [own + p for p in bulk]
                    bulk_entries, wrs = plan_cuda_tensors(bulk, tensors, storage_paths, flags, None, is_async, max_chunk,
                                                          table)
                    take_plan_cache.put(config, bulk, tensors, bulk_entries, wrs, table)
                bulk_write_reqs.extend(wrs)
                if len(bulk_entries) == len(flattened):  # every leaf was planned above, in the order of `flattened`
                    return dict(bulk_entries), wrs_by_path, primitives
            for logical_path, obj in flattened.items():
                if logical_path in bulk_entries:
                    objects[logical_path] = bulk_entries[logical_path]
                    continue
                entry, wrs = prepare_write(
                    obj,
                    logical_path,
                    rank,
                    logical_path in replicated_paths,
                    is_async,
                    functools.partial(prepare_func, logical_path) if prepare_func else None,
                    max_chunk,
                )
                if isinstance(entry, PrimitiveEntry):
                    primitives[logical_path] = entry
                else:
                    objects[logical_path] = entry
                    wrs_by_path[logical_path] = wrs
            return objects, wrs_by_path, primitives

        # ---- collective 2: which candidate paths exist on every rank, and who writes what ---------
        # Replicated state has the same paths on every rank, so the plan is made for "all my candidates are
        # replicated" and the partitioner's one exchange also carries a digest of the candidates.  Only when
        # the ranks turn out to differ are the candidate lists themselves exchanged (the reference's way,
        # snapshot.py:796-822) and the plan redone for the paths every rank has.
        replicated_paths = set(candidates)
        # (`globs` is the same on every rank after the first exchange: without any, no rank has replicated
        # state, and there is nothing to agree on or to split)
        agree_on = _paths_digest(candidates) if world > 1 and globs else None
        # A multi-rank job that checkpoints the same replicated tensors again (same paths, addresses, shapes) and
        # whose previous split of the writes still stands repeats its previous plan: the exchange still happens
        # (it carries this take's throughputs, and tells everybody whether everybody can repeat), the planning
        # does not (`_RepeatTake`).
        repeat = (_RepeatTake.candidate(flattened, rank, world, globs, prepare_func, max_chunk, leaf_table)
                  if world > 1 and candidates else None)
        negotiated = None
        repeated = False
        object_entries: Dict[str, Entry] = {}
        write_reqs: List[WriteReq] = []
        try:
            if repeat is not None and _RepeatTake.last is not None and _RepeatTake.last.matches(repeat):
                from .partitioner import negotiate, previous_assignment_stands

                negotiated = negotiate(pg, _RepeatTake.last.nonrep, _RepeatTake.last.digest, agree_on, can_repeat=True)
                if previous_assignment_stands(negotiated):
                    object_entries, write_reqs, primitive_entries = _RepeatTake.last.reuse(
                        flattened, is_async, replicated_paths)
                    repeated = True
                    mark("partition_gather")
            if not repeated:
                object_entries, write_reqs_by_path, primitive_entries = plan(replicated_paths)
                mark("prepare_write")
                if world > 1 and globs:
                    object_entries, write_reqs_by_path = partition_write_reqs(
                        object_entries, write_reqs_by_path, pg, agree_on=agree_on, negotiated=negotiated
                    )
        except RanksDisagree:
            repeat = None
            seen: List[Any] = [None] * world
            pg.all_gather_object(seen, candidates, large=True)
            counts: Dict[str, int] = defaultdict(int)
            for paths in seen:
                for p in paths:
                    counts[p] += 1
            replicated_paths = {p for p in seen[0] if counts[p] == world}
            object_entries, write_reqs_by_path, primitive_entries = plan(replicated_paths)
            object_entries, write_reqs_by_path = partition_write_reqs(object_entries, write_reqs_by_path, pg)
        if not repeated:
            mark("partition_gather")
            write_reqs = [wr for wrs in write_reqs_by_path.values() for wr in wrs]
            if not knobs.is_batching_disabled():
                _, write_reqs = batch_write_requests(entries=list(object_entries.values()), write_reqs=write_reqs)
            write_reqs += bulk_write_reqs  # GPU slabs last, like the reference (batcher.py:319)
            if repeat is not None:
                _RepeatTake.remember(repeat, flattened, object_entries, write_reqs)
        manifest.update(primitive_entries)
        rest_of_manifest = dict(manifest)  # containers and inline primitives: what is not an object entry
        manifest.update(object_entries)
        mark("batch")

        # ---- collective 3: the manifest ----------------------------------------------------------
        # Every rank's entries travel as columns of strings and ints (manifest.encode_manifest), and nobody
        # rebuilds other ranks' entries before somebody needs them -- rank 0 in the thread that writes the
        # metadata file, after async_take has returned.  Only replicated entries have to be looked at right
        # away: they are merged and listed once, under rank 0 (partitioner.consolidate_replicated_entries).
        metadata: SnapshotMetadata
        if world == 1:  # with one rank every replicated entry already sits under rank 0
            metadata = LazySnapshotMetadata(VERSION, 1, [manifest])
        else:
            from .batcher import take_plan_cache

            columns = take_plan_cache.columns_for(manifest)
            if columns is None and repeat is not None and _RepeatTake.last is not None:
                columns = _RepeatTake.last.columns_for(manifest)
            # when the object entries are exactly the cached plan's (same objects, nothing else), only the rest
            # of the manifest is walked
            rest = None
            if columns is not None and len(columns[0]) == len(object_entries) and len(object_entries) > 0:
                last = next(reversed(object_entries))
                if columns[0].get(last) is object_entries[last] and len(rest_of_manifest) + len(object_entries) == len(manifest):
                    rest = rest_of_manifest
            # First a digest of what this rank would send, next to what it remembers of the previous exchange:
            # when every rank's manifest is what it was last time (a job that checkpoints the same state again),
            # the previous metadata document stands and no manifest travels.  Decided from gathered data only,
            # so every rank decides alike; a rank that remembers nothing (its caches were cleared) makes all
            # of them exchange the manifests.
            import hashlib
            import pickle

            global _last_gathered_manifests
            head, tail = encode_manifest(manifest, columns, rest, parts=True)
            mine = manifest_digest(head, columns)
            remembered = _last_gathered_manifests[0] if _last_gathered_manifests is not None else None
            said: List[Any] = [None] * world
            pg.all_gather_object(said, (bool(replicated_paths), mine, remembered))
            exchange = hashlib.blake2b(
                b"".join((b"R" if flag else b"-") + digest for flag, digest, _r in said), digest_size=16).digest()
            if all(r == exchange for _flag, _digest, r in said):
                metadata = _last_gathered_manifests[1]  # every rank would send what it sent last time: same document
            else:
                blobs: List[Any] = [None] * world
                pg.all_gather_object(blobs, pickle.dumps((head, tail), protocol=5), large=True)
                if any(flag for flag, _digest, _r in said):
                    manifests = consolidate_replicated_entries([decode_manifest(blob) for blob in blobs])
                    global_manifest: Manifest = {}
                    for r, m in enumerate(manifests):
                        prefix = f"{r}/"
                        for logical_path, entry in m.items():
                            global_manifest[prefix + logical_path] = entry
                    metadata = SnapshotMetadata(version=VERSION, world_size=world, manifest=global_manifest)
                else:
                    metadata = LazySnapshotMetadata(VERSION, world, blobs)
                _last_gathered_manifests = (exchange, metadata)
        mark("manifest_gather")

        pending = sync_execute_write_reqs(
            write_reqs=write_reqs,
            storage=storage,
            memory_budget_bytes=_memory_budget(local_world),
            rank=rank,
            event_loop=event_loop,
        )
        mark("stage")
        global last_take_phases
        last_take_phases = phases
        return path, pending, metadata, storage

    # ------------------------------------------------------------------------------------------
    # restore / read_object
    # ------------------------------------------------------------------------------------------
    def restore(self, app_state: AppState) -> None:
        with _gc_paused():  # same reason as in take: planning allocates a few objects per tensor
            self._restore_impl(app_state)

    def _restore_impl(self, app_state: AppState) -> None:
        _validate_app_state(app_state)
        event_loop = asyncio.new_event_loop()
        pg = PGWrapper(self.pg)
        rank, world = pg.get_rank(), pg.get_world_size()
        storage = url_to_storage_plugin(self.path, self._storage_options)
        app_state = dict(app_state)
        rng_items = [(k, v) for k, v in app_state.items() if type(v).__name__ == "RNGState"]
        keys: List[Any] = [None] * world
        pg.all_gather_object(keys, (list(app_state.keys()), socket.gethostname()))
        global_keys = sorted(set(itertools.chain.from_iterable(k[0] for k in keys)))
        local_world = sum(1 for k in keys if k[1] == keys[rank][1])
        set_local_world_size(local_world)
        budget = _memory_budget(local_world)
        rng_key = rng_items[0][0] if rng_items else None
        # keys every rank restores: only those can use the collective peer-to-peer reshard path -- and only
        # inside one host: the payload moves through CUDA IPC mappings and NVLink, which end at the box.  A job
        # that spans hosts restores the reference's way, every rank reading from storage what it needs.
        on_all_ranks = set.intersection(*(set(k[0]) for k in keys))
        one_host = len({k[1] for k in keys}) == 1
        for key in global_keys:
            if key != rng_key:
                p2p_pg = pg if (world > 1 and one_host and key in on_all_ranks and knobs.is_p2p_reshard_enabled()
                                and torch.cuda.is_available()) else None
                self._load_stateful(key, app_state.get(key), storage, rank, budget, event_loop, p2p_pg)
            if world > 1:
                pg.barrier()
        if rng_key is not None:  # RNG state last, so loading has no side effect on it
            self._load_stateful(rng_key, app_state[rng_key], storage, rank, budget, event_loop)
        storage.sync_close(event_loop=event_loop)
        event_loop.close()

    def _load_stateful(
        self, key: str, stateful: Optional[Any], storage: StoragePlugin, rank: int, budget: int,
        event_loop: asyncio.AbstractEventLoop, p2p_pg: Optional[PGWrapper] = None,
    ) -> None:
        if stateful is None:
            return
        from . import tensor_table as tt

        prefix = _encode(key)
        # load straight into the tensors the stateful already owns (no second copy of the state)
        _, flattened = flatten(stateful.state_dict(), prefix=key)
        # address, device, dtype, sizes and strides of every leaf, read once; only tensors and sharded tensors
        # are destinations
        paths, values = list(flattened.keys()), list(flattened.values())
        table = tt.scan(values)
        keep = table[:, tt.KIND] != tt.NOT_A_TENSOR
        for i in np.flatnonzero(~keep).tolist():
            if _is_sharded_tensor(values[i]):
                keep[i] = True
        if not keep.all():
            idx = np.flatnonzero(keep)
            table = table[idx]
            idx = idx.tolist()
            paths, values = [paths[i] for i in idx], [values[i] for i in idx]
            flattened = dict(zip(paths, values))
        read_reqs: List[ReadReq] = []
        futs: Dict[str, Any] = {}
        # Sharded tensors and replicated state restored into CUDA tensors by every rank: each persisted
        # piece / slab range is read from storage by ONE rank and reaches the others over NVLink
        # (reshard.py); everything else takes the ordinary path below.
        device = torch.cuda.current_device() if (p2p_pg is not None and torch.cuda.is_available()) else -1
        sharded_p2p: List[Tuple[str, ShardedTensorEntry, Any]] = []
        shared_read_reqs: List[ReadReq] = []

        def plan_entry(logical_path: str, entry: Entry, obj_out: Any) -> None:
            if isinstance(entry, ShardedTensorEntry):
                if obj_out is None:
                    return  # sharded entries can only be loaded into a runtime object
                if device >= 0:
                    from . import reshard

                    if reshard.eligible(entry, obj_out, device):
                        sharded_p2p.append((logical_path, entry, obj_out))
                        from .io_types import Future

                        futs[logical_path] = Future(obj=obj_out)
                        return
            rrs, fut = prepare_read(entry=entry, obj_out=obj_out)
            if (
                device >= 0
                and is_replicated(entry)
                and rrs
                and all(getattr(rr.buffer_consumer, "is_device_consumable", lambda: False)() for rr in rrs)
            ):
                shared_read_reqs.extend(rrs)
            else:
                read_reqs.extend(rrs)
            futs[logical_path] = fut

        # Slab members restored in place into plain CUDA tensors of the saved dtype and shape are planned in one
        # pass over columns (batcher.plan_cuda_reads): with thousands of small tensors the per-entry planner is
        # most of the time before the first read starts.  The entries' side of those columns is made once per
        # Snapshot object, rank and stateful (`_RestoreView`).
        view = self._restore_view(rank, prefix)
        bulk_reads: List[ReadReq] = []
        if view.merged_sd or view.has_sharded:
            # sharded tensors follow what this state dict asks for: requested ones this rank did not save are
            # added, saved ones nobody asks for are dropped -- on a copy of the manifest, walked entry by entry
            manifest = dict(view.manifest)
            handle_sharded_tensor_elasticity(manifest, view.merged_sd, paths)
            containers: Manifest = {}
            prefix_slash = prefix + "/"
            walk = []
            for logical_path, entry in manifest.items():
                if not logical_path.startswith(prefix_slash) and logical_path != prefix:
                    continue
                if is_container_entry(entry):
                    containers[logical_path] = entry
                elif logical_path not in view.index:
                    walk.append((logical_path, entry))
        else:
            containers = view.containers
            walk = view.rest
        if view.paths and paths and not knobs.is_batching_disabled():
            # rows of the destination table for the members, in the order of the view
            if view.paths == paths:
                rows = np.arange(len(paths), dtype=np.int64)
            else:
                position = {p: i for i, p in enumerate(paths)}
                rows = np.fromiter((position.get(p, -1) for p in view.paths), dtype=np.int64, count=len(view.paths))
            dest = table[np.maximum(rows, 0)]
            eligible = (
                (rows >= 0)
                & (dest[:, tt.KIND] == tt.PLAIN)
                & (dest[:, tt.DEVICE] >= 0)
                & (dest[:, tt.DTYPE] == view.dtype_codes)
                & (dest[:, tt.NDIM] == view.ndims)
                & (dest[:, tt.SIZES:tt.SIZES + tt.MAX_DIMS] == view.shapes).all(axis=1)
            )
            if device >= 0:
                eligible &= ~view.replicated  # replicated state of a multi-rank restore: read once, shared over NVLink
            chosen = np.flatnonzero(eligible)
            if chosen.size:
                from .batcher import plan_cuda_read_columns

                chosen_rows = rows[chosen]
                tensors = [values[i] for i in chosen_rows.tolist()]
                bulk_reads = plan_cuda_read_columns(
                    view.locations, view.file_ids[chosen], view.los[chosen], view.his[chosen], tensors,
                    table[chosen_rows])
                futs.update(zip([view.paths[i] for i in chosen.tolist()], map(_Ready, tensors)))
            left = np.flatnonzero(~eligible).tolist()
        else:
            left = range(len(view.paths))
        for i in left:
            logical_path = view.paths[i]
            plan_entry(logical_path, view.entries[i], flattened.get(logical_path))
        for logical_path, entry in walk:
            plan_entry(logical_path, entry, flattened.get(logical_path))
        if p2p_pg is not None and device >= 0:
            from . import reshard

            shared_batched = shared_read_reqs
            if not knobs.is_batching_disabled():
                shared_batched = batch_read_requests(read_reqs=shared_read_reqs)
            st = reshard.p2p_reshard_restore(
                sharded_p2p, storage, p2p_pg, event_loop, device, extra_units=reshard.read_req_units(shared_batched)
            )
            # what every rank alike left out of the shared-read path (reshard.py) takes the ordinary path
            # below, each rank reading from storage what it needs
            if st.fell_back or "replicated" in st.skipped_groups:
                read_reqs += shared_read_reqs
            if st.fell_back or "sharded" in st.skipped_groups:
                for _logical_path, entry, obj_out in sharded_p2p:
                    rrs, _fut = prepare_read(entry=entry, obj_out=obj_out)
                    read_reqs += rrs
        if not knobs.is_batching_disabled():
            read_reqs = batch_read_requests(read_reqs=read_reqs)
        read_reqs += bulk_reads
        sync_execute_read_reqs(
            read_reqs=read_reqs, storage=storage, memory_budget_bytes=budget, rank=rank, event_loop=event_loop
        )
        state_dict = inflate(containers, {k: fut.obj for k, fut in futs.items()}, prefix=key)
        stateful.load_state_dict(state_dict)

    def read_object(self, path: str, obj_out: Optional[Any] = None, memory_budget_bytes: Optional[int] = None) -> Any:
        """Random access to one persisted object, `path` = "<rank>/<logical path>"; no collectives
        (reference snapshot.py:501-594)."""
        rank_str, _, logical_path = path.partition("/")
        try:
            rank = int(rank_str)
        except ValueError:
            raise RuntimeError(f'The path must be in the form "RANK/LOGICAL_PATH" (got {path}).') from None
        manifest, merged_sd = get_manifest_for_rank(self.metadata, rank)
        entry = merged_sd.get(logical_path) or manifest.get(logical_path)
        if entry is None:
            raise RuntimeError(f'The supplied path "{path}" does not exist in the snapshot\'s manifest.')
        if isinstance(entry, PrimitiveEntry):
            return entry.get_value()
        read_reqs, fut = prepare_read(entry=entry, obj_out=obj_out, buffer_size_limit_bytes=memory_budget_bytes)
        if not knobs.is_batching_disabled():
            # merged reads never exceed the caller's budget, so the tiles stay tiles
            read_reqs = batch_read_requests(read_reqs=read_reqs, max_merged_bytes=memory_budget_bytes)
        event_loop = asyncio.new_event_loop()
        storage = url_to_storage_plugin(self.path, self._storage_options)
        sync_execute_read_reqs(
            read_reqs=read_reqs,
            storage=storage,
            memory_budget_bytes=memory_budget_bytes or _MAX_PER_RANK_MEMORY_BUDGET_BYTES,
            rank=0,
            event_loop=event_loop,
        )
        storage.sync_close(event_loop=event_loop)
        event_loop.close()
        return fut.obj

    def _restore_view(self, rank: int, prefix: str) -> _RestoreView:
        cache = self.__dict__.setdefault("_restore_views", {})
        view = cache.get((rank, prefix))
        if view is None:
            rank_cache = self.__dict__.setdefault("_rank_views", {})
            if rank not in rank_cache:
                rank_cache[rank] = get_manifest_for_rank(self.metadata, rank)
            manifest, merged = rank_cache[rank]
            view = cache[(rank, prefix)] = _RestoreView(manifest, merged, prefix)
        return view

    def _manifest_for_rank(self, rank: int) -> Tuple[Manifest, Dict[str, ShardedTensorEntry]]:
        """Per-rank view of the manifest, computed once per Snapshot object and rank (a restore with several
        statefuls would otherwise re-partition the whole manifest for each of them).  Callers get a shallow
        copy: `_load_stateful` adds elastic sharded entries to it."""
        cache = self.__dict__.setdefault("_rank_views", {})
        if rank not in cache:
            cache[rank] = get_manifest_for_rank(self.metadata, rank)
        manifest, merged = cache[rank]
        return dict(manifest), merged

    # ------------------------------------------------------------------------------------------
    # metadata
    # ------------------------------------------------------------------------------------------
    @property
    def metadata(self) -> SnapshotMetadata:
        if self._metadata is None:
            event_loop = asyncio.new_event_loop()
            storage = url_to_storage_plugin(self.path, self._storage_options)
            read_io = ReadIO(path=SNAPSHOT_METADATA_FNAME)
            storage.sync_read(read_io=read_io, event_loop=event_loop)
            self._metadata = SnapshotMetadata.from_yaml(read_io.buf.getvalue().decode("utf-8"))
            storage.sync_close(event_loop=event_loop)
            event_loop.close()
        return self._metadata

    def get_manifest(self) -> Dict[str, Entry]:
        import copy

        return copy.deepcopy(self.metadata.manifest)

    @staticmethod
    def _write_snapshot_metadata(
        metadata: SnapshotMetadata, storage: StoragePlugin, event_loop: asyncio.AbstractEventLoop
    ) -> None:
        storage.sync_write(
            WriteIO(path=SNAPSHOT_METADATA_FNAME, buf=metadata.to_yaml().encode("utf-8")), event_loop=event_loop
        )


class PendingSnapshot:
    """Handle of an async_take: the background thread finishes D2H + storage I/O, then the ranks
    agree (through the default store, never through collectives in this thread) and rank 0
    commits the metadata (reference snapshot.py:856-947)."""

    DEFAULT_BARRIER_TIMEOUT_S = 1800
    _seq = 0

    def __init__(
        self,
        path: str,
        pending_io_work: PendingIOWork,
        pg: PGWrapper,
        metadata: SnapshotMetadata,
        storage: StoragePlugin,
        event_loop: asyncio.AbstractEventLoop,
        storage_options: Optional[Dict[str, Any]],
    ) -> None:
        self.path = path
        self.pg_obj = pg.pg
        self._storage_options = storage_options
        self.exc_info: Optional[Any] = None
        self._done = False
        self._rank, self._world = pg.get_rank(), pg.get_world_size()
        self._store = None
        if self._world > 1:
            import torch.distributed as dist

            self._store = dist.distributed_c10d._get_default_store()
        PendingSnapshot._seq += 1  # every rank takes the same sequence of async snapshots
        self._key = f"tsb200_{path}_{PendingSnapshot._seq}"
        self.thread = threading.Thread(
            target=self._complete, args=(pending_io_work, metadata, storage, event_loop), daemon=True
        )
        # The thread's first act is to wait for this event, which gives the GIL straight back: the caller of
        # async_take gets its handle before the thread starts queuing a few hundred copies (otherwise the two
        # fight over the GIL for a millisecond or more of the caller's blocked time).
        self._go = threading.Event()
        self.thread.start()
        self._go.set()

    def _complete(self, pending: PendingIOWork, metadata: SnapshotMetadata, storage: StoragePlugin,
                  event_loop: asyncio.AbstractEventLoop) -> None:
        err: Optional[str] = None
        self._go.wait(1.0)
        try:
            pending.sync_complete(event_loop)
        except BaseException as e:  # noqa: B902
            self.exc_info = sys.exc_info()
            err = repr(e)
        try:
            if self._store is not None:
                # two-phase barrier through the store; a failed rank poisons the commit
                self._store.set(f"{self._key}/arrive/{self._rank}", err or "ok")
                if self._rank == 0:
                    states = [self._store.get(f"{self._key}/arrive/{r}").decode() for r in range(self._world)]
                    bad = [s for s in states if s != "ok"]
                    if not bad:
                        Snapshot._write_snapshot_metadata(metadata, storage, event_loop)
                    self._store.set(f"{self._key}/commit", "ok" if not bad else "; ".join(bad))
                verdict = self._store.get(f"{self._key}/commit").decode()
                if verdict != "ok" and self.exc_info is None:
                    raise RuntimeError(f"async snapshot failed on a peer rank: {verdict}")
            elif err is None:
                Snapshot._write_snapshot_metadata(metadata, storage, event_loop)
            storage.sync_close(event_loop=event_loop)
        except BaseException:  # noqa: B902
            if self.exc_info is None:
                self.exc_info = sys.exc_info()
        finally:
            self._done = True

    def wait(self) -> Snapshot:
        self.thread.join()
        if self.exc_info is not None:
            # same error type and message as the reference (snapshot.py:935-941), original as the cause
            import traceback

            formatted = "".join(traceback.format_exception(*self.exc_info))
            raise RuntimeError(
                f"Encountered exception while taking snapshot asynchronously:\n{formatted}"
            ) from self.exc_info[1]
        return Snapshot(path=self.path, pg=self.pg_obj, storage_options=self._storage_options)

    def done(self) -> bool:
        return self._done
