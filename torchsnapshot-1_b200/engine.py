"""Per-device engine: the libtsb200 context plus the small amount of Python state around it.

One engine per (process, device).  It owns the pinned host arena the staged buffers live in and
the device staging arena the pack kernel writes to; both replace per-call allocations in the
reference (torch.cuda.ByteTensor at batcher.py:144, a fresh pageable host tensor per .cpu() /
.to("cpu") at batcher.py:156 and io_preparers/tensor.py:252-254).
"""

import ctypes
import threading
import weakref
from typing import Dict, Optional, Sequence

import numpy as np
import psutil
import torch

from . import _native, knobs
from ._native import ArenaFull, STREAM_D2H, STREAM_H2D, STREAM_PACK  # noqa: F401
from .regions import RegionTuple, regions_to_array

_engines: Dict[int, "Engine"] = {}
_engines_lock = threading.Lock()

_GiB = 1 << 30


class PinnedBuffer:
    """A block of the pinned arena exposed through the buffer protocol.

    `memoryview(buf)` is what BufferStager.stage_buffer hands to the scheduler (reference
    io_types.py:16-26); the block goes back to the arena when the last view is dropped, which is
    when the reference's scheduler reclaims buffer memory (scheduler.py:87-90).
    """

    def __init__(self, engine: "Engine", nbytes: int, grow: bool = True) -> None:
        self.engine = engine
        self.nbytes = nbytes
        # raises ArenaFull when at the bound (or, with grow=False, when new memory would have to be pinned)
        self.ptr = engine.ctx.host_alloc(max(nbytes, 1)) if grow else engine.ctx.host_alloc_nogrow(max(nbytes, 1))
        self._array = (ctypes.c_uint8 * nbytes).from_address(self.ptr)
        self._finalizer = weakref.finalize(self, engine.ctx.host_free, self.ptr)

    def __buffer__(self, flags: int) -> memoryview:  # PEP 688
        return memoryview(self._array)

    def view(self) -> memoryview:
        # the memoryview keeps `self` (and with it the arena block) alive
        return memoryview(self)

    def release(self) -> None:
        """Return the block now (all views must be gone)."""
        self._array = None
        self._finalizer()


class Engine:
    def __init__(self, device: int) -> None:
        if not torch.cuda.is_available():
            raise _native.TsbError("The B200 path needs a CUDA device; there is no CPU fallback.")
        free_hbm, _total = torch.cuda.mem_get_info(device)
        # pinned arena bound: a quarter of the available host memory shared by the GPUs of the box, at most 16 GiB
        # per rank -- room for a batch of the largest objects the writer produces (16 x 512 MiB) to be in flight
        # together, which is what keeps every writer thread busy (scheduler.drain_wave)
        gpus = max(torch.cuda.device_count(), 1)
        host_default = min(16 * _GiB, int(psutil.virtual_memory().available * 0.25) // gpus)
        host_bound = knobs.get_pinned_bytes(host_default)
        dev_bound = knobs.get_device_staging_bytes(int(free_hbm * 0.5))
        self.device = device
        self.ctx = _native.Context(device, host_bound, dev_bound)
        self.host_bound = host_bound
        self.dev_bound = dev_bound
        self.launches = 0  # pack/unpack kernel launches issued (bench.py's gpu_launches)

    # -- launches -------------------------------------------------------------------------------
    def copy_regions(self, regions: Sequence[RegionTuple], stream: int = STREAM_PACK) -> None:
        if not regions:
            return
        self.copy_region_array(regions_to_array(regions), stream)

    def copy_region_array(self, arr: np.ndarray, stream: int = STREAM_PACK) -> None:
        if arr.shape[0] == 0:
            return
        self.ctx.copy_regions(stream, arr)
        self.launches += 1

    def copy_region_array_timed(self, arr: np.ndarray, stream: int = STREAM_PACK):
        """Launch; returns (start, stop) events around the copy kernels only (caller releases)."""
        ev = self.ctx.copy_regions_timed(stream, arr)
        self.launches += 1
        return ev

    def wait_current_stream(self, stream: int) -> None:
        """Order a context stream after the work already queued on torch's current stream."""
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream().cuda_stream
        self.ctx.stream_wait_external(stream, cur)

    def current_stream_wait(self, stream: int) -> None:
        """Order torch's current stream after a context stream."""
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream().cuda_stream
        self.ctx.external_wait_stream(cur, stream)

    def pinned(self, nbytes: int, grow: bool = True) -> PinnedBuffer:
        return PinnedBuffer(self, nbytes, grow)

    def host_in_use(self) -> int:
        return self.ctx.arena_stats(0)[0]

    def dev_in_use(self) -> int:
        return self.ctx.arena_stats(1)[0]


def get_engine(device: Optional[int] = None) -> Engine:
    if device is None:
        device = torch.cuda.current_device()
    with _engines_lock:
        eng = _engines.get(device)
        if eng is None:
            eng = Engine(device)
            _engines[device] = eng
        return eng


def shutdown_engines() -> None:
    with _engines_lock:
        for eng in _engines.values():
            eng.ctx.close()
        _engines.clear()


def buffer_address(buf) -> int:
    """Host address of a writable buffer-protocol object (memoryview, bytearray, ...)."""
    return ctypes.addressof(ctypes.c_char.from_buffer(buf))


def readonly_buffer_address(buf) -> int:
    """Host address of any contiguous buffer, including immutable `bytes`."""
    arr = np.frombuffer(buf, dtype=np.uint8)
    return int(arr.ctypes.data)
