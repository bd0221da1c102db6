"""ctypes binding of libtsb200.so (the C-ABI declared in include/tsb200.h).

The product path has no CPU fallback: if the shared library is missing or no sm_100 device is
visible, loading / context creation raises.  Every call releases the GIL (ctypes.CDLL), so
blocking waits may run in executor threads the way the reference runs its blocking D2H copies
(torchsnapshot/io_preparers/tensor.py:216-222, 252-254).
"""

import ctypes
import os
import threading
from typing import Any, Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libtsb200.so")

TSB_OK = 0
TSB_EAGAIN = 1
TSB_MAX_DIMS = 4

STREAM_PACK = 0
STREAM_D2H = 1
STREAM_H2D = 2
STREAM_AUX = 3

IPC_HANDLE_BYTES = 64

# numpy mirror of `struct TsbRegion` (include/tsb200.h); 128 bytes
REGION_DTYPE = np.dtype(
    [
        ("src", "<u8"),
        ("dst", "<u8"),
        ("run_bytes", "<u8"),
        ("ndim", "<u4"),
        ("reserved", "<u4"),
        ("size", "<u8", (TSB_MAX_DIMS,)),
        ("src_stride", "<i8", (TSB_MAX_DIMS,)),
        ("dst_stride", "<i8", (TSB_MAX_DIMS,)),
    ],
    align=False,
)
assert REGION_DTYPE.itemsize == 128

# every symbol include/tsb200.h declares: (name, restype, argtypes)
_u64 = ctypes.c_uint64
_p = ctypes.c_void_p
_i = ctypes.c_int
_SIGNATURES = {
    "tsb_version": (_i, []),
    "tsb_last_error": (ctypes.c_char_p, []),
    "tsb_device_count": (_i, [ctypes.POINTER(_i)]),
    "tsb_ctx_create": (_i, [_i, _u64, _u64, ctypes.POINTER(_p)]),
    "tsb_ctx_destroy": (_i, [_p]),
    "tsb_ctx_device": (_i, [_p, ctypes.POINTER(_i)]),
    "tsb_ctx_sm_count": (_i, [_p, ctypes.POINTER(_i)]),
    "tsb_ctx_counters": (_i, [_p, ctypes.POINTER(_u64), ctypes.POINTER(_u64), ctypes.POINTER(_u64), ctypes.POINTER(_u64)]),
    "tsb_dev_alloc": (_i, [_p, _u64, ctypes.POINTER(_u64)]),
    "tsb_dev_alloc_exportable": (_i, [_p, _u64, ctypes.POINTER(_u64)]),
    "tsb_dev_free": (_i, [_p, _u64]),
    "tsb_host_alloc": (_i, [_p, _u64, ctypes.POINTER(_p)]),
    "tsb_host_alloc_nogrow": (_i, [_p, _u64, ctypes.POINTER(_p)]),
    "tsb_host_free": (_i, [_p, _p]),
    "tsb_arena_stats": (_i, [_p, _i, ctypes.POINTER(_u64), ctypes.POINTER(_u64), ctypes.POINTER(_u64)]),
    "tsb_arena_trim": (_i, [_p]),
    "tsb_event_record": (_i, [_p, _i, ctypes.POINTER(_u64)]),
    "tsb_stream_wait_event": (_i, [_p, _i, _u64]),
    "tsb_event_query": (_i, [_p, _u64]),
    "tsb_event_sync": (_i, [_p, _u64]),
    "tsb_event_elapsed_ms": (_i, [_p, _u64, _u64, ctypes.POINTER(ctypes.c_float)]),
    "tsb_event_release": (_i, [_p, _u64]),
    "tsb_stream_wait_external": (_i, [_p, _i, _u64]),
    "tsb_external_wait_stream": (_i, [_p, _u64, _i]),
    "tsb_stream_sync": (_i, [_p, _i]),
    "tsb_copy_regions": (_i, [_p, _i, _p, ctypes.c_uint32]),
    "tsb_copy_regions_timed": (_i, [_p, _i, _p, ctypes.c_uint32, ctypes.POINTER(_u64), ctypes.POINTER(_u64)]),
    "tsb_copy_regions_static": (_i, [_i, _u64, _p, ctypes.c_uint32]),
    "tsb_memcpy_d2h": (_i, [_p, _i, _p, _u64, _u64]),
    "tsb_memcpy_h2d": (_i, [_p, _i, _u64, _p, _u64]),
    "tsb_memcpy_d2h_interleaved": (_i, [_p, _i, ctypes.c_uint32, _p, _p, _p, _u64, _p]),
    "tsb_file_write_streamed": (_i, [_p, ctypes.c_char_p, _p, _u64, _p, _p, ctypes.c_uint32]),
    "tsb_memcpy_d2d": (_i, [_p, _i, _u64, _u64, _u64]),
    "tsb_host_memcpy": (_i, [_p, _p, _u64, _i]),
    "tsb_host_gather": (_i, [_p, _p, _p, _u64, _i]),
    "tsb_ipc_export": (_i, [_p, _u64, _p, ctypes.POINTER(_u64)]),
    "tsb_ipc_open": (_i, [_p, _p, ctypes.POINTER(_u64)]),
    "tsb_ipc_close": (_i, [_p, _u64]),
    "tsb_peer_enable": (_i, [_p, _i]),
    "tsb_file_write": (_i, [ctypes.c_char_p, _p, _u64, _i, ctypes.c_uint32]),
    "tsb_file_read": (_i, [ctypes.c_char_p, _u64, _p, _u64, _i, ctypes.c_uint32]),
}

_lib: Optional[ctypes.CDLL] = None
_lib_lock = threading.Lock()


class TsbError(RuntimeError):
    """A nonzero return from libtsb200 (message from tsb_last_error)."""


class ArenaFull(TsbError):
    """TSB_EAGAIN: an arena is at its bound; retry after a release."""


def declared_symbols():
    return sorted(_SIGNATURES)


def load_library(path: Optional[str] = None) -> ctypes.CDLL:
    """Load libtsb200.so and bind every declared symbol; raises if anything is missing."""
    global _lib
    with _lib_lock:
        if _lib is not None and path is None:
            return _lib
        lib_path = path or LIB_PATH
        if not os.path.exists(lib_path):
            raise TsbError(
                f"{lib_path} not found: build it with `make -C {os.path.dirname(lib_path)}` "
                "(or __graft_entry__.build()). There is no CPU fallback for the GPU path."
            )
        lib = ctypes.CDLL(lib_path)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype = restype
            fn.argtypes = argtypes
        if path is None:
            _lib = lib
        return lib


def _check(lib: ctypes.CDLL, rc: int) -> None:
    if rc == TSB_OK:
        return
    msg = lib.tsb_last_error().decode("utf-8", "replace")
    if rc == TSB_EAGAIN:
        raise ArenaFull(msg)
    raise TsbError(f"libtsb200 error {rc}: {msg}")


class Context:
    """Thin object wrapper over a TsbCtx*; one per (process, device)."""

    def __init__(self, device: int, host_arena_max: int = 0, dev_arena_max: int = 0) -> None:
        self.lib = load_library()
        handle = _p()
        _check(self.lib, self.lib.tsb_ctx_create(device, host_arena_max, dev_arena_max, ctypes.byref(handle)))
        self._h = handle
        self.device = device
        sms = _i()
        _check(self.lib, self.lib.tsb_ctx_sm_count(self._h, ctypes.byref(sms)))
        self.sm_count = sms.value

    def close(self) -> None:
        if self._h is not None and self._h.value:
            self.lib.tsb_ctx_destroy(self._h)
            self._h = None

    def counters(self) -> dict:
        a, b, c, d = _u64(), _u64(), _u64(), _u64()
        _check(self.lib, self.lib.tsb_ctx_counters(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c), ctypes.byref(d)))
        return {"kernel_launches": a.value, "copy_calls": b.value, "desc_upload_bytes": c.value,
                "kernel_payload_bytes": d.value}

    # -- arenas
    def dev_alloc(self, nbytes: int) -> int:
        out = _u64()
        _check(self.lib, self.lib.tsb_dev_alloc(self._h, nbytes, ctypes.byref(out)))
        return out.value

    def dev_alloc_exportable(self, nbytes: int) -> int:
        """A device block that is its own driver allocation (what a peer maps is exactly this block)."""
        out = _u64()
        _check(self.lib, self.lib.tsb_dev_alloc_exportable(self._h, nbytes, ctypes.byref(out)))
        return out.value

    def dev_free(self, dptr: int) -> None:
        _check(self.lib, self.lib.tsb_dev_free(self._h, dptr))

    def host_alloc(self, nbytes: int) -> int:
        out = _p()
        _check(self.lib, self.lib.tsb_host_alloc(self._h, nbytes, ctypes.byref(out)))
        return out.value or 0

    def host_alloc_nogrow(self, nbytes: int) -> int:
        """Pinned block out of the segments the arena already holds; ArenaFull if that takes new pinned memory."""
        out = _p()
        _check(self.lib, self.lib.tsb_host_alloc_nogrow(self._h, nbytes, ctypes.byref(out)))
        return out.value or 0

    def host_free(self, hptr: int) -> None:
        _check(self.lib, self.lib.tsb_host_free(self._h, hptr))

    def arena_stats(self, which: int) -> Tuple[int, int, int]:
        a, b, c = _u64(), _u64(), _u64()
        _check(self.lib, self.lib.tsb_arena_stats(self._h, which, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return a.value, b.value, c.value

    def arena_trim(self) -> None:
        _check(self.lib, self.lib.tsb_arena_trim(self._h))

    # -- events / streams
    def event_record(self, stream: int) -> int:
        ev = _u64()
        _check(self.lib, self.lib.tsb_event_record(self._h, stream, ctypes.byref(ev)))
        return ev.value

    def stream_wait_event(self, stream: int, ev: int) -> None:
        _check(self.lib, self.lib.tsb_stream_wait_event(self._h, stream, ev))

    def event_done(self, ev: int) -> bool:
        rc = self.lib.tsb_event_query(self._h, ev)
        if rc == TSB_EAGAIN:
            return False
        _check(self.lib, rc)
        return True

    def event_sync(self, ev: int) -> None:
        _check(self.lib, self.lib.tsb_event_sync(self._h, ev))

    def event_elapsed_ms(self, start: int, stop: int) -> float:
        ms = ctypes.c_float()
        _check(self.lib, self.lib.tsb_event_elapsed_ms(self._h, start, stop, ctypes.byref(ms)))
        return ms.value

    def event_release(self, ev: int) -> None:
        _check(self.lib, self.lib.tsb_event_release(self._h, ev))

    def stream_wait_external(self, stream: int, cuda_stream: int) -> None:
        _check(self.lib, self.lib.tsb_stream_wait_external(self._h, stream, cuda_stream))

    def external_wait_stream(self, cuda_stream: int, stream: int) -> None:
        _check(self.lib, self.lib.tsb_external_wait_stream(self._h, cuda_stream, stream))

    def stream_sync(self, stream: int) -> None:
        _check(self.lib, self.lib.tsb_stream_sync(self._h, stream))

    # -- hot path
    def copy_regions(self, stream: int, regions: np.ndarray) -> None:
        if regions.dtype != REGION_DTYPE or not regions.flags["C_CONTIGUOUS"]:
            raise ValueError("regions must be a C-contiguous array of REGION_DTYPE")
        n = int(regions.shape[0])
        if n == 0:
            return
        _check(self.lib, self.lib.tsb_copy_regions(self._h, stream, regions.ctypes.data, n))

    def copy_regions_timed(self, stream: int, regions: np.ndarray) -> Tuple[int, int]:
        """Launch and return (start, stop) events that bracket only the copy kernels."""
        if regions.dtype != REGION_DTYPE or not regions.flags["C_CONTIGUOUS"]:
            raise ValueError("regions must be a C-contiguous array of REGION_DTYPE")
        a, b = _u64(), _u64()
        _check(
            self.lib,
            self.lib.tsb_copy_regions_timed(
                self._h, stream, regions.ctypes.data, int(regions.shape[0]), ctypes.byref(a), ctypes.byref(b)
            ),
        )
        return a.value, b.value

    def memcpy_d2h(self, stream: int, hptr: int, dptr: int, nbytes: int) -> None:
        _check(self.lib, self.lib.tsb_memcpy_d2h(self._h, stream, hptr, dptr, nbytes))

    def memcpy_d2h_interleaved(self, stream: int, hptrs: Any, dptrs: Any, nbytes: Any, chunk: int) -> Any:
        """Interleaved chunked D2H copies of several buffers; returns, per buffer, (event handles, end offsets) as
        uint64 numpy arrays (one event per chunk; the caller releases them)."""
        h = np.ascontiguousarray(hptrs, dtype=np.uint64)
        d = np.ascontiguousarray(dptrs, dtype=np.uint64)
        n = np.ascontiguousarray(nbytes, dtype=np.uint64)
        counts = [(int(x) + chunk - 1) // chunk for x in n]
        events = np.zeros(max(sum(counts), 1), dtype=np.uint64)
        _check(self.lib, self.lib.tsb_memcpy_d2h_interleaved(self._h, stream, len(counts), h.ctypes.data, d.ctypes.data,
                                                             n.ctypes.data, chunk, events.ctypes.data))
        out, at = [], 0
        for c, size in zip(counts, n):
            ends = np.minimum(np.arange(1, c + 1, dtype=np.uint64) * np.uint64(chunk), size)
            out.append((events[at : at + c].copy(), ends))
            at += c
        return out

    def file_write_streamed(self, path: str, hptr: int, nbytes: int, events: Any, ends: Any) -> None:
        """Write a pinned buffer while it is being filled (events[i] done => bytes [0, ends[i]) valid)."""
        ev = np.ascontiguousarray(events, dtype=np.uint64)
        en = np.ascontiguousarray(ends, dtype=np.uint64)
        _check(self.lib, self.lib.tsb_file_write_streamed(self._h, os.fsencode(path), hptr, nbytes, ev.ctypes.data,
                                                          en.ctypes.data, int(ev.shape[0])))

    def memcpy_h2d(self, stream: int, dptr: int, hptr: int, nbytes: int) -> None:
        _check(self.lib, self.lib.tsb_memcpy_h2d(self._h, stream, dptr, hptr, nbytes))

    def memcpy_d2d(self, stream: int, dst: int, src: int, nbytes: int) -> None:
        _check(self.lib, self.lib.tsb_memcpy_d2d(self._h, stream, dst, src, nbytes))

    # -- peer access
    def ipc_export(self, dptr: int) -> Tuple[bytes, int]:
        buf = (ctypes.c_uint8 * IPC_HANDLE_BYTES)()
        off = _u64()
        _check(self.lib, self.lib.tsb_ipc_export(self._h, dptr, buf, ctypes.byref(off)))
        return bytes(buf), off.value

    def ipc_open(self, handle: bytes) -> int:
        buf = (ctypes.c_uint8 * IPC_HANDLE_BYTES).from_buffer_copy(handle)
        base = _u64()
        _check(self.lib, self.lib.tsb_ipc_open(self._h, buf, ctypes.byref(base)))
        return base.value

    def ipc_close(self, base: int) -> None:
        _check(self.lib, self.lib.tsb_ipc_close(self._h, base))

    def peer_enable(self, peer_device: int) -> None:
        _check(self.lib, self.lib.tsb_peer_enable(self._h, peer_device))


def copy_regions_static(device: int, cuda_stream: int, regions: np.ndarray) -> None:
    """tsb_copy_regions_static: the multi-region copy for callers that may not allocate (no context,
    scratch in a fixed array inside the library; limits in include/tsb200.h).  Ordered on `cuda_stream`."""
    if regions.dtype != REGION_DTYPE or not regions.flags["C_CONTIGUOUS"]:
        raise ValueError("regions must be a C-contiguous array of REGION_DTYPE")
    lib = load_library()
    _check(lib, lib.tsb_copy_regions_static(device, cuda_stream, regions.ctypes.data, int(regions.shape[0])))


def host_memcpy(dst: int, src: int, nbytes: int, nthreads: int = 0) -> None:
    lib = load_library()
    _check(lib, lib.tsb_host_memcpy(dst, src, nbytes, nthreads))


def host_gather(dst: Any, src: Any, nbytes: Any, nthreads: int = 0) -> None:
    """dst[i] <- src[i] (nbytes[i] bytes) for every i: three equally long uint64 numpy arrays of host addresses
    and lengths.  Runs without the GIL, on `nthreads` workers balanced by bytes."""
    import numpy as np

    d = np.ascontiguousarray(dst, dtype=np.uint64)
    s_ = np.ascontiguousarray(src, dtype=np.uint64)
    n = np.ascontiguousarray(nbytes, dtype=np.uint64)
    if not (d.shape == s_.shape == n.shape and d.ndim == 1):
        raise ValueError("host_gather: dst, src and nbytes must be 1-d arrays of one length")
    lib = load_library()
    _check(lib, lib.tsb_host_gather(d.ctypes.data, s_.ctypes.data, n.ctypes.data, d.shape[0], nthreads))


FILE_DIRECT = 1  # TSB_FILE_DIRECT


def file_write(path: str, hptr: int, nbytes: int, nthreads: int = 0, flags: int = 0) -> None:
    lib = load_library()
    _check(lib, lib.tsb_file_write(os.fsencode(path), hptr, nbytes, nthreads, flags))


def file_read(path: str, offset: int, hptr: int, nbytes: int, nthreads: int = 0, flags: int = 0) -> None:
    lib = load_library()
    _check(lib, lib.tsb_file_read(os.fsencode(path), offset, hptr, nbytes, nthreads, flags))
