"""Local-filesystem storage plugin.

Same StoragePlugin ABI and file layout as the reference plugin (torchsnapshot/storage_plugins/
fs.py:18-56): one file per write request under `root`, whole-file or ranged reads into a BytesIO.
Differences are below the ABI: file I/O runs in libtsb200's pwrite/pread workers instead of
aiofiles, and the optional `read_into` lets the scheduler read straight into pinned memory
 -- plugins without it are served through `read` as in the reference.
"""

import asyncio
import io
import os
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Any, Dict, Optional, Set, Tuple

from .. import _native, knobs
from ..engine import buffer_address, readonly_buffer_address
from ..io_types import ReadIO, StoragePlugin, WriteIO


_executors: Dict[Tuple[int, int], ThreadPoolExecutor] = {}
_executors_lock = threading.Lock()


def _shared_executor(workers: int) -> ThreadPoolExecutor:
    key = (os.getpid(), workers)  # threads do not survive a fork: a forked child gets a pool of its own
    with _executors_lock:
        ex = _executors.get(key)
        if ex is None:
            ex = _executors[key] = ThreadPoolExecutor(max_workers=workers, thread_name_prefix="tsb-fs")
        return ex


class FSStoragePlugin(StoragePlugin):
    def __init__(self, root: str, storage_options: Optional[Dict[str, Any]] = None) -> None:
        self.root = root
        self._dir_cache: Set[str] = set()
        # one pool of I/O threads per process and size: every Snapshot call makes a plugin of its own, and
        # starting (then joining) 32 threads per call is milliseconds on the critical path of a take / restore
        self._executor = _shared_executor(int(os.environ.get("TSB200_IO_CONCURRENCY", "32")))
        # TSB200_IO_THREADS pins the pwrite/pread workers per file; by default they follow the load
        self._io_threads = knobs.get_io_threads()
        from ..scheduler import io_concurrency

        self._thread_budget = io_concurrency()
        self._inflight = 0
        self._lock = threading.Lock()
        # TSB200_FS_DIRECT=1: page-cache bypass (O_DIRECT) for the aligned body of every transfer -- for real
        # block devices; tmpfs has no use for it
        self._flags = _native.FILE_DIRECT if knobs.is_fs_direct_enabled() else 0
        if self._flags or type(self).write is not FSStoragePlugin.write:
            # O_DIRECT transfers want whole, aligned buffers; and a subclass that customises `write` sees every
            # write there, as whole buffers
            self.write_streamed = None  # type: ignore[assignment]

    def _prepare(self, rel: str) -> str:
        path = os.path.join(self.root, rel)
        parent = os.path.dirname(path)
        if parent not in self._dir_cache:
            os.makedirs(parent, exist_ok=True)
            self._dir_cache.add(parent)
        return path

    def _threads_for_one_op(self) -> int:
        """Workers for one file operation: the rank's I/O thread budget split over the operations in
        flight, so 32 concurrent slab writes use one thread each and a lone 512 MiB read uses eight."""
        if self._io_threads:
            return self._io_threads
        with self._lock:
            inflight = max(self._inflight, 1)
        return max(1, min(int(os.environ.get("TSB200_IO_THREADS_MAX", "8")), self._thread_budget // inflight))

    def _write_sync(self, path: str, buf: Any) -> None:
        n = len(buf)
        addr = readonly_buffer_address(buf) if n else 0
        with self._lock:
            self._inflight += 1
        try:
            _native.file_write(path, addr, n, self._threads_for_one_op(), self._flags)
        finally:
            with self._lock:
                self._inflight -= 1

    async def write(self, write_io: WriteIO) -> None:
        path = self._prepare(write_io.path)
        await asyncio.get_running_loop().run_in_executor(self._executor, self._write_sync, path, write_io.buf)

    def _write_streamed_sync(self, path: str, buf: Any, wait_more: Any) -> None:
        n = len(buf)
        native = getattr(wait_more, "native", None)
        if native is not None and n:
            ctx, events, ends = native  # wait for the copies and write, all inside libtsb200
            ctx.file_write_streamed(path, readonly_buffer_address(buf), n, events, ends)
            return
        fd = os.open(path, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
        try:
            view = memoryview(buf)
            done = 0
            while done < n:
                upto = wait_more(done)  # blocks until more of the buffer is valid
                while done < upto:
                    done += os.pwrite(fd, view[done:upto], done)
        finally:
            os.close(fd)

    async def write_streamed(self, path: str, buf: Any, wait_more: Any) -> None:
        """Optional extension next to the ABI (like `read_into`): write a buffer whose bytes become valid front to
        back while the call runs.  `wait_more(done) -> upto` blocks until bytes beyond `done` are valid.  One
        writer thread per file follows the device->host copies, so a file is on storage shortly after its
        last byte has left the GPU.  Same file as `write` produces."""
        full = self._prepare(path)
        await asyncio.get_running_loop().run_in_executor(self._executor, self._write_streamed_sync, full, buf, wait_more)

    def _read_sync(self, path: str, byte_range: Optional[Tuple[int, int]], out: Optional[memoryview]) -> bytearray:
        if byte_range is None:
            offset, n = 0, os.path.getsize(path)
        else:
            offset, n = byte_range[0], byte_range[1] - byte_range[0]
        if out is None:
            dst = bytearray(n)
        else:
            if len(out) != n:
                raise AssertionError(f"{path}: destination holds {len(out)} bytes, the read is {n} bytes.")
            dst = out
        if n:
            with self._lock:
                self._inflight += 1
            try:
                _native.file_read(path, offset, buffer_address(dst), n, self._threads_for_one_op(), self._flags)
            finally:
                with self._lock:
                    self._inflight -= 1
        return dst

    async def read(self, read_io: ReadIO) -> None:
        path = os.path.join(self.root, read_io.path)
        data = await asyncio.get_running_loop().run_in_executor(
            self._executor, self._read_sync, path, read_io.byte_range, None
        )
        read_io.buf = io.BytesIO(data)

    async def read_into(self, path: str, byte_range: Optional[Tuple[int, int]], out: memoryview) -> None:
        """Read a file (range) directly into a caller-provided writable buffer."""
        full = os.path.join(self.root, path)
        await asyncio.get_running_loop().run_in_executor(self._executor, self._read_sync, full, byte_range, out)

    async def delete(self, path: str) -> None:
        os.remove(os.path.join(self.root, path))

    async def close(self) -> None:
        return None  # the I/O threads are shared by the plugins of the process; every operation has been awaited
