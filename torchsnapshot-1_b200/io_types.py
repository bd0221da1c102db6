"""The drop-in boundary of the hot path, as Python types.

These are the reference's interface names with the reference's fields and call signatures
(torchsnapshot/io_types.py:16-111), because everything on either side of the path is written against
them: planners emit `WriteReq` / `ReadReq`, schedulers drive `BufferStager` / `BufferConsumer`, and the
sink is any `StoragePlugin`.  Objects are matched structurally (duck typing) everywhere in this package,
so stagers, consumers and plugins made by the reference interoperate without subclassing these.
"""

import asyncio
import io
from abc import ABC, abstractmethod
from concurrent.futures import Executor
from dataclasses import dataclass, field
from typing import Any, Awaitable, Generic, Optional, Tuple, TypeVar, Union

# what a stager hands to the scheduler and the scheduler to StoragePlugin.write: immutable bytes, or a
# memoryview (on this path: a view of a pinned arena block)
BufferType = Union[bytes, memoryview]
ByteRange = Tuple[int, int]
T = TypeVar("T")


# ---- producers and consumers of one storage object ------------------------------------------------
class BufferStager(ABC):
    """Produces the bytes of one storage object.  `get_staging_cost_bytes` is what the scheduler charges
    against the memory budget while the buffer is alive (reference io_types.py:19-26)."""

    @abstractmethod
    def get_staging_cost_bytes(self) -> int:
        raise NotImplementedError

    @abstractmethod
    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        raise NotImplementedError


class BufferConsumer(ABC):
    """Consumes the bytes of one storage object, or of the byte range its ReadReq names (reference
    io_types.py:35-44)."""

    @abstractmethod
    def get_consuming_cost_bytes(self) -> int:
        raise NotImplementedError

    @abstractmethod
    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        raise NotImplementedError


@dataclass
class WriteReq:
    path: str
    buffer_stager: BufferStager


@dataclass
class ReadReq:
    path: str
    buffer_consumer: BufferConsumer
    byte_range: Optional[ByteRange] = None  # None = the whole object


@dataclass
class Future(Generic[T]):
    """Where a read deposits the object it produced (the reference's minimal future, io_types.py:57-59)."""

    obj: Optional[T] = None


# ---- the storage sink --------------------------------------------------------------------------------
@dataclass
class WriteIO:
    path: str
    buf: BufferType


@dataclass
class ReadIO:
    path: str
    buf: io.BytesIO = field(default_factory=io.BytesIO)
    byte_range: Optional[ByteRange] = None


def _run_sync(coro: Awaitable[Any], event_loop: Optional[asyncio.AbstractEventLoop]) -> None:
    loop = event_loop if event_loop is not None else asyncio.new_event_loop()
    loop.run_until_complete(coro)


class StoragePlugin(ABC):
    """Unchanged sink ABI (reference io_types.py:75-111): four coroutines plus blocking conveniences.
    A plugin may additionally offer `async read_into(path, byte_range, out: memoryview)`; the scheduler
    then reads straight into pinned memory (see storage_plugins/fs.py)."""

    @abstractmethod
    async def write(self, write_io: WriteIO) -> None:
        raise NotImplementedError

    @abstractmethod
    async def read(self, read_io: ReadIO) -> None:
        raise NotImplementedError

    @abstractmethod
    async def delete(self, path: str) -> None:
        raise NotImplementedError

    @abstractmethod
    async def close(self) -> None:
        raise NotImplementedError

    def sync_write(self, write_io: WriteIO, event_loop: Optional[asyncio.AbstractEventLoop] = None) -> None:
        _run_sync(self.write(write_io=write_io), event_loop)

    def sync_read(self, read_io: ReadIO, event_loop: Optional[asyncio.AbstractEventLoop] = None) -> None:
        _run_sync(self.read(read_io=read_io), event_loop)

    def sync_close(self, event_loop: Optional[asyncio.AbstractEventLoop] = None) -> None:
        _run_sync(self.close(), event_loop)
