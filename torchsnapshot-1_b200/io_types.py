"""The drop-in boundary types of the hot path.

Mirror of the reference's torchsnapshot/io_types.py:16-111 (same names, fields, argument
meaning): BufferStager / BufferConsumer are what the schedulers drive, WriteReq / ReadReq what
the planners emit, StoragePlugin the unchanged sink.  Any object with these methods works
(duck typing), so reference stagers and third-party storage plugins interoperate.
"""

import abc
import asyncio
import io
from concurrent.futures import Executor
from dataclasses import dataclass, field
from typing import Generic, Optional, Tuple, TypeVar, Union

BufferType = Union[bytes, memoryview]


class BufferStager(abc.ABC):
    """Produces the bytes of one storage object (reference io_types.py:19-26)."""

    @abc.abstractmethod
    async def stage_buffer(self, executor: Optional[Executor] = None) -> BufferType:
        ...

    @abc.abstractmethod
    def get_staging_cost_bytes(self) -> int:
        ...


@dataclass
class WriteReq:
    path: str
    buffer_stager: BufferStager


class BufferConsumer(abc.ABC):
    """Consumes the bytes of one storage object or byte range (reference io_types.py:35-44)."""

    @abc.abstractmethod
    async def consume_buffer(self, buf: BufferType, executor: Optional[Executor] = None) -> None:
        ...

    @abc.abstractmethod
    def get_consuming_cost_bytes(self) -> int:
        ...


@dataclass
class ReadReq:
    path: str
    buffer_consumer: BufferConsumer
    byte_range: Optional[Tuple[int, int]] = None


T = TypeVar("T")


@dataclass
class Future(Generic[T]):
    obj: Optional[T] = None


@dataclass
class WriteIO:
    path: str
    buf: BufferType


@dataclass
class ReadIO:
    path: str
    buf: io.BytesIO = field(default_factory=io.BytesIO)
    byte_range: Optional[Tuple[int, int]] = None


class StoragePlugin(abc.ABC):
    """Unchanged storage sink ABI (reference io_types.py:75-111)."""

    @abc.abstractmethod
    async def write(self, write_io: WriteIO) -> None:
        ...

    @abc.abstractmethod
    async def read(self, read_io: ReadIO) -> None:
        ...

    @abc.abstractmethod
    async def delete(self, path: str) -> None:
        ...

    @abc.abstractmethod
    async def close(self) -> None:
        ...

    def sync_write(self, write_io: WriteIO, event_loop: Optional[asyncio.AbstractEventLoop] = None) -> None:
        (event_loop or asyncio.new_event_loop()).run_until_complete(self.write(write_io=write_io))

    def sync_read(self, read_io: ReadIO, event_loop: Optional[asyncio.AbstractEventLoop] = None) -> None:
        (event_loop or asyncio.new_event_loop()).run_until_complete(self.read(read_io=read_io))

    def sync_close(self, event_loop: Optional[asyncio.AbstractEventLoop] = None) -> None:
        (event_loop or asyncio.new_event_loop()).run_until_complete(self.close())
