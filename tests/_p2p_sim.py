"""CPU stand-ins that let tests/test_reshard_sim.py run the peer-to-peer restore ORCHESTRATION of
torchsnapshot_b200/reshard.py (ownership, request exchange, per-destination layouts, region tables, error
agreement) without a GPU: ranks are threads of one process, "device memory" is host memory, the copy
"kernel" is the memmove interpreter of tests/_sim.py, IPC handles are addresses.  Test-side only; the CUDA
path itself is covered by tests/test_gpu_multi.py on real GPUs."""

import copy
import threading
from typing import Any, Dict, List

import numpy as np

import _sim


class ThreadGroup:
    """Object collectives between threads (the subset of PGWrapper that reshard.py uses)."""

    def __init__(self, world: int) -> None:
        self.world = world
        self._barrier = threading.Barrier(world)
        self._slots: List[Any] = [None] * world

    def view(self, rank: int) -> "RankView":
        return RankView(self, rank)


class RankView:
    def __init__(self, group: ThreadGroup, rank: int) -> None:
        self.group, self.rank = group, rank

    def get_rank(self) -> int:
        return self.rank

    def get_world_size(self) -> int:
        return self.group.world

    def barrier(self) -> None:
        self.group._barrier.wait(timeout=60)

    def all_gather_object(self, obj_list: List[Any], obj: Any) -> None:
        g = self.group
        g._slots[self.rank] = copy.deepcopy(obj)  # what pickling through a real collective would do
        g._barrier.wait(timeout=60)
        for i in range(g.world):
            obj_list[i] = copy.deepcopy(g._slots[i])
        g._barrier.wait(timeout=60)


class FakePinned:
    def __init__(self, nbytes: int) -> None:
        self._arr = np.empty(max(nbytes, 1), dtype=np.uint8)
        self.ptr = self._arr.ctypes.data
        self.nbytes = nbytes

    def view(self) -> memoryview:
        return memoryview(self._arr)[: self.nbytes]

    def release(self) -> None:
        pass


class FakeCtx:
    """The C-ABI context calls reshard.py makes, on host memory.  One per rank; `registry` is shared, so an
    "IPC handle" can be resolved to the exporting rank's block."""

    def __init__(self, rank: int, registry: Dict[bytes, int], log: List[Any]) -> None:
        self.rank, self.registry, self.log = rank, registry, log
        self.blocks: Dict[int, np.ndarray] = {}
        self.exportable: set = set()
        # peer mappings of every rank, shared through the registry: base address -> number of open mappings
        self.mapped: Dict[int, int] = registry.setdefault(b"__mapped__", {})  # type: ignore[assignment]
        self.fail_exportable = False  # dev_alloc_exportable raises (a rank that is out of HBM)
        self.fail_open = False  # ipc_open raises (a peer allocation that cannot be mapped)

    # -- memory --
    def _alloc(self, nbytes: int) -> int:
        arr = np.full(max(nbytes, 1) + 64, 0xEE, dtype=np.uint8)
        addr = arr.ctypes.data
        self.blocks[addr] = arr
        return addr

    def dev_alloc(self, nbytes: int) -> int:
        return self._alloc(nbytes)

    def dev_alloc_exportable(self, nbytes: int) -> int:
        if self.fail_exportable:
            raise MemoryError("out of memory")
        self.log.append(("alloc_exportable", self.rank, nbytes))
        addr = self._alloc(nbytes)
        self.exportable.add(addr)
        return addr

    def dev_free(self, addr: int) -> None:
        assert not self.mapped.get(addr), "freed while a peer has it mapped"
        del self.blocks[addr]
        self.exportable.discard(addr)

    # -- peer access: the handle names the ALLOCATION, like cudaIpcGetMemHandle --
    def ipc_export(self, dptr: int):
        for base, arr in self.blocks.items():
            if base <= dptr < base + arr.nbytes:
                handle = (b"R%03d" % self.rank) + base.to_bytes(8, "little") + bytes(52)
                self.registry[handle] = base
                self.log.append(("export", self.rank, arr.nbytes - 64, base in self.exportable))
                return handle, dptr - base
        raise AssertionError("export of an address that is not device memory")

    def ipc_open(self, handle: bytes) -> int:
        if self.fail_open:
            raise RuntimeError("invalid device context")
        base = self.registry[handle]
        self.log.append(("open", self.rank, handle[:4]))
        self.mapped[base] = self.mapped.get(base, 0) + 1
        return base

    def ipc_close(self, base: int) -> None:
        self.log.append(("close", self.rank))
        self.mapped[base] -= 1

    # -- streams / events: everything is synchronous here --
    def memcpy_h2d(self, stream: int, dst: int, src: int, nbytes: int) -> None:
        import ctypes

        ctypes.memmove(dst, src, nbytes)

    def event_record(self, stream: int) -> int:
        return 1

    def event_sync(self, ev: int) -> None:
        pass

    def event_release(self, ev: int) -> None:
        pass

    def event_elapsed_ms(self, a: int, b: int) -> float:
        return 0.0

    def stream_wait_event(self, stream: int, ev: int) -> None:
        pass

    def stream_sync(self, stream: int) -> None:
        pass


class FakeEngine:
    def __init__(self, rank: int, registry: Dict[bytes, int], log: List[Any]) -> None:
        self.ctx = FakeCtx(rank, registry, log)
        self.device = rank
        self.launches = 0

    def pinned(self, nbytes: int, grow: bool = True) -> FakePinned:
        return FakePinned(nbytes)

    def wait_current_stream(self, stream: int) -> None:
        pass

    def copy_region_array(self, arr: np.ndarray, stream: int = 0) -> None:
        regions = []
        for rec in arr:
            nd = int(rec["ndim"])
            regions.append((int(rec["src"]), int(rec["dst"]), int(rec["run_bytes"]),
                            tuple(int(x) for x in rec["size"][:nd]),
                            tuple(int(x) for x in rec["src_stride"][:nd]),
                            tuple(int(x) for x in rec["dst_stride"][:nd])))
        _sim.run_regions(regions)
        self.launches += 1

    def copy_region_array_timed(self, arr: np.ndarray, stream: int = 0):
        self.copy_region_array(arr, stream)
        return 1, 1
