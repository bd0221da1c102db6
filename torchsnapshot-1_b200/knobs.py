"""Environment knobs of the hot path; same variable names and defaults as the reference
(torchsnapshot/knobs.py:21-55, scheduler.py:27-30, 46-56) so one environment drives both."""

import os
from contextlib import contextmanager
from typing import Any, Generator

_MAX_CHUNK_SIZE_ENV_VAR = "TORCHSNAPSHOT_MAX_CHUNK_SIZE_BYTES_OVERRIDE"
_MAX_SHARD_SIZE_ENV_VAR = "TORCHSNAPSHOT_MAX_SHARD_SIZE_BYTES_OVERRIDE"
_SLAB_SIZE_THRESHOLD_ENV_VAR = "TORCHSNAPSHOT_SLAB_SIZE_THRESHOLD_BYTES_OVERRIDE"
_DISABLE_BATCHING_ENV_VAR = "TORCHSNAPSHOT_DISABLE_BATCHING"
_MEMORY_BUDGET_ENV_VAR = "TORCHSNAPSHOT_PER_RANK_MEMORY_BUDGET_BYTES"
# B200-path additions (no reference counterpart)
_DEVICE_STAGING_ENV_VAR = "TSB200_DEVICE_STAGING_BYTES"  # bound of the device staging arena
_PINNED_ENV_VAR = "TSB200_PINNED_BYTES"  # bound of the pinned host arena
_IO_THREADS_ENV_VAR = "TSB200_IO_THREADS"  # pwrite/pread workers per file in the fs fast path

_MiB = 1024 * 1024


def _int_env(name: str, default: int) -> int:
    val = os.environ.get(name)
    return int(val) if val is not None else default


def get_max_chunk_size_bytes() -> int:
    return _int_env(_MAX_CHUNK_SIZE_ENV_VAR, 512 * _MiB)


def get_max_shard_size_bytes() -> int:
    return _int_env(_MAX_SHARD_SIZE_ENV_VAR, 512 * _MiB)


def get_slab_size_threshold_bytes() -> int:
    return _int_env(_SLAB_SIZE_THRESHOLD_ENV_VAR, 128 * _MiB)


def is_batching_disabled() -> bool:
    return os.getenv(_DISABLE_BATCHING_ENV_VAR, "False").lower() in ("true", "1")


def get_device_staging_bytes(default: int) -> int:
    return _int_env(_DEVICE_STAGING_ENV_VAR, default)


def get_pinned_bytes(default: int) -> int:
    return _int_env(_PINNED_ENV_VAR, default)


def is_p2p_reshard_enabled() -> bool:
    """Peer-to-peer reshard-restore of sharded CUDA tensors (reshard.py); TSB200_P2P_RESHARD=0 forces the
    reference's storage-only path."""
    return os.getenv("TSB200_P2P_RESHARD", "1").lower() not in ("0", "false")


def is_p2p_scatter_enabled() -> bool:
    """Reshards (units whose bytes are partitioned among the ranks): the owner packs one buffer per
    destination rank and the destination maps only that buffer (reshard.py); TSB200_P2P_SCATTER=0 makes every
    destination map the owner's whole staging block instead (the replicated-state scheme)."""
    return os.getenv("TSB200_P2P_SCATTER", "1").lower() not in ("0", "false")


def trim_arenas_after_use() -> bool:
    """TSB200_TRIM_ARENAS=1: give the device staging and pinned host segments back to the driver when a
    take / restore has drained (a trainer short of HBM wants its memory back; the next take then pays for
    cudaMalloc / cudaHostAlloc again).  Default: keep them for the next checkpoint."""
    return os.getenv("TSB200_TRIM_ARENAS", "0").lower() in ("1", "true")


def is_fs_direct_enabled() -> bool:
    """TSB200_FS_DIRECT=1: the FS plugin moves the 4 KiB-aligned body of every file transfer with O_DIRECT
    (no page-cache copy, no dirty-page throttling) -- for checkpoints onto real block devices."""
    return os.getenv("TSB200_FS_DIRECT", "0").lower() in ("1", "true")


def get_io_threads() -> int:
    return _int_env(_IO_THREADS_ENV_VAR, 0)


@contextmanager
def _override(name: str, value: Any) -> Generator[None, None, None]:
    prev = os.environ.get(name)
    os.environ[name] = str(value)
    try:
        yield
    finally:
        if prev is None:
            del os.environ[name]
        else:
            os.environ[name] = prev


def override_max_chunk_size_bytes(n: int):
    return _override(_MAX_CHUNK_SIZE_ENV_VAR, n)


def override_max_shard_size_bytes(n: int):
    return _override(_MAX_SHARD_SIZE_ENV_VAR, n)


def override_slab_size_threshold_bytes(n: int):
    # the reference's helper of this name sets the *shard* variable by mistake
    # (torchsnapshot/knobs.py:92-96); this one sets the slab threshold it is named after.
    return _override(_SLAB_SIZE_THRESHOLD_ENV_VAR, n)


def override_is_batching_disabled(disabled: bool):
    return _override(_DISABLE_BATCHING_ENV_VAR, disabled)
