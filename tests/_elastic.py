"""Sharded state whose every element is a function of its GLOBAL index, and the world-size-change round trip
built on it: taken by N ranks, restored by N/2, taken again by N/2, restored by N (BASELINE configs 3 and 4 in
shape: FSDP-style flat 1-D parameters and row-wise embedding tables; the reference's workloads are
benchmarks/fsdp/main.py:75-103 and benchmarks/torchrec/main.py:115-151, its reshard plan
io_preparers/sharded_tensor.py:196-268 after manifest_ops.py:118-176).

Because the expected bytes of any box can be recomputed anywhere, every restored shard is compared bit for
bit with the source data on whatever rank and in whatever cut it ends up -- no reference copy travels.

Device-agnostic: the gloo tests run it on CPU tensors (tests/test_dist_cpu.py), tests/test_gpu_multi.py and
bench.py's pre-flight on CUDA tensors over NCCL.  Test infrastructure; the product never imports it.
"""

import os
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

_CHUNK = 1 << 24  # elements per step of fill / compare: bounds the int64 index temporaries


def _as_int(dtype: torch.dtype) -> torch.dtype:
    return {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}[torch.empty(0, dtype=dtype).element_size()]


def pattern(idx: torch.Tensor, seed: int, dtype: torch.dtype) -> torch.Tensor:
    """Bit pattern of element `idx` (int64 global linear indices) of object `seed`, as `dtype`."""
    h = (idx * 0x9E3779B1 + (seed + 1) * 0x85EBCA6B) & 0xFFFFFFFF
    h = h ^ (h >> 15)
    it = _as_int(dtype)
    mask = {torch.uint8: 0xFF, torch.int16: 0x7FFF, torch.int32: 0x7FFFFFFF, torch.int64: 0xFFFFFFFF}[it]
    return (h & mask).to(it).view(dtype)


def _box_rows(global_shape: Sequence[int], offsets: Sequence[int], sizes: Sequence[int], device: Any):
    """Yields (row slice of the box, int64 global linear indices of those rows): the box is walked along its
    first dimension in steps of about _CHUNK elements."""
    shape = list(global_shape) if len(global_shape) else [1]
    offsets = list(offsets) if len(offsets) else [0]
    sizes = list(sizes) if len(sizes) else [1]
    strides = [1] * len(shape)
    for d in range(len(shape) - 2, -1, -1):
        strides[d] = strides[d + 1] * shape[d + 1]
    inner = 1
    for s in sizes[1:]:
        inner *= s
    step = max(_CHUNK // max(inner, 1), 1)
    tail = torch.zeros(sizes[1:], dtype=torch.int64, device=device)
    for d in range(1, len(shape)):
        view = [1] * (len(shape) - 1)
        view[d - 1] = sizes[d]
        tail = tail + ((torch.arange(sizes[d], device=device, dtype=torch.int64) + offsets[d]) * strides[d]).view(view)
    for r0 in range(0, sizes[0], step):
        r1 = min(r0 + step, sizes[0])
        rows = (torch.arange(r0, r1, device=device, dtype=torch.int64) + offsets[0]) * strides[0]
        yield slice(r0, r1), rows.view([r1 - r0] + [1] * (len(shape) - 1)) + tail


def fill_box(t: torch.Tensor, global_shape: Sequence[int], offsets: Sequence[int], sizes: Sequence[int], seed: int) -> None:
    v = t.view([1]) if t.ndim == 0 else t
    for rows, idx in _box_rows(global_shape, offsets, sizes, t.device):
        v[rows].copy_(pattern(idx, seed, t.dtype))


def box_equals(t: torch.Tensor, global_shape: Sequence[int], offsets: Sequence[int], sizes: Sequence[int], seed: int) -> bool:
    v = t.view([1]) if t.ndim == 0 else t
    it = _as_int(t.dtype)
    for rows, idx in _box_rows(global_shape, offsets, sizes, t.device):
        if not torch.equal(v[rows].contiguous().view(it), pattern(idx, seed, t.dtype).view(it)):
            return False
    return True


def make_sharded(global_shape: Sequence[int], dtype: torch.dtype, dim: int, pg: Optional[Any], device: str,
                 seed: Optional[int]) -> Any:
    """A ShardedTensor cut along `dim` over the ranks of `pg` (None: the world); filled with the pattern of
    object `seed`, or with zeros when seed is None."""
    from torch.distributed._shard import sharded_tensor
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec

    n = dist.get_world_size(group=pg)
    if device == "cpu":
        placements = [f"rank:{r}/cpu" for r in range(n)]
    else:  # one process per GPU on one box: the device index is the rank's index in the world
        placements = [f"rank:{r}/cuda:{dist.get_global_rank(pg, r) if pg is not None else r}" for r in range(n)]
    st = sharded_tensor.empty(ChunkShardingSpec(dim=dim, placements=placements), *global_shape, dtype=dtype,
                              process_group=pg)
    for shard in st.local_shards():
        if seed is None:
            shard.tensor.zero_()
        else:
            fill_box(shard.tensor, global_shape, shard.metadata.shard_offsets, shard.metadata.shard_sizes, seed)
    return st


def sharded_equals(st: Any, global_shape: Sequence[int], seed: int) -> bool:
    return all(
        box_equals(s.tensor, global_shape, s.metadata.shard_offsets, s.metadata.shard_sizes, seed)
        for s in st.local_shards()
    )


# name -> (global shape, dtype, sharding dim)
Spec = Dict[str, Tuple[List[int], torch.dtype, int]]


def small_spec() -> Spec:
    """Miniature of configs 3 and 4: flat parameters of awkward lengths in two dtypes (new shard boundaries fall
    inside saved pieces) and a row-wise table."""
    return {
        "flat_bf16": ([1_000_003], torch.bfloat16, 0),
        "flat_fp32": ([700_001], torch.float32, 0),
        "table": ([10_007, 128], torch.float32, 0),
    }


def build(spec: Spec, pg: Optional[Any], device: str, filled: bool, rep_shape: Sequence[int] = (333, 7)) -> Dict[str, Any]:
    state: Dict[str, Any] = {}
    for i, (name, (shape, dtype, dim)) in enumerate(spec.items()):
        state[name] = make_sharded(shape, dtype, dim, pg, device, seed=i if filled else None)
    rep = torch.zeros(list(rep_shape), dtype=torch.float32, device=device)
    if filled:
        fill_box(rep, list(rep_shape), [0] * len(rep_shape), list(rep_shape), 1000)
    state["rep"] = rep
    state["step"] = 5 if filled else 0
    return state


def verify(state: Dict[str, Any], spec: Spec, rep_shape: Sequence[int] = (333, 7)) -> List[str]:
    """Names of the objects whose restored bytes differ from the source pattern (empty = all good)."""
    bad = [name for i, (name, (shape, _dt, _dim)) in enumerate(spec.items()) if not sharded_equals(state[name], shape, i)]
    if not box_equals(state["rep"], list(rep_shape), [0] * len(rep_shape), list(rep_shape), 1000):
        bad.append("rep")
    if state["step"] != 5:
        bad.append("step")
    return bad


def elastic_roundtrip(tmp: str, device: str, spec: Optional[Spec] = None, max_shard_bytes: Optional[int] = 100_000,
                      sub: Optional[Any] = None) -> Dict[str, Any]:
    """Collective over the world: take at N -> restore at N/2 -> take at N/2 -> restore at N, every restored
    shard checked against the pattern.  `sub` is the process group of ranks [0, N/2) (created here when not
    given; creating a group is collective over the world).  Raises AssertionError on a mismatch; returns what
    path each restore took (for the caller's record)."""
    import tsb200_loader

    tsb200_loader.load()
    from torchsnapshot_b200 import knobs, reshard
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    rank, world = dist.get_rank(), dist.get_world_size()
    half = max(world // 2, 1)
    if sub is None:
        sub = dist.new_group(ranks=list(range(half)))
    spec = spec or small_spec()
    info: Dict[str, Any] = {"world": world, "half": half}

    def barrier() -> None:
        if device == "cpu":
            dist.barrier()
        else:
            torch.cuda.synchronize()
            dist.barrier(device_ids=[torch.cuda.current_device()])

    def override():
        import contextlib

        return knobs.override_max_shard_size_bytes(max_shard_bytes) if max_shard_bytes else contextlib.nullcontext()

    def note(tag: str) -> None:
        st = reshard.last_stats
        info[tag] = None if st is None else {
            "p2p": not st.fell_back and "sharded" not in st.skipped_groups, "skipped": list(st.skipped_groups),
            "peer_bytes": st.peer_bytes, "read_bytes": st.read_bytes}

    path_a, path_b = os.path.join(tmp, "taken_at_n"), os.path.join(tmp, "taken_at_half")
    src = StateDict(build(spec, None, device, filled=True))
    with override():
        Snapshot.take(path_a, {"m": src}, replicated=["m/rep", "m/step"])
    barrier()
    if rank < half:
        mid = StateDict(build(spec, sub, device, filled=False))
        reshard.last_stats = None
        Snapshot(path_a, pg=sub).restore({"m": mid})
        note("restore_at_half")
        bad = verify(mid, spec)
        assert not bad, f"rank {rank}: restored at {half} ranks, differs from the source: {bad}"
        with override():
            Snapshot.take(path_b, {"m": mid}, pg=sub, replicated=["m/rep", "m/step"])
    barrier()
    dst = StateDict(build(spec, None, device, filled=False))
    reshard.last_stats = None
    Snapshot(path_b).restore({"m": dst})
    note("restore_at_n")
    bad = verify(dst, spec)
    assert not bad, f"rank {rank}: restored at {world} ranks, differs from the source: {bad}"
    barrier()
    return info
