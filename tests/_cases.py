"""Seeded synthetic inputs shared by the fixture generator (oracle/gen_golden.py, run against the
real reference in the build container) and the tests (run anywhere, including the GPU box where
the reference does not exist).  Values come from numpy's RandomState byte stream, so they are
identical on every machine and torch build."""

from typing import Any, Dict, List, Sequence, Tuple

import numpy as np
import torch

DTYPES = {
    "float64": torch.float64,
    "float32": torch.float32,
    "float16": torch.float16,
    "bfloat16": torch.bfloat16,
    "int64": torch.int64,
    "int32": torch.int32,
    "int16": torch.int16,
    "int8": torch.int8,
    "uint8": torch.uint8,
    "bool": torch.bool,
}


def raw_tensor(seed: int, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
    """Dense CPU tensor of `shape` filled with reproducible bytes (floats may hold any bit pattern
    except that NaN payloads are avoided so value comparisons stay meaningful)."""
    n = int(np.prod(shape)) if len(shape) else 1
    es = torch.empty((), dtype=dtype).element_size()
    rs = np.random.RandomState(seed)
    if dtype == torch.bool:
        return torch.from_numpy(rs.randint(0, 2, size=n).astype(np.uint8)).view(torch.bool).reshape(list(shape))
    if dtype.is_floating_point:
        vals = rs.standard_normal(n).astype(np.float64)
        return torch.from_numpy(vals).to(dtype).reshape(list(shape))
    raw = rs.randint(0, 256, size=n * es).astype(np.uint8)
    return torch.from_numpy(raw).view(dtype).reshape(list(shape))


def build_view(spec: Dict[str, Any], device: str = "cpu") -> torch.Tensor:
    """spec: {seed, shape, dtype, view}.  `shape` is the shape of the dense base tensor; `view`
    selects how the saved tensor looks at it."""
    base = raw_tensor(spec["seed"], spec["shape"], DTYPES[spec["dtype"]]).to(device)
    view = spec.get("view", "dense")
    if view == "dense":
        return base
    if view == "scalar":  # 0-d
        return base.reshape(-1)[0]
    if view == "transpose":
        return base.transpose(0, 1)
    if view == "stride2":  # every other column
        return base[:, ::2]
    if view == "offset":  # storage offset, still contiguous
        return base.reshape(-1)[3:]
    if view == "inner":  # a box strictly inside a matrix
        return base[1:-1, 2:-1]
    if view == "narrow1":  # the kind of view shard subdivision along dim 1 makes
        return base.narrow(1, 1, base.shape[1] - 2)
    if view == "permute3":
        return base.permute(2, 0, 1)
    if view == "empty":
        return base[:0]
    raise ValueError(view)


def _spec(seed: int, shape: Sequence[int], dtype: str, view: str = "dense") -> Dict[str, Any]:
    return {"seed": seed, "shape": list(shape), "dtype": dtype, "view": view}


def slab_case_mixed() -> Tuple[int, List[Tuple[str, Dict[str, Any]]]]:
    """Every buffer-protocol dtype, odd byte counts (so later members start at unaligned slab
    offsets), 0-d scalars, strided views, an empty tensor."""
    items: List[Tuple[str, Dict[str, Any]]] = []
    seed = 100
    for dt in DTYPES:
        # the reference drops the last 2 bytes of a bf16 tensor whose byte count is not a multiple
        # of 4, so bf16 members of the reference-made fixture have even numel
        bf16 = dt == "bfloat16"
        items.append((f"dense_{dt}", _spec(seed, [6, 7] if bf16 else [5, 7], dt)))
        if not bf16:
            items.append((f"scalar_{dt}", _spec(seed + 1, [4], dt, "scalar")))
        items.append((f"odd_{dt}", _spec(seed + 2, [14] if bf16 else [13], dt)))
        seed += 3
    items.append(("u8_1", _spec(900, [1], "uint8")))
    items.append(("f32_after_u8", _spec(901, [33, 9], "float32")))
    items.append(("u8_3", _spec(902, [3], "uint8")))
    items.append(("i64_after_u8", _spec(903, [17], "int64")))
    items.append(("bf16_even", _spec(904, [6, 10], "bfloat16")))
    items.append(("transpose_f32", _spec(905, [9, 14], "float32", "transpose")))
    items.append(("stride2_f16", _spec(906, [8, 10], "float16", "stride2")))
    items.append(("offset_f64", _spec(907, [40], "float64", "offset")))
    items.append(("inner_i32", _spec(908, [9, 11], "int32", "inner")))
    items.append(("narrow1_f32", _spec(909, [16, 12], "float32", "narrow1")))
    items.append(("permute3_i16", _spec(910, [3, 4, 5], "int16", "permute3")))
    # (no empty tensor here: the reference cannot stage one -- memoryview.cast rejects a zero in
    # the shape, serialization.py:188; tests/test_edge_cases.py covers it for this build)
    items.append(("big_f32", _spec(912, [300, 129], "float32")))
    items.append(("tail_u8", _spec(913, [5], "uint8")))
    return 128 * 1024 * 1024, items


def slab_case_threshold() -> Tuple[int, List[Tuple[str, Dict[str, Any]]]]:
    """A small threshold: several slabs, tensors at/over the threshold written on their own."""
    items = []
    rs = np.random.RandomState(7)
    for i in range(40):
        n = int(rs.randint(1, 700))
        dt = ["float32", "bfloat16", "uint8", "int64"][i % 4]
        if dt == "bfloat16":
            n += n % 2
        items.append((f"p{i}", _spec(2000 + i, [n], dt)))
    items.append(("exact_threshold", _spec(2100, [1024], "float32")))  # 4096 B == threshold
    items.append(("over_threshold", _spec(2101, [3000], "float32")))
    items.append(("just_under", _spec(2102, [1023], "float32")))
    return 4096, items


def chunk_cases() -> List[Dict[str, Any]]:
    return [
        {"name": "rows", "spec": _spec(3000, [7, 10], "float32"), "chunk_sz_bytes": 120},
        {"name": "nondiv", "spec": _spec(3001, [7, 10], "float32"), "chunk_sz_bytes": 180},
        {"name": "noncontig", "spec": _spec(3002, [10, 12], "float32", "narrow1"), "chunk_sz_bytes": 84},
        {"name": "bf16", "spec": _spec(3003, [64, 6], "bfloat16"), "chunk_sz_bytes": 100},
        {"name": "one", "spec": _spec(3004, [5, 5], "int64"), "chunk_sz_bytes": 4096},
        {"name": "scalar", "spec": _spec(3005, [1], "int64", "scalar"), "chunk_sz_bytes": 100},
    ]


Box = Tuple[List[int], List[int]]  # (offsets, sizes)


def chunk_boxes(shape: Sequence[int], dim: int, n: int) -> List[Box]:
    """ChunkShardingSpec(dim, n placements) boxes: ceil split, trailing empty shards dropped."""
    per = -(-shape[dim] // n)
    out = []
    for i in range(n):
        lo = i * per
        hi = min(lo + per, shape[dim])
        if hi <= lo:
            break
        off = [0] * len(shape)
        off[dim] = lo
        sizes = list(shape)
        sizes[dim] = hi - lo
        out.append((off, sizes))
    return out


def grid_boxes(shape: Sequence[int], rows: int, cols: int) -> List[Box]:
    out = []
    rh, cw = shape[0] // rows, shape[1] // cols
    for r in range(rows):
        for c in range(cols):
            out.append(([r * rh, c * cw], [rh, cw]))
    return out


def shard_layouts(shape: Sequence[int]) -> Dict[str, Tuple[int, List[Box]]]:
    """name -> (sharding dim the reference would subdivide along, boxes)."""
    return {
        "dim0x3": (0, chunk_boxes(shape, 0, 3)),
        "dim0x5": (0, chunk_boxes(shape, 0, 5)),
        "dim1x3": (1, chunk_boxes(shape, 1, 3)),
        "dim1x5": (1, chunk_boxes(shape, 1, 5)),
        "grid2x2": (0, grid_boxes(shape, 2, 2)),
    }


SHARD_GLOBAL_SHAPE = [24, 20]
SHARD_DTYPE = "float32"
SHARD_SEED = 4000
SHARD_MAX_BYTES = 2 * 20 * 4 + 10  # forces subdivision: two full rows and a bit


def box_slice(t: torch.Tensor, box: Box) -> torch.Tensor:
    off, sizes = box
    for d, (o, s) in enumerate(zip(off, sizes)):
        t = t.narrow(d, o, s)
    return t


def random_view(rs, base: torch.Tensor) -> torch.Tensor:
    """A random non-overlapping view of `base`: per-dim slices with steps, then a random permutation, then
    (sometimes) a dropped leading index -- the shapes torch hands a stager in practice, and then some."""
    idx = []
    for n in base.shape:
        lo = int(rs.randint(0, max(n // 3, 1)))
        hi = int(rs.randint(lo + 1, n + 1))
        step = int(rs.choice([1, 1, 1, 2, 3]))
        idx.append(slice(lo, hi, step))
    v = base[tuple(idx)]
    if v.ndim > 1 and rs.rand() < 0.6:
        v = v.permute(*[int(i) for i in rs.permutation(v.ndim)])
    if v.ndim > 1 and rs.rand() < 0.2:
        v = v[int(rs.randint(0, v.shape[0]))]
    return v


def random_base_shape(rs, rank: int, max_log2_elems: float) -> List[int]:
    """A random shape of `rank` dims with about 2^U(4, max_log2_elems) elements."""
    budget = int(2 ** rs.uniform(4, max_log2_elems))
    shape = []
    for d in range(rank):
        n = max(1, int(round(budget ** (1.0 / (rank - d)) * rs.uniform(0.5, 1.5)))) if budget > 1 else 1
        n = min(n, budget)
        shape.append(n)
        budget = max(budget // n, 1)
    return shape
