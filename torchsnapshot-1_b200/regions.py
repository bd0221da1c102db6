"""Host half of the GPU-side planner: tensors -> TsbRegion descriptor tables.

A region is a strided box (include/tsb200.h, struct TsbRegion).  The three producers are
  pack_regions    tensor (any strides)  -> contiguous bytes at `dst_addr`; the byte order is the
                  one tensor_as_memoryview gives after .contiguous() (reference
                  torchsnapshot/serialization.py:162-188), i.e. row-major over the logical shape;
  unpack_regions  contiguous bytes at `src_addr` -> tensor (reference tensor.py:329-338);
  box_regions     a sub-box of one strided view -> the same-shaped sub-box of another (reference
                  sharded_tensor.py:271-284: narrow() on both sides, then copy_).
Everything here is integer arithmetic on shapes and strides; no tensor data is touched.
"""

from typing import List, Sequence, Tuple

import numpy as np
import torch

from ._native import REGION_DTYPE, TSB_MAX_DIMS

# (src, dst, run_bytes, sizes, src_strides, dst_strides) with byte strides, len(sizes) <= TSB_MAX_DIMS
RegionTuple = Tuple[int, int, int, Tuple[int, ...], Tuple[int, ...], Tuple[int, ...]]


def contiguous_strides(sizes: Sequence[int], elsize: int) -> List[int]:
    """Row-major byte strides of a dense tensor of `sizes`."""
    strides = [0] * len(sizes)
    acc = elsize
    for i in range(len(sizes) - 1, -1, -1):
        strides[i] = acc
        acc *= max(int(sizes[i]), 1)
    return strides


def box_regions(
    src_addr: int,
    src_strides: Sequence[int],
    dst_addr: int,
    dst_strides: Sequence[int],
    sizes: Sequence[int],
    elsize: int,
) -> List[RegionTuple]:
    """Regions that copy the box `sizes` from one strided view to another (byte strides)."""
    dims = []
    for n, ss, ds in zip(sizes, src_strides, dst_strides):
        n = int(n)
        if n == 0:
            return []
        if n != 1:
            dims.append([n, int(ss), int(ds)])
    # merge neighbours that are dense with respect to each other on both sides
    merged: List[List[int]] = []
    for d in dims:
        if merged and merged[-1][1] == d[1] * d[0] and merged[-1][2] == d[2] * d[0]:
            merged[-1] = [merged[-1][0] * d[0], d[1], d[2]]
        else:
            merged.append(d)
    run_bytes = elsize
    if merged and merged[-1][1] == elsize and merged[-1][2] == elsize:
        run_bytes = merged[-1][0] * elsize
        merged.pop()
    out: List[RegionTuple] = []

    def emit(src: int, dst: int, outer: List[List[int]]) -> None:
        if len(outer) <= TSB_MAX_DIMS:
            out.append(
                (src, dst, run_bytes, tuple(d[0] for d in outer), tuple(d[1] for d in outer), tuple(d[2] for d in outer))
            )
            return
        n, ss, ds = outer[0]
        for i in range(n):
            emit(src + i * ss, dst + i * ds, outer[1:])

    emit(int(src_addr), int(dst_addr), merged)
    return out


def _byte_strides(t: torch.Tensor) -> List[int]:
    es = t.element_size()
    return [s * es for s in t.stride()]


def pack_regions(t: torch.Tensor, dst_addr: int) -> List[RegionTuple]:
    """tensor -> its serialized bytes at dst_addr."""
    es = t.element_size()
    if t.is_contiguous():
        n = t.numel() * es
        return [(t.data_ptr(), dst_addr, n, (), (), ())] if n else []
    sizes = list(t.shape)
    return box_regions(t.data_ptr(), _byte_strides(t), dst_addr, contiguous_strides(sizes, es), sizes, es)


def unpack_regions(src_addr: int, t: torch.Tensor) -> List[RegionTuple]:
    """serialized bytes at src_addr -> tensor."""
    es = t.element_size()
    if t.is_contiguous():
        n = t.numel() * es
        return [(src_addr, t.data_ptr(), n, (), (), ())] if n else []
    sizes = list(t.shape)
    return box_regions(src_addr, contiguous_strides(sizes, es), t.data_ptr(), _byte_strides(t), sizes, es)


def regions_to_array(regions: Sequence[RegionTuple]) -> np.ndarray:
    """Pack region tuples into the C layout (numpy structured array of REGION_DTYPE)."""
    n = len(regions)
    arr = np.zeros(n, dtype=REGION_DTYPE)
    if n == 0:
        return arr
    arr["src"] = np.fromiter((r[0] for r in regions), dtype=np.uint64, count=n)
    arr["dst"] = np.fromiter((r[1] for r in regions), dtype=np.uint64, count=n)
    arr["run_bytes"] = np.fromiter((r[2] for r in regions), dtype=np.uint64, count=n)
    for i, r in enumerate(regions):
        nd = len(r[3])
        if nd:
            rec = arr[i]
            rec["ndim"] = nd
            rec["size"][:nd] = r[3]
            rec["src_stride"][:nd] = r[4]
            rec["dst_stride"][:nd] = r[5]
    return arr


def regions_payload_bytes(regions: Sequence[RegionTuple]) -> int:
    total = 0
    for r in regions:
        n = r[2]
        for s in r[3]:
            n *= s
        total += n
    return total
