"""Host interpreter of TsbRegion tuples: executes a region list with memmove on HOST addresses.
Test-side only -- it lets the CPU suite check the descriptor math of regions.py (strides, run
collapsing, dimension splitting) without a GPU.  The CUDA kernel is checked separately, on the GPU."""

import ctypes
import itertools
from typing import Sequence


def run_regions(regions: Sequence[tuple]) -> int:
    moved = 0
    for src, dst, run_bytes, sizes, sstr, dstr in regions:
        for idx in itertools.product(*[range(n) for n in sizes]):
            so = sum(i * s for i, s in zip(idx, sstr))
            do = sum(i * s for i, s in zip(idx, dstr))
            ctypes.memmove(dst + do, src + so, run_bytes)
            moved += run_bytes
    return moved
