"""TEST INFRASTRUCTURE -- CPU restatement of the reference's planning arithmetic for the hot path.

Pure Python integer math, no torch, each function citing the reference lines it follows.  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; it is the checker, never the thing measured or shipped.

Pinned by tests/test_oracle_golden.py against (a) the golden vectors the reference's own tests hold
(chunking offsets/sizes, sub-shard counts, tiled-read count) and (b) fixtures produced by running
the unmodified reference in the build container (oracle/gen_golden.py -> tests/golden/*.json).
"""

import math
from typing import Dict, List, Optional, Sequence, Tuple


def numel(shape: Sequence[int]) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


# --------------------------------------------------------------------------------------------
# slab planning -- torchsnapshot/batcher.py:201-352
# --------------------------------------------------------------------------------------------
def plan_slabs(
    items: Sequence[Tuple[str, int, bool, bool]], threshold: int
) -> Tuple[List[Tuple[str, List[Tuple[str, int, int]]]], List[str]]:
    """items: (path, nbytes, is_cuda, batchable) in write-request order.

    Returns (slabs, passthrough): slabs as (kind, [(path, lo, hi), ...]) in emission order -- CPU
    slabs first, then GPU slabs (batcher.py:319) -- and the paths that are written on their own.
      batcher.py:281-286  non-batchable stagers pass through
      batcher.py:291-293  tensors with nbytes >= threshold pass through
      batcher.py:300-302  a new slab starts when slab_size + nbytes >= threshold
      batcher.py:304-316  byte range = (slab_size, slab_size + nbytes), no padding
      batcher.py:320-321  empty slabs are dropped
    """
    cpu: List[List[Tuple[str, int, int]]] = [[]]
    gpu: List[List[Tuple[str, int, int]]] = [[]]
    cpu_sz = [0]
    gpu_sz = [0]
    passthrough: List[str] = []
    for path, nbytes, is_cuda, batchable in items:
        if not batchable or nbytes >= threshold:
            passthrough.append(path)
            continue
        slabs, sizes = (gpu, gpu_sz) if is_cuda else (cpu, cpu_sz)
        if sizes[-1] + nbytes >= threshold:
            slabs.append([])
            sizes.append(0)
        slabs[-1].append((path, sizes[-1], sizes[-1] + nbytes))
        sizes[-1] += nbytes
    out = [("cpu", s) for s in cpu if s] + [("gpu", s) for s in gpu if s]
    return out, passthrough


# --------------------------------------------------------------------------------------------
# read batching -- torchsnapshot/batcher.py:384-474
# --------------------------------------------------------------------------------------------
def merge_reads(
    reads: Sequence[Tuple[str, Optional[Tuple[int, int]]]]
) -> Tuple[List[int], Dict[str, Tuple[Tuple[int, int], List[Tuple[int, Tuple[int, int]]]]]]:
    """reads: (path, byte_range or None) by index.  Returns (passthrough indices, {path: (merged
    range, [(index, range relative to the merged lower bound)])}).
      batcher.py:433-435  byte_range None passes through
      batcher.py:443-448  merged range = (min lower, max upper)
      batcher.py:459-464  member range is shifted by the merged lower bound
    """
    passthrough: List[int] = []
    merged: Dict[str, Tuple[int, int]] = {}
    members: Dict[str, List[Tuple[int, Tuple[int, int]]]] = {}
    for i, (path, br) in enumerate(reads):
        if br is None:
            passthrough.append(i)
            continue
        lo, hi = merged.get(path, br)
        merged[path] = (min(lo, br[0]), max(hi, br[1]))
        members.setdefault(path, []).append((i, br))
    out = {}
    for path, (lo, hi) in merged.items():
        out[path] = ((lo, hi), [(i, (br[0] - lo, br[1] - lo)) for i, br in members[path]])
    return passthrough, out


# --------------------------------------------------------------------------------------------
# chunking -- torchsnapshot/io_preparers/chunked_tensor.py:35-62 (torch.chunk semantics)
# --------------------------------------------------------------------------------------------
def chunk_tensor(
    shape: Sequence[int], elsize: int, chunk_sz_bytes: int, dim: int = 0
) -> List[Tuple[List[int], List[int]]]:
    """(offsets, sizes) of each chunk.
      chunked_tensor.py:44-45  0-d tensors are viewed as 1-d
      chunked_tensor.py:47-49  n_chunks = ceil(nbytes / chunk_sz); torch.chunk(tensor, n_chunks, dim)
    torch.chunk: every chunk holds ceil(extent / n_chunks) slices, the last one the remainder, and
    fewer than n_chunks chunks come back when that uses the extent up.
    """
    shape = list(shape) if len(shape) else [1]
    nbytes = numel(shape) * elsize
    n_chunks = math.ceil(nbytes / chunk_sz_bytes)
    per = math.ceil(shape[dim] / n_chunks)
    out = []
    start = 0
    while start < shape[dim]:
        length = min(per, shape[dim] - start)
        offsets = [0] * len(shape)
        offsets[dim] = start
        sizes = list(shape)
        sizes[dim] = length
        out.append((offsets, sizes))
        start += length
    return out


# --------------------------------------------------------------------------------------------
# shard subdivision -- torchsnapshot/io_preparers/sharded_tensor.py:47-76
# --------------------------------------------------------------------------------------------
def subdivide_shard(
    offsets: Sequence[int], sizes: Sequence[int], elsize: int, dim: int, max_shard_sz_bytes: int
) -> List[Tuple[List[int], List[int]]]:
    """
      sharded_tensor.py:61     slice_sz = prod(sizes) // sizes[dim] * element_size
      sharded_tensor.py:62     chunk_length = max(floor(max_shard_sz / slice_sz), 1)
      sharded_tensor.py:63     n_chunks = ceil(sizes[dim] / chunk_length)
      sharded_tensor.py:66-75  piece i: start = i * chunk_length, length = min((i+1)*cl, size) - start
    """
    slice_sz = numel(sizes) // sizes[dim] * elsize
    chunk_length = max(math.floor(max_shard_sz_bytes / slice_sz), 1)
    n_chunks = math.ceil(sizes[dim] / chunk_length)
    out = []
    for i in range(n_chunks):
        start = i * chunk_length
        length = min((i + 1) * chunk_length, sizes[dim]) - start
        sub_off = list(offsets)
        sub_off[dim] += start
        sub_sizes = list(sizes)
        sub_sizes[dim] = length
        out.append((sub_off, sub_sizes))
    return out


# --------------------------------------------------------------------------------------------
# reshard overlap -- torchsnapshot/io_preparers/sharded_tensor.py:79-125, 229-248
# --------------------------------------------------------------------------------------------
def boxes_overlap(a_off: Sequence[int], a_sizes: Sequence[int], b_off: Sequence[int], b_sizes: Sequence[int]) -> bool:
    """torch.distributed._shard.sharding_spec._internals._check_shard_metadata_pair_overlap, which
    sharded_tensor.py:236 calls: two boxes are disjoint iff they are disjoint along some dim."""
    for ao, asz, bo, bsz in zip(a_off, a_sizes, b_off, b_sizes):
        if ao >= bo + bsz or bo >= ao + asz:
            return False
    return True


def overlap_region(
    saved_off: Sequence[int], saved_sizes: Sequence[int], cur_off: Sequence[int], cur_sizes: Sequence[int]
) -> List[Tuple[int, int, int, int]]:
    """(dim, offset in saved shard, offset in current shard, length) per dim
      sharded_tensor.py:106-111  length = min(ends) - max(starts)
      sharded_tensor.py:113-118  the side that starts later has offset 0, the other the difference
    """
    out = []
    for dim, (so, co, ss, cs) in enumerate(zip(saved_off, cur_off, saved_sizes, cur_sizes)):
        min_end = min(so + ss, co + cs)
        length = min_end - max(co, so)
        if so > co:
            off_saved, off_cur = 0, so - co
        else:
            off_saved, off_cur = co - so, 0
        out.append((dim, off_saved, off_cur, length))
    return out


def global_shape(shards: Sequence[Tuple[Sequence[int], Sequence[int]]]) -> List[int]:
    """sharded_tensor.py:173-180: max over shards of offset + size, per dim."""
    shape = [0] * len(shards[0][1])
    for off, sizes in shards:
        for d in range(len(off)):
            shape[d] = max(shape[d], off[d] + sizes[d])
    return shape


# --------------------------------------------------------------------------------------------
# tiled reads -- torchsnapshot/io_preparers/tensor.py:127-179
# --------------------------------------------------------------------------------------------
def tile_byte_ranges(
    total_elems: int, elsize: int, limit_bytes: int, base: Optional[int]
) -> List[Tuple[int, int]]:
    """Byte ranges of the tiles of a flattened (view(-1)) tensor.
      tensor.py:134-136  num_chunks = ceil(nbytes / limit)
      tensor.py:144      torch.chunk(flat, num_chunks, dim=0)
      tensor.py:162-171  range = (base + offset, base + offset + chunk bytes), base 0 if no byte_range
    """
    nbytes = total_elems * elsize
    n = math.ceil(nbytes / limit_bytes)
    per = math.ceil(total_elems / n)
    out = []
    start = 0
    b = base or 0
    while start < total_elems:
        length = min(per, total_elems - start)
        out.append((b + start * elsize, b + (start + length) * elsize))
        start += length
    return out


# --------------------------------------------------------------------------------------------
# replicated-write partitioning -- torchsnapshot/partitioner.py:42-79
# --------------------------------------------------------------------------------------------
def partition_write_loads(
    loads: Sequence[Tuple[str, int]], rank_to_size: Sequence[int]
) -> List[List[str]]:
    """Greedy assignment of non-subpartitionable logical paths, in rank-0 dict order, to the rank
    with the least bytes so far (np.argmin: the lowest index wins ties).
      partitioner.py:62-69  chosen_rank = argmin(rank_to_size); rank_to_size[chosen] += size
    """
    sizes = list(rank_to_size)
    out: List[List[str]] = [[] for _ in sizes]
    for path, size in loads:
        chosen = min(range(len(sizes)), key=lambda r: (sizes[r], r))
        out[chosen].append(path)
        sizes[chosen] += size
    return out
