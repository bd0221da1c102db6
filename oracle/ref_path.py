"""TEST INFRASTRUCTURE -- CPU port of the reference's take / restore path for CPU tensors.

Restates, step for step and with the reference's own concurrency, what `Snapshot.take` and
`Snapshot.restore` do to tensor bytes when the state dict lives on CPU tensors and the sink is the
FS plugin.  It is the timed CPU baseline of bench.py (`cpu_baseline`, kind "port") and of
`bench.py --impl reference`; it is also compared against the reference-made fixtures in
tests/test_oracle_golden.py.  The product never imports it.

  take     per tensor: entry + stager objects            io_preparers/tensor.py:49-87
           slab planning                                 batcher.py:201-352 (oracle.planner.plan_slabs)
           per slab, ON THE EVENT-LOOP THREAD: slab = bytearray(size); slab[lo:hi] = memoryview of
           the tensor                                    batcher.py:63-90, serialization.py:162-188
           non-contiguous tensors: .contiguous() first   serialization.py:182-185
           <= 16 concurrent whole-file writes in worker threads   scheduler.py:30, 277-289; fs.py:26-36
           manifest json.dumps + metadata write          manifest.py:283-289, snapshot.py:731-741
  restore  metadata load + per-rank manifest             manifest.py:291-314, manifest_ops.py:24-52
           one ranged read per slab, <= 16 in flight: f.read -> BytesIO -> getvalue()
                                                         batcher.py:436-448, fs.py:38-49, scheduler.py:367-372
           per member: buf[lo:hi] (bytes copy) -> frombuffer -> dst.copy_(src) in a 4-thread pool
                                                         batcher.py:364-375, tensor.py:317-338, scheduler.py:29, 394
"""

import io
import json
import os
import uuid
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, List, Tuple

import torch

from . import planner

IO_THREADS = 16  # scheduler.py:30 _MAX_PER_RANK_IO_CONCURRENCY
CPU_THREADS = 4  # scheduler.py:29 _MAX_PER_RANK_CPU_CONCURRENCY
THREADS_USED = 1 + CPU_THREADS + IO_THREADS


def _as_memoryview(t: torch.Tensor) -> memoryview:
    # serialization.py:162-212 (bf16 has no numpy dtype: reinterpret the same bytes)
    if not t.is_contiguous():
        t = t.contiguous()
    if t.dtype == torch.bfloat16:
        t = t.view(torch.int16)
    if t.ndim == 0:
        t = t.reshape(1)
    return memoryview(t.numpy()).cast("b")


def _write_file(root: str, rel: str, buf) -> None:
    path = os.path.join(root, rel)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "wb+") as f:  # fs.py:34-36
        f.write(buf)


def take(root: str, tensors: Dict[str, torch.Tensor], threshold: int = 128 << 20) -> Dict[str, dict]:
    """Write `tensors` ({logical path: CPU tensor}) the way the reference does; returns the manifest."""
    manifest: Dict[str, dict] = {}
    plan_in = []
    for path, t in tensors.items():
        nbytes = t.numel() * t.element_size()
        manifest[f"0/{path}"] = {
            "type": "Tensor", "location": f"0/{path}", "serializer": "buffer_protocol", "dtype": str(t.dtype),
            "shape": list(t.shape), "replicated": False, "byte_range": None,
        }
        plan_in.append((path, nbytes, False, True))
    slabs, passthrough = planner.plan_slabs(plan_in, threshold)
    pool = ThreadPoolExecutor(max_workers=IO_THREADS)
    futures = []
    for path in passthrough:
        futures.append(pool.submit(_write_file, root, f"0/{path}", _as_memoryview(tensors[path])))
    for _kind, members in slabs:
        location = os.path.join("batched", str(uuid.uuid4()))
        slab = bytearray(members[-1][2])
        for path, lo, hi in members:
            slab[lo:hi] = _as_memoryview(tensors[path])
            e = manifest[f"0/{path}"]
            e["location"] = location
            e["byte_range"] = [lo, hi]
        futures.append(pool.submit(_write_file, root, location, memoryview(slab)))
    for f in futures:
        f.result()
    pool.shutdown()
    doc = json.dumps({"version": "0.1.0", "world_size": 1, "manifest": manifest}, sort_keys=False, indent=2)
    _write_file(root, ".snapshot_metadata", doc.encode("utf-8"))
    return manifest


def _read_range(root: str, rel: str, lo: int, hi: int) -> bytes:
    with open(os.path.join(root, rel), "rb") as f:  # fs.py:42-49
        f.seek(lo)
        return io.BytesIO(f.read(hi - lo)).getvalue()  # scheduler.py:367-372


def restore(root: str, outputs: Dict[str, torch.Tensor]) -> None:
    """Load the snapshot at `root` into `outputs` (same dtypes/shapes) the way the reference does."""
    with open(os.path.join(root, ".snapshot_metadata")) as f:
        manifest = json.loads(f.read())["manifest"]
    by_file: Dict[str, List[Tuple[str, int, int]]] = {}
    whole: List[str] = []
    for path in outputs:
        e = manifest[f"0/{path}"]
        if e["byte_range"] is None:
            whole.append(path)
        else:
            by_file.setdefault(e["location"], []).append((path, e["byte_range"][0], e["byte_range"][1]))
    io_pool = ThreadPoolExecutor(max_workers=IO_THREADS)
    cpu_pool = ThreadPoolExecutor(max_workers=CPU_THREADS)

    def load_member(path: str, buf: bytes) -> None:
        dst = outputs[path]
        dtype = torch.int16 if dst.dtype == torch.bfloat16 else dst.dtype
        n = dst.numel()
        src = torch.frombuffer(bytearray(buf), dtype=dtype) if n else torch.empty(0, dtype=dtype)
        if dst.dtype == torch.bfloat16:
            src = src.view(torch.bfloat16)
        dst.detach().copy_(src.reshape(dst.shape))  # tensor.py:356-358

    reads = []
    for location, members in by_file.items():
        lo = min(m[1] for m in members)
        hi = max(m[2] for m in members)
        reads.append((io_pool.submit(_read_range, root, location, lo, hi), lo, members))
    for path in whole:
        e = manifest[f"0/{path}"]
        size = os.path.getsize(os.path.join(root, e["location"]))
        reads.append((io_pool.submit(_read_range, root, e["location"], 0, size), 0, [(path, 0, size)]))
    copies = []
    for fut, base, members in reads:
        buf = fut.result()
        for path, lo, hi in members:
            copies.append(cpu_pool.submit(load_member, path, buf[lo - base : hi - base]))  # batcher.py:368-371
    for c in copies:
        c.result()
    io_pool.shutdown()
    cpu_pool.shutdown()


def take_gpu(root: str, tensors: Dict[str, torch.Tensor], threshold: int = 128 << 20) -> Dict[str, dict]:
    """The reference's own GPU take path, restated with the torch calls it makes (baseline only; this is
    what the CUDA kernels replace):
      per GPU slab, on the event-loop thread: gpu_buf = torch.cuda.ByteTensor(slab)      batcher.py:144
        one D2D copy_ per member tensor (after .contiguous())                            batcher.py:151-155
        gpu_buf.cpu(): blocking D2H into pageable memory                                 batcher.py:156
      tensors >= threshold: tensor.to("cpu") in the 4-thread pool                        tensor.py:247-254
      <= 16 concurrent whole-file writes in worker threads                               scheduler.py:277-289
    """
    manifest: Dict[str, dict] = {}
    plan_in = []
    for path, t in tensors.items():
        manifest[f"0/{path}"] = {
            "type": "Tensor", "location": f"0/{path}", "serializer": "buffer_protocol", "dtype": str(t.dtype),
            "shape": list(t.shape), "replicated": False, "byte_range": None,
        }
        plan_in.append((path, t.numel() * t.element_size(), True, True))
    slabs, passthrough = planner.plan_slabs(plan_in, threshold)
    io_pool = ThreadPoolExecutor(max_workers=IO_THREADS)
    cpu_pool = ThreadPoolExecutor(max_workers=CPU_THREADS)
    futures = []
    big = {path: cpu_pool.submit(lambda t: t.detach().to("cpu"), tensors[path]) for path in passthrough}
    for _kind, members in slabs:
        location = os.path.join("batched", str(uuid.uuid4()))
        gpu_buf = torch.empty(members[-1][2], dtype=torch.uint8, device=next(iter(tensors.values())).device)
        for path, lo, hi in members:
            t = tensors[path].detach().contiguous()
            gpu_buf[lo:hi].copy_(t.reshape(-1).view(torch.uint8))
            e = manifest[f"0/{path}"]
            e["location"] = location
            e["byte_range"] = [lo, hi]
        host = gpu_buf.cpu()
        futures.append(io_pool.submit(_write_file, root, location, memoryview(host.numpy())))
    for path, fut in big.items():
        futures.append(io_pool.submit(_write_file, root, f"0/{path}", _as_memoryview(fut.result())))
    for f in futures:
        f.result()
    io_pool.shutdown()
    cpu_pool.shutdown()
    doc = json.dumps({"version": "0.1.0", "world_size": 1, "manifest": manifest}, sort_keys=False, indent=2)
    _write_file(root, ".snapshot_metadata", doc.encode("utf-8"))
    return manifest
