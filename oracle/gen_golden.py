#!/usr/bin/env python3
"""TEST INFRASTRUCTURE -- generates tests/golden/ by running the UNMODIFIED reference.

    PYTHONPATH=/root/reference python oracle/gen_golden.py

Runs only in the build container (the reference is mounted at /root/reference and cannot travel to
the GPU box); what it writes is committed:

  tests/golden/slab_<case>.json    reference batch_write_requests + staged slabs: per logical path
                                   (slab index, byte_range, sha256 of payload), per slab (size, sha256)
  tests/golden/chunked.json        reference chunk_tensor / ChunkedTensorIOPreparer.prepare_write
  tests/golden/sharded.json        reference ShardedTensorIOPreparer.prepare_write (with subdivision)
                                   and prepare_read overlap plans for every src x dst layout
  tests/golden/tiled.json          reference prepare_read_tiled byte ranges
  tests/golden/partition.json      reference _partition_write_loads (non-chunked, deterministic part)
  tests/golden/c1_simple/          a complete snapshot taken by the reference Snapshot.take of the
                                   simple_example.py-style MLP + Adam state (BASELINE config 0), plus
                                   the flattened input tensors (inputs.pt)
"""

import asyncio
import hashlib
import json
import os
import shutil
import sys
import uuid
from typing import Any, Dict, List

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference")

import _cases  # noqa: E402
import torchsnapshot  # noqa: E402  (the reference)
from torchsnapshot import knobs  # noqa: E402
from torchsnapshot.batcher import batch_write_requests  # noqa: E402
from torchsnapshot.io_preparer import (  # noqa: E402
    ChunkedTensorIOPreparer,
    ShardedTensorIOPreparer,
    TensorIOPreparer,
    prepare_write,
)
from torchsnapshot.manifest import SnapshotMetadata  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def run(coro):
    return asyncio.new_event_loop().run_until_complete(coro)


def gen_slab_case(name: str, threshold: int, items) -> None:
    entries = {}
    wrs = []
    for path, spec in items:
        t = _cases.build_view(spec)
        entry, w = prepare_write(obj=t, logical_path=path, rank=0, replicated=False)
        entries[path] = entry
        wrs += w
    _, batched = batch_write_requests(list(entries.values()), wrs, slab_size_threshold_bytes=threshold)
    staged = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in batched}
    slab_ids: Dict[str, int] = {}
    slabs = []
    for wr in batched:
        if wr.path.startswith("batched/"):
            slab_ids[wr.path] = len(slabs)
            slabs.append({"size": len(staged[wr.path]), "sha256": sha(staged[wr.path])})
    objects = {}
    for path, entry in entries.items():
        buf = staged[entry.location]
        br = entry.byte_range
        payload = buf if br is None else buf[br[0] : br[1]]
        objects[path] = {
            "slab": slab_ids.get(entry.location),
            "location": None if entry.location in slab_ids else entry.location,
            "byte_range": br,
            "dtype": entry.dtype,
            "shape": entry.shape,
            "serializer": entry.serializer,
            "nbytes": len(payload),
            "sha256": sha(payload),
        }
    with open(os.path.join(GOLDEN, f"slab_{name}.json"), "w") as f:
        json.dump({"threshold": threshold, "order": [p for p, _ in items], "slabs": slabs, "objects": objects}, f,
                  indent=1)


def gen_chunked() -> None:
    out = []
    for case in _cases.chunk_cases():
        t = _cases.build_view(case["spec"])
        chunks = ChunkedTensorIOPreparer.chunk_tensor(t, chunk_sz_bytes=case["chunk_sz_bytes"])
        entry, wrs = ChunkedTensorIOPreparer.prepare_write("0/foo", t, chunks)
        staged = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in wrs}
        out.append(
            {
                "name": case["name"],
                "dtype": entry.dtype,
                "shape": entry.shape,
                "chunks": [
                    {
                        "offsets": c.offsets,
                        "sizes": c.sizes,
                        "location": c.tensor.location,
                        "shape": c.tensor.shape,
                        "sha256": sha(staged[c.tensor.location]),
                    }
                    for c in entry.chunks
                ],
            }
        )
    with open(os.path.join(GOLDEN, "chunked.json"), "w") as f:
        json.dump(out, f, indent=1)


def make_sharded(boxes, global_shape, full=None):
    import torch.distributed as dist
    from torch.distributed._shard import sharded_tensor
    from torch.distributed._shard.metadata import ShardMetadata
    from torch.distributed._shard.sharded_tensor import Shard

    shards = []
    for off, sizes in boxes:
        if full is not None:
            local = _cases.box_slice(full, (off, sizes)).contiguous().clone()
        else:
            local = torch.zeros(sizes, dtype=_cases.DTYPES[_cases.SHARD_DTYPE])
        shards.append(Shard(tensor=local, metadata=ShardMetadata(shard_offsets=off, shard_sizes=sizes,
                                                                 placement="rank:0/cpu")))
    return sharded_tensor.init_from_local_shards(shards, global_shape)


def gen_sharded() -> None:
    import torch.distributed as dist
    from torch.distributed._shard.sharding_spec import ChunkShardingSpec

    dist.init_process_group(backend="gloo", init_method=f"file:///tmp/{uuid.uuid4()}", rank=0, world_size=1)
    shape = _cases.SHARD_GLOBAL_SHAPE
    full = _cases.raw_tensor(_cases.SHARD_SEED, shape, _cases.DTYPES[_cases.SHARD_DTYPE])
    layouts = _cases.shard_layouts(shape)
    out: Dict[str, Any] = {"writes": {}, "reshard": {}}
    entries = {}
    stageds = {}
    for name, (dim, boxes) in layouts.items():
        st = make_sharded(boxes, shape, full)
        # subdivision follows the ChunkShardingSpec dim; init_from_local_shards builds an
        # EnumerableShardingSpec (dim 0), so call subdivide_shard with the layout's own dim the way
        # prepare_write does for a ChunkShardingSpec (sharded_tensor.py:139-151)
        with knobs.override_max_shard_size_bytes(_cases.SHARD_MAX_BYTES):
            shards = []
            wrs = []
            for local in st.local_shards():
                for tensor, offsets, sizes in ShardedTensorIOPreparer.subdivide_shard(
                    shard=local.tensor, offsets=local.metadata.shard_offsets, sizes=local.metadata.shard_sizes,
                    dim=dim, max_shard_sz_bytes=knobs.get_max_shard_size_bytes(),
                ):
                    suffix = "_".join(str(i) for i in offsets)
                    entry, w = TensorIOPreparer.prepare_write(storage_path=f"sharded/foo_{suffix}", tensor=tensor)
                    wrs += w
                    from torchsnapshot.manifest import Shard as MShard

                    shards.append(MShard(offsets=offsets, sizes=sizes, tensor=entry))
            from torchsnapshot.manifest import ShardedTensorEntry

            entry = ShardedTensorEntry(shards=shards)
        staged = {wr.path: bytes(run(wr.buffer_stager.stage_buffer())) for wr in wrs}
        entries[name] = entry
        stageds[name] = staged
        out["writes"][name] = {
            "dim": dim,
            "boxes": [[o, s] for o, s in boxes],
            "pieces": [
                {"offsets": s.offsets, "sizes": s.sizes, "location": s.tensor.location,
                 "sha256": sha(staged[s.tensor.location])}
                for s in entry.shards
            ],
        }
    for src_name, dst_name in [(a, b) for a in layouts for b in layouts]:
        dst = make_sharded(layouts[dst_name][1], shape, None)
        read_reqs, _ = ShardedTensorIOPreparer.prepare_read(entry=entries[src_name], obj_out=dst)
        plan = []
        dst_ptrs = [s.tensor.data_ptr() for s in dst.local_shards()]
        for rr in read_reqs:
            regions = []
            for reg in rr.buffer_consumer.overlapping_regions:
                regions.append({"dst_shard": dst_ptrs.index(reg.dst_tensor.data_ptr()),
                                "overlap": [list(x) for x in reg.overlap_region]})
            plan.append({"location": rr.path, "regions": regions})
            run(rr.buffer_consumer.consume_buffer(buf=stageds[src_name][rr.path]))
        out["reshard"][f"{src_name}->{dst_name}"] = {
            "plan": plan,
            "dst_sha256": [sha(s.tensor.contiguous().numpy().tobytes()) for s in dst.local_shards()],
        }
        for s, box in zip(dst.local_shards(), layouts[dst_name][1]):
            assert torch.equal(s.tensor, _cases.box_slice(full, box))
    # sharded -> plain tensor (read_object path, sharded_tensor.py:211-221)
    dense = torch.zeros(shape, dtype=full.dtype)
    read_reqs, _ = ShardedTensorIOPreparer.prepare_read(entry=entries["dim1x3"], obj_out=dense)
    for rr in read_reqs:
        run(rr.buffer_consumer.consume_buffer(buf=stageds["dim1x3"][rr.path]))
    assert torch.equal(dense, full)
    out["dense_sha256"] = sha(dense.numpy().tobytes())
    with open(os.path.join(GOLDEN, "sharded.json"), "w") as f:
        json.dump(out, f, indent=1)
    dist.destroy_process_group()


def gen_tiled() -> None:
    from torchsnapshot.manifest import TensorEntry

    out = []
    for shape, dtype, limit, base in [([1009, 1009], "float32", 1009 * 1009, None), ([100, 7], "bfloat16", 333, [64, 64 + 1400]),
                                      ([5], "int64", 1000, [8, 48]), ([1000], "uint8", 77, None)]:
        t = torch.zeros(shape, dtype=_cases.DTYPES[dtype])
        entry = TensorEntry(location="x", serializer="buffer_protocol", dtype=str(t.dtype), shape=shape,
                            replicated=False, byte_range=base)
        rrs, _ = TensorIOPreparer.prepare_read(entry, t, buffer_size_limit_bytes=limit)
        out.append({"shape": shape, "dtype": dtype, "limit": limit, "base": base,
                    "ranges": [list(rr.byte_range) for rr in rrs],
                    "tile_shapes": [list(rr.buffer_consumer.tensor.shape) for rr in rrs]})
    with open(os.path.join(GOLDEN, "tiled.json"), "w") as f:
        json.dump(out, f, indent=1)


def gen_partition() -> None:
    import numpy as np
    from torchsnapshot.manifest import TensorEntry
    from torchsnapshot.partitioner import _partition_write_loads, _WriteLoad

    cases = []
    rs = np.random.RandomState(11)
    for world, n, seed_sizes in [(4, 23, [0, 0, 0, 0]), (8, 100, [0] * 8), (3, 10, [500, 0, 100])]:
        sizes = [int(rs.randint(1, 10000)) for _ in range(n)]
        paths = [f"p{i}" for i in range(n)]
        entries = [{p: TensorEntry(location=p, serializer="buffer_protocol", dtype="torch.float32", shape=[1],
                                   replicated=True) for p in paths} for _ in range(world)]
        loads = [{p: [_WriteLoad(logical_path=p, write_req_idx=0, size=s)] for p, s in zip(paths, sizes)}
                 for _ in range(world)]
        res = _partition_write_loads(entries, loads, list(seed_sizes), world)
        cases.append({"world": world, "paths": paths, "sizes": sizes, "rank_to_size": seed_sizes,
                      "assignment": [[wl.logical_path for wl in r] for r in res]})
    with open(os.path.join(GOLDEN, "partition.json"), "w") as f:
        json.dump(cases, f, indent=1)


def gen_c1() -> None:
    """BASELINE config 0: examples/simple_example.py-style MLP + Adam on CPU, reference Snapshot.take."""
    from torchsnapshot.flatten import flatten

    dst = os.path.join(GOLDEN, "c1_simple")
    shutil.rmtree(dst, ignore_errors=True)
    torch.manual_seed(1234)
    model = torch.nn.Sequential(torch.nn.Linear(128, 64), torch.nn.Linear(64, 32), torch.nn.Linear(32, 1))
    optim = torch.optim.Adam(model.parameters(), lr=1e-3)
    model(torch.randn(16, 128)).sum().backward()
    optim.step()
    app_state = {"model": model, "optim": optim, "progress": torchsnapshot.StateDict(current_epoch=0)}
    torchsnapshot.Snapshot.take(path=os.path.join(dst, "snapshot"), app_state=app_state)
    flat: Dict[str, Any] = {}
    for key, stateful in app_state.items():
        _, f = flatten(stateful.state_dict(), prefix=key)
        flat.update(f)
    tensors = {k: v.detach().clone() for k, v in flat.items() if isinstance(v, torch.Tensor)}
    others = {k: v for k, v in flat.items() if not isinstance(v, torch.Tensor)}
    torch.save({"tensors": tensors, "others": others, "order": list(flat.keys())}, os.path.join(dst, "inputs.pt"))
    with open(os.path.join(dst, "snapshot", ".snapshot_metadata")) as f:
        meta = SnapshotMetadata.from_yaml(f.read())
    assert meta.world_size == 1


def main() -> None:
    os.makedirs(GOLDEN, exist_ok=True)
    th, items = _cases.slab_case_mixed()
    gen_slab_case("mixed", th, items)
    th, items = _cases.slab_case_threshold()
    gen_slab_case("threshold", th, items)
    gen_chunked()
    gen_tiled()
    gen_partition()
    gen_sharded()
    gen_c1()
    print("golden fixtures written to", GOLDEN)


if __name__ == "__main__":
    main()
