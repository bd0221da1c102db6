"""Pins the CPU oracle (oracle/) against the reference: fixtures written by the unmodified
reference (oracle/gen_golden.py) and the golden vectors of the reference's own tests."""

import hashlib
import json
import os

import pytest
import torch

import _cases
import oracle
from oracle import planner


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, name)) as f:
        return json.load(f)


@pytest.mark.parametrize("case", ["mixed", "threshold"])
def test_slab_layout_and_bytes(golden_dir, case):
    gold = _load(golden_dir, f"slab_{case}.json")
    threshold, items = getattr(_cases, f"slab_case_{case}")()
    assert threshold == gold["threshold"]
    tensors = {path: _cases.build_view(spec) for path, spec in items}
    plan_in = [(p, t.numel() * t.element_size(), False, True) for p, t in tensors.items()]
    slabs, passthrough = planner.plan_slabs(plan_in, threshold)
    assert len(slabs) == len(gold["slabs"])
    for k, (kind, members) in enumerate(slabs):
        size = members[-1][2]
        assert size == gold["slabs"][k]["size"]
        blob = oracle.pack_slab([(tensors[p], lo, hi) for p, lo, hi in members], size)
        assert sha(blob) == gold["slabs"][k]["sha256"]
        for p, lo, hi in members:
            g = gold["objects"][p]
            assert g["slab"] == k and g["byte_range"] == [lo, hi]
            assert sha(blob[lo:hi]) == g["sha256"]
    for p in passthrough:
        g = gold["objects"][p]
        assert g["slab"] is None and g["byte_range"] is None
        assert sha(oracle.serialize(tensors[p])) == g["sha256"]
    assert len(passthrough) + sum(len(m) for _k, m in slabs) == len(gold["objects"])


def test_chunking_against_reference_run(golden_dir):
    gold = {c["name"]: c for c in _load(golden_dir, "chunked.json")}
    for case in _cases.chunk_cases():
        t = _cases.build_view(case["spec"])
        chunks = planner.chunk_tensor(list(t.shape), t.element_size(), case["chunk_sz_bytes"])
        g = gold[case["name"]]["chunks"]
        assert [[o, s] for o, s in chunks] == [[c["offsets"], c["sizes"]] for c in g]
        base = t.reshape(-1) if t.ndim == 0 else t
        for (off, sizes), c in zip(chunks, g):
            assert c["location"] == "0/foo_" + "_".join(str(x) for x in off)
            assert sha(oracle.serialize(_cases.box_slice(base, (off, sizes)))) == c["sha256"]


def test_chunking_golden_vectors_of_reference_tests():
    # tests/test_chunked_tensor_io_preparer.py:50-101 in the reference
    assert planner.chunk_tensor([], 8, 100) == [([0], [1])]
    assert planner.chunk_tensor([7, 10], 4, 30 * 4) == [([0, 0], [3, 10]), ([3, 0], [3, 10]), ([6, 0], [1, 10])]
    assert planner.chunk_tensor([7, 10], 4, 180) == [([0, 0], [4, 10]), ([4, 0], [3, 10])]
    assert planner.chunk_tensor([10, 10], 4, 21 * 4) == [([2 * i, 0], [2, 10]) for i in range(5)]
    assert planner.chunk_tensor([6, 2, 4], 4, 24 * 4) == [([0, 0, 0], [3, 2, 4]), ([3, 0, 0], [3, 2, 4])]


def test_subdivision_golden_counts_of_reference_tests():
    # tests/test_sharded_tensor_io_preparer.py:210-295 in the reference: 256/256/86/86/1 pieces
    for max_bytes, dim, offsets, want in [(77, 0, [512, 0], 256), (1999, 0, [512, 0], 256), (4001, 0, [512, 0], 86),
                                          (4001, 1, [0, 512], 86), (300000, 0, [512, 0], 1)]:
        pieces = planner.subdivide_shard(offsets, [256, 256], 4, dim, max_bytes)
        assert len(pieces) == want
        assert pieces[0][0] == offsets
        assert sum(p[1][dim] for p in pieces) == 256


def test_sharded_write_and_reshard_plans(golden_dir):
    gold = _load(golden_dir, "sharded.json")
    shape = _cases.SHARD_GLOBAL_SHAPE
    full = _cases.raw_tensor(_cases.SHARD_SEED, shape, _cases.DTYPES[_cases.SHARD_DTYPE])
    layouts = _cases.shard_layouts(shape)
    pieces_by_layout = {}
    for name, (dim, boxes) in layouts.items():
        g = gold["writes"][name]
        assert g["boxes"] == [[o, s] for o, s in boxes]
        pieces = []
        for off, sizes in boxes:
            pieces += planner.subdivide_shard(off, sizes, 4, dim, _cases.SHARD_MAX_BYTES)
        assert [[o, s] for o, s in pieces] == [[p["offsets"], p["sizes"]] for p in g["pieces"]]
        for (off, sizes), p in zip(pieces, g["pieces"]):
            assert p["location"] == "sharded/foo_" + "_".join(str(x) for x in off)
            assert sha(oracle.serialize(_cases.box_slice(full, (off, sizes)))) == p["sha256"]
        pieces_by_layout[name] = pieces
        assert planner.global_shape(pieces) == shape
    for key, g in gold["reshard"].items():
        src_name, dst_name = key.split("->")
        dst_boxes = layouts[dst_name][1]
        dst_shards = [torch.zeros(s, dtype=full.dtype) for _o, s in dst_boxes]
        plan = []
        for off, sizes in pieces_by_layout[src_name]:
            regions = []
            for k, (d_off, d_sizes) in enumerate(dst_boxes):
                if not planner.boxes_overlap(d_off, d_sizes, off, sizes):
                    continue
                ov = planner.overlap_region(off, sizes, d_off, d_sizes)
                regions.append({"dst_shard": k, "overlap": [list(x) for x in ov]})
                payload = oracle.serialize(_cases.box_slice(full, (off, sizes)))
                oracle.copy_box(dst_shards[k], payload, sizes, ov)
            if regions:
                plan.append({"location": "sharded/foo_" + "_".join(str(x) for x in off), "regions": regions})
        assert plan == g["plan"], key
        assert [sha(oracle.serialize(s)) for s in dst_shards] == g["dst_sha256"], key


def test_tiled_reads(golden_dir):
    for g in _load(golden_dir, "tiled.json"):
        es = torch.empty((), dtype=_cases.DTYPES[g["dtype"]]).element_size()
        n = planner.numel(g["shape"])
        base = g["base"][0] if g["base"] else None
        assert [list(r) for r in planner.tile_byte_ranges(n, es, g["limit"], base)] == g["ranges"]
    # tests/test_tensor_io_preparer.py:148 in the reference: budget = nbytes / 4 -> exactly 4 reads
    assert len(planner.tile_byte_ranges(1009 * 1009, 4, 1009 * 1009, None)) == 4


def test_partition(golden_dir):
    for g in _load(golden_dir, "partition.json"):
        got = planner.partition_write_loads(list(zip(g["paths"], g["sizes"])), g["rank_to_size"])
        assert got == g["assignment"]


def test_merge_reads():
    reads = [("a", (10, 20)), ("b", None), ("a", (40, 50)), ("c", (0, 8)), ("a", (20, 30))]
    passthrough, merged = planner.merge_reads(reads)
    assert passthrough == [1]
    assert merged["a"] == ((10, 50), [(0, (0, 10)), (2, (30, 40)), (4, (10, 20))])
    assert merged["c"] == ((0, 8), [(3, (0, 8))])


def test_c1_snapshot_payloads(golden_dir):
    """BASELINE config 0: the oracle reproduces the bytes of the snapshot the reference took."""
    root = os.path.join(golden_dir, "c1_simple")
    inputs = torch.load(os.path.join(root, "inputs.pt"), weights_only=False)
    meta = json.load(open(os.path.join(root, "snapshot", ".snapshot_metadata")))
    tensors = inputs["tensors"]
    plan_in = [(p, tensors[p].numel() * tensors[p].element_size(), False, True) for p in inputs["order"] if p in tensors]
    slabs, passthrough = planner.plan_slabs(plan_in, 128 << 20)
    assert passthrough == [] and len(slabs) == 1
    members = slabs[0][1]
    blob = oracle.pack_slab([(tensors[p], lo, hi) for p, lo, hi in members], members[-1][2])
    locations = set()
    for p, lo, hi in members:
        e = meta["manifest"]["0/" + p]
        assert e["type"] == "Tensor" and e["byte_range"] == [lo, hi]
        locations.add(e["location"])
    assert len(locations) == 1
    with open(os.path.join(root, "snapshot", locations.pop()), "rb") as f:
        assert f.read() == blob


def test_cpu_port_of_take_and_restore(golden_dir, tmp_path):
    """oracle/ref_path.py (the timed CPU baseline) writes the bytes the reference wrote and reads them back."""
    from oracle import ref_path

    root = os.path.join(golden_dir, "c1_simple")
    inputs = torch.load(os.path.join(root, "inputs.pt"), weights_only=False)
    tensors = {p: inputs["tensors"][p] for p in inputs["order"] if p in inputs["tensors"]}
    manifest = ref_path.take(str(tmp_path / "s"), tensors)
    ref_meta = json.load(open(os.path.join(root, "snapshot", ".snapshot_metadata")))["manifest"]
    for path, e in manifest.items():
        r = ref_meta[path]
        assert e["byte_range"] == r["byte_range"] and e["dtype"] == r["dtype"] and e["shape"] == r["shape"]
        with open(os.path.join(root, "snapshot", r["location"]), "rb") as f:
            want = f.read()[r["byte_range"][0] : r["byte_range"][1]]
        with open(os.path.join(str(tmp_path / "s"), e["location"]), "rb") as f:
            assert f.read()[e["byte_range"][0] : e["byte_range"][1]] == want
    outs = {p: torch.zeros_like(t) for p, t in tensors.items()}
    ref_path.restore(os.path.join(root, "snapshot"), outs)  # reads the reference-written snapshot
    for p, t in tensors.items():
        assert oracle.serialize(outs[p]) == oracle.serialize(t)
    big = {"a": _cases.raw_tensor(1, [300, 7], torch.bfloat16), "b": _cases.raw_tensor(2, [5000], torch.float32),
           "c": _cases.raw_tensor(3, [64, 64], torch.float64)[:, ::2]}
    ref_path.take(str(tmp_path / "t"), big, threshold=4096)
    outs = {p: torch.zeros_like(t.contiguous()) for p, t in big.items()}
    ref_path.restore(str(tmp_path / "t"), outs)
    for p, t in big.items():
        assert oracle.serialize(outs[p]) == oracle.serialize(t)
