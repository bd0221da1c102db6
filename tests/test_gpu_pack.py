"""Parity of the CUDA multi-region copy kernel with the CPU oracle, through the C-ABI.

Bit-exact: the path moves bytes, so every comparison is on raw bytes (no tolerance)."""

import hashlib
import json
import os

import numpy as np
import pytest
import torch

import _cases
import oracle
from oracle import planner
from torchsnapshot_b200 import _native
from torchsnapshot_b200.regions import box_regions, pack_regions, regions_to_array, unpack_regions

pytestmark = pytest.mark.gpu


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def _launch(engine, regions):
    engine.wait_current_stream(_native.STREAM_PACK)
    engine.copy_regions(regions, _native.STREAM_PACK)
    engine.ctx.stream_sync(_native.STREAM_PACK)


def _to_host(t: torch.Tensor) -> bytes:
    return t.cpu().numpy().tobytes()


@pytest.mark.parametrize("case", ["mixed", "threshold"])
def test_pack_matches_oracle_and_reference_fixture(engine, golden_dir, case):
    gold = json.load(open(os.path.join(golden_dir, f"slab_{case}.json")))
    threshold, items = getattr(_cases, f"slab_case_{case}")()
    cpu = {p: _cases.build_view(spec) for p, spec in items}
    gpu = {p: _cases.build_view(spec, "cuda") for p, spec in items}
    plan_in = [(p, t.numel() * t.element_size(), True, True) for p, t in cpu.items()]
    slabs, _ = planner.plan_slabs(plan_in, threshold)
    for k, (_kind, members) in enumerate(slabs):
        size = members[-1][2]
        # guard bytes on both sides: the kernel must not write outside the slab
        dev = torch.full((size + 64,), 0xAB, dtype=torch.uint8, device="cuda")
        base = dev.data_ptr() + 32
        regions = []
        for p, lo, _hi in members:
            regions += pack_regions(gpu[p], base + lo)
        _launch(engine, regions)
        got = _to_host(dev)
        assert got[:32] == b"\xab" * 32 and got[-32:] == b"\xab" * 32
        want = oracle.pack_slab([(cpu[p], lo, hi) for p, lo, hi in members], size)
        assert got[32:-32] == want
        assert sha(want) == gold["slabs"][k]["sha256"]  # and the reference wrote the same bytes


def test_unpack_matches_oracle(engine):
    _threshold, items = _cases.slab_case_mixed()
    cpu = {p: _cases.build_view(spec) for p, spec in items}
    plan_in = [(p, t.numel() * t.element_size(), True, True) for p, t in cpu.items()]
    slabs, _ = planner.plan_slabs(plan_in, 128 << 20)
    members = slabs[0][1]
    size = members[-1][2]
    slab = oracle.pack_slab([(cpu[p], lo, hi) for p, lo, hi in members], size)
    dev_slab = torch.from_numpy(np.frombuffer(slab, dtype=np.uint8).copy()).cuda()
    # destinations with the same (strided) geometry, pre-filled with a marker
    outs = {}
    bases = {}
    regions = []
    for p, spec in items:
        base = torch.full(spec["shape"], 0, dtype=_cases.DTYPES[spec["dtype"]]).view(torch.uint8).fill_(0xCD)
        base = base.view(_cases.DTYPES[spec["dtype"]]).reshape(spec["shape"]).cuda()
        bases[p] = base
    for p, lo, _hi in members:
        spec = dict(items)[p]
        view = _view_of(bases[p], spec["view"])
        outs[p] = view
        regions += unpack_regions(dev_slab.data_ptr() + lo, view)
    _launch(engine, regions)
    for p, lo, hi in members:
        spec = dict(items)[p]
        # oracle: deserialize into an identically shaped CPU view of a marker-filled base
        cbase = torch.zeros(spec["shape"], dtype=_cases.DTYPES[spec["dtype"]]).view(torch.uint8).fill_(0xCD)
        cbase = cbase.view(_cases.DTYPES[spec["dtype"]]).reshape(spec["shape"])
        oracle.deserialize_into(_view_of(cbase, spec["view"]), slab[lo:hi])
        assert _to_host(bases[p].view(torch.uint8)) == cbase.view(torch.uint8).numpy().tobytes(), p


def _view_of(base, view):
    if view == "dense":
        return base
    if view == "scalar":
        return base.reshape(-1)[0]
    if view == "transpose":
        return base.transpose(0, 1)
    if view == "stride2":
        return base[:, ::2]
    if view == "offset":
        return base.reshape(-1)[3:]
    if view == "inner":
        return base[1:-1, 2:-1]
    if view == "narrow1":
        return base.narrow(1, 1, base.shape[1] - 2)
    if view == "permute3":
        return base.permute(2, 0, 1)
    raise ValueError(view)


def test_alignment_sweep(engine):
    """Every (src, dst) alignment pair modulo 16 and lengths around every vector boundary, in one launch."""
    lengths = [0, 1, 2, 3, 4, 5, 15, 16, 17, 31, 32, 33, 47, 63, 64, 65, 127, 255, 256, 257, 1000, 4095, 4096, 4097,
               16383, 16385, 65535, 65536, 65537, 200001]
    rs = np.random.RandomState(3)
    jobs = []
    src_total = dst_total = 0
    for n in lengths:
        for sa in range(16):
            for da in (0, 1, 2, 3, 4, 7, 8, 12, 13, 15):
                s_off = (src_total + 15) // 16 * 16 + sa
                d_off = (dst_total + 15) // 16 * 16 + 16 + da
                jobs.append((s_off, d_off, n))
                src_total = s_off + n
                dst_total = d_off + n
    src_h = rs.randint(0, 256, size=src_total + 64).astype(np.uint8)
    src = torch.from_numpy(src_h).cuda()
    dst = torch.full((dst_total + 64,), 0xEE, dtype=torch.uint8, device="cuda")
    regions = [(src.data_ptr() + s, dst.data_ptr() + d, n, (), (), ()) for s, d, n in jobs if n]
    _launch(engine, regions)
    want = np.full(dst_total + 64, 0xEE, dtype=np.uint8)
    for s, d, n in jobs:
        want[d : d + n] = src_h[s : s + n]
    got = dst.cpu().numpy()
    assert np.array_equal(got, want)


@pytest.mark.parametrize("run_elems,rows,pitch_extra", [(8192, 9, 40), (4100, 33, 3), (512, 300, 7), (37, 1000, 5),
                                                         (16, 5000, 1), (3, 7000, 2), (1, 9000, 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.int16, torch.uint8])
def test_strided_rows_all_classes(engine, run_elems, rows, pitch_extra, dtype):
    """Pitched rows of every run-length class (CTA / warp / element), packed and unpacked."""
    es = torch.empty((), dtype=dtype).element_size()
    base = _cases.raw_tensor(11, [rows, run_elems + pitch_extra], dtype)
    view_c = base[:, 1 : 1 + run_elems] if pitch_extra > 1 else base[:, :run_elems]
    gbase = base.cuda()
    view_g = gbase[:, 1 : 1 + run_elems] if pitch_extra > 1 else gbase[:, :run_elems]
    n = rows * run_elems * es
    slab = torch.full((n + 23,), 0x11, dtype=torch.uint8, device="cuda")
    _launch(engine, pack_regions(view_g, slab.data_ptr() + 5))  # odd slab offset
    got = _to_host(slab)
    assert got[5 : 5 + n] == oracle.serialize(view_c)
    assert got[:5] == b"\x11" * 5 and got[5 + n :] == b"\x11" * 18
    # unpack into a marker-filled twin
    out = torch.zeros_like(gbase).view(torch.uint8).fill_(0x77).view(dtype).reshape(gbase.shape)
    out_v = out[:, 1 : 1 + run_elems] if pitch_extra > 1 else out[:, :run_elems]
    _launch(engine, unpack_regions(slab.data_ptr() + 5, out_v))
    want = torch.zeros_like(base).view(torch.uint8).fill_(0x77).view(dtype).reshape(base.shape)
    want_v = want[:, 1 : 1 + run_elems] if pitch_extra > 1 else want[:, :run_elems]
    oracle.deserialize_into(want_v, oracle.serialize(view_c))
    assert _to_host(out.view(torch.uint8)) == want.view(torch.uint8).numpy().tobytes()


@pytest.mark.parametrize("dtype", [torch.uint8, torch.int16, torch.float32, torch.float64])
@pytest.mark.parametrize("shape", [(1000, 777), (64, 64), (17, 4099), (3, 130, 257), (5, 2, 70, 33)])
def test_transposed_views_pack_and_unpack(engine, dtype, shape):
    """Views whose fastest dimension differs between source and destination (x.t(), permutes of the last two
    dims under batch dims): the tiled transpose kernel, packed against the oracle and unpacked into a
    marker-filled twin; extents that are not multiples of the tile; and the element-by-element path
    (TSB200_TRANSPOSE=0 is read once per process, so the plain strided kernel is checked through a region the
    transpose class does not take: an extent below 16)."""
    base = _cases.raw_tensor(31, list(shape), dtype)
    perm = list(range(len(shape)))
    perm[-1], perm[-2] = perm[-2], perm[-1]
    view_c = base.permute(perm)
    gbase = base.cuda()
    view_g = gbase.permute(perm)
    n = view_c.numel() * view_c.element_size()
    slab = torch.full((n + 41,), 0x11, dtype=torch.uint8, device="cuda")
    before = engine.ctx.counters()["kernel_launches"]
    _launch(engine, pack_regions(view_g, slab.data_ptr() + 16))
    assert engine.ctx.counters()["kernel_launches"] == before + 1  # the transpose kernel alone
    got = _to_host(slab)
    assert got[16 : 16 + n] == oracle.serialize(view_c)
    assert got[:16] == b"\x11" * 16 and got[16 + n :] == b"\x11" * 25
    out = torch.zeros_like(gbase).view(torch.uint8).fill_(0x77).view(dtype).reshape(gbase.shape)
    _launch(engine, unpack_regions(slab.data_ptr() + 16, out.permute(perm)))
    assert _to_host(out.view(torch.uint8)) == base.view(torch.uint8).numpy().tobytes()
    # a strided column slice of the transposed view as the source, a pitched destination: both sides strided
    if len(shape) == 2 and shape[0] >= 64:
        src = gbase.t()[3:50:2]  # rows 3, 5, ... of the transposed view
        dst = torch.zeros(src.shape[0], src.shape[1] + 5, dtype=dtype, device="cuda")
        from torchsnapshot_b200.regions import box_regions

        es = src.element_size()
        _launch(engine, box_regions(src.data_ptr(), [x * es for x in src.stride()], dst.data_ptr(),
                                    [x * es for x in dst.stride()], list(src.shape), es))
        assert torch.equal(dst[:, : src.shape[1]].cpu(), base.t()[3:50:2]) and not dst[:, src.shape[1] :].any()


def test_high_rank_views(engine):
    """5-d permuted view: more outer dims than one region holds, so the planner splits it."""
    base = _cases.raw_tensor(21, [3, 4, 5, 6, 7, 8], torch.int32)
    view_c = base.permute(5, 3, 1, 0, 2, 4)[:, 1:5]
    view_g = base.cuda().permute(5, 3, 1, 0, 2, 4)[:, 1:5]
    n = view_c.numel() * 4
    slab = torch.zeros(n, dtype=torch.uint8, device="cuda")
    regions = pack_regions(view_g, slab.data_ptr())
    assert len(regions) > 1
    _launch(engine, regions)
    assert _to_host(slab) == oracle.serialize(view_c)


def test_reshard_boxes_match_reference_fixture(engine, golden_dir):
    """Every src x dst layout of the reshard fixture, executed as box regions on the GPU."""
    from torchsnapshot_b200.regions import contiguous_strides

    gold = json.load(open(os.path.join(golden_dir, "sharded.json")))
    shape = _cases.SHARD_GLOBAL_SHAPE
    full = _cases.raw_tensor(_cases.SHARD_SEED, shape, _cases.DTYPES[_cases.SHARD_DTYPE])
    gfull = full.cuda()
    layouts = _cases.shard_layouts(shape)
    for key, g in gold["reshard"].items():
        src_name, dst_name = key.split("->")
        pieces = {p["location"]: p for p in gold["writes"][src_name]["pieces"]}
        dst_boxes = layouts[dst_name][1]
        dst = [torch.zeros(s, dtype=full.dtype, device="cuda") for _o, s in dst_boxes]
        regions = []
        keep = []
        for item in g["plan"]:
            piece = pieces[item["location"]]
            dense = _cases.box_slice(gfull, (piece["offsets"], piece["sizes"])).contiguous()  # the persisted bytes
            keep.append(dense)
            sstr = contiguous_strides(piece["sizes"], 4)
            for reg in item["regions"]:
                d = dst[reg["dst_shard"]]
                dstr = [s * 4 for s in d.stride()]
                sa, da, lens = dense.data_ptr(), d.data_ptr(), []
                for dim, so, do, ln in reg["overlap"]:
                    sa += so * sstr[dim]
                    da += do * dstr[dim]
                    lens.append(ln)
                regions += box_regions(sa, sstr, da, dstr, lens, 4)
        _launch(engine, regions)
        assert [sha(_to_host(d)) for d in dst] == g["dst_sha256"], key


def test_full_size_round_trip_c3(engine):
    """BASELINE config 2 (many-small-tensors stress, ~8 GB per rank): pack -> unpack round trip and a
    checksum of checksums, properties that do not need the oracle to touch 8 GB on the host."""
    free, _ = torch.cuda.mem_get_info()
    target = min(8 << 30, int(free * 0.25))
    rs = np.random.RandomState(5)
    sizes = []
    total = 0
    while total < target:
        n = 4 * int(round(2 ** rs.uniform(10, 20)))
        sizes.append(n)
        total += n
    g = torch.Generator(device="cuda").manual_seed(5)
    pool = torch.randint(0, 256, (total + 64,), dtype=torch.uint8, device="cuda", generator=g)
    offs = np.concatenate([[0], np.cumsum(sizes)[:-1]])
    # every tenth tensor starts at an odd source address; slab offsets are unaligned after the first odd size
    srcs = [pool[o + (i % 10 == 0) : o + (i % 10 == 0) + n - 1] if i % 10 == 0 else pool[o : o + n]
            for i, (o, n) in enumerate(zip(offs, sizes))]
    lens = [t.numel() for t in srcs]
    slab_bytes = int(sum(lens))
    slab = torch.empty(slab_bytes, dtype=torch.uint8, device="cuda")
    dst_offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    regions = [(t.data_ptr(), slab.data_ptr() + int(d), n, (), (), ()) for t, d, n in zip(srcs, dst_offs, lens)]
    arr = regions_to_array(regions)
    engine.wait_current_stream(_native.STREAM_PACK)
    engine.copy_region_array(arr)
    engine.ctx.stream_sync(_native.STREAM_PACK)
    # checksum of checksums: the packed slab holds exactly the members' bytes, in order
    want_sum = sum(int(t.sum(dtype=torch.int64)) for t in srcs[::97])
    got_sum = sum(int(slab[int(d) : int(d) + n].sum(dtype=torch.int64)) for d, n in list(zip(dst_offs, lens))[::97])
    assert want_sum == got_sum
    for t, d, n in list(zip(srcs, dst_offs, lens))[::211]:
        assert torch.equal(slab[int(d) : int(d) + n], t)
    # round trip
    out = torch.zeros_like(pool)
    outs = [out[o + (i % 10 == 0) : o + (i % 10 == 0) + n - 1] if i % 10 == 0 else out[o : o + n]
            for i, (o, n) in enumerate(zip(offs, sizes))]
    back = [(slab.data_ptr() + int(d), t.data_ptr(), n, (), (), ()) for t, d, n in zip(outs, dst_offs, lens)]
    engine.copy_region_array(regions_to_array(back))
    engine.ctx.stream_sync(_native.STREAM_PACK)
    # bytes not covered by a member (one per odd tensor, plus the tail) stay zero in `out`
    mask = torch.ones_like(pool, dtype=torch.bool)
    for i, (o, n) in enumerate(zip(offs, sizes)):
        if i % 10 == 0:
            mask[o] = False
    mask[total:] = False
    assert torch.equal(out[mask], pool[mask])
    assert int(out[~mask].sum()) == 0


@pytest.mark.parametrize("use_tma", ["1", "0"])
def test_tma_and_lsu_variants_of_the_flat_kernel(golden_dir, monkeypatch, use_tma):
    """Default: tiles aligned on both sides go through cp.async.bulk (TMA), everything else through the LSU
    path of the same kernel.  TSB200_USE_TMA=0: the pure LSU kernel.  Both must give the oracle's bytes in
    every alignment case."""
    from torchsnapshot_b200 import engine as engmod

    engmod.shutdown_engines()
    monkeypatch.setenv("TSB200_USE_TMA", use_tma)
    try:
        eng = engmod.get_engine(0)
        rs = np.random.RandomState(8)
        src_h = rs.randint(0, 256, size=(6 << 20) + 64).astype(np.uint8)
        src = torch.from_numpy(src_h).cuda()
        jobs = []
        d_off = 0
        for n in (16, 4096, 32768, 32768 + 16, 1 << 20, (1 << 20) + 5, 100000):
            for sa, da in ((0, 0), (0, 16), (16, 0), (0, 4), (4, 0), (1, 1), (3, 8)):
                jobs.append((64 + sa, d_off + 32 + da, n))
                d_off += n + 96
        dst = torch.full((d_off + 64,), 0xEE, dtype=torch.uint8, device="cuda")
        _launch(eng, [(src.data_ptr() + s, dst.data_ptr() + d, n, (), (), ()) for s, d, n in jobs])
        want = np.full(d_off + 64, 0xEE, dtype=np.uint8)
        for s, d, n in jobs:
            want[d : d + n] = src_h[s : s + n]
        assert np.array_equal(dst.cpu().numpy(), want)
        # and the reference-made slab fixture
        gold = json.load(open(os.path.join(golden_dir, "slab_mixed.json")))
        threshold, items = _cases.slab_case_mixed()
        cpu = {p: _cases.build_view(spec) for p, spec in items}
        gpu = {p: _cases.build_view(spec, "cuda") for p, spec in items}
        slabs, _ = planner.plan_slabs([(p, t.numel() * t.element_size(), True, True) for p, t in cpu.items()], threshold)
        members = slabs[0][1]
        slab = torch.zeros(members[-1][2], dtype=torch.uint8, device="cuda")
        regions = []
        for p, lo, _hi in members:
            regions += pack_regions(gpu[p], slab.data_ptr() + lo)
        _launch(eng, regions)
        assert sha(_to_host(slab)) == gold["slabs"][0]["sha256"]
    finally:
        engmod.shutdown_engines()


def _fill_pattern(t: torch.Tensor, seed: int) -> None:
    # cheap, position-dependent bytes for multi-GiB buffers (a 256 MiB random block repeated with a twist
    # per repetition, so a copy that lands 2^32 bytes off does not go unnoticed)
    g = torch.Generator(device="cuda").manual_seed(seed)
    block = torch.randint(0, 256, (256 << 20,), dtype=torch.uint8, device="cuda", generator=g)
    n = t.numel()
    for i, lo in enumerate(range(0, n, block.numel())):
        hi = min(n, lo + block.numel())
        t[lo:hi] = block[: hi - lo] + (i * 37 & 0xFF)


def test_regions_beyond_4GiB(engine):
    """Maximum sizes: 64-bit byte arithmetic inside ONE region -- a flat run longer than 4 GiB from an odd
    address, long pitched rows spanning more than 4 GiB of source, and short runs whose flat payload
    offset passes 2^32.  Compared on the device against torch's own indexing of the same bytes."""
    free, _ = torch.cuda.mem_get_info()
    if free < (24 << 30):
        pytest.skip("needs 24 GiB of free device memory")
    # (1) flat: 4 GiB + 4099 bytes, source at an odd address, destination 3 bytes into its buffer
    n = (4 << 30) + 4099
    src = torch.empty(5_500_000_001, dtype=torch.uint8, device="cuda")
    _fill_pattern(src, 1)
    dst = torch.zeros(n + 8, dtype=torch.uint8, device="cuda")
    _launch(engine, [(src.data_ptr() + 1, dst.data_ptr() + 3, n, (), (), ())])
    assert torch.equal(dst[3 : 3 + n], src[1 : 1 + n])
    assert int(dst[:3].sum()) == 0 and int(dst[3 + n :].sum()) == 0
    del dst
    # (2) long runs: 40 000 rows of 64 KiB taken from rows 128 KiB apart (source span 5.2 GB)
    rows, run, pitch = 40000, 64 << 10, 128 << 10
    assert rows * pitch <= src.numel() - 1 and rows * pitch > (4 << 30)
    view = src[1 : 1 + rows * pitch].view(rows, pitch)[:, 100 : 100 + run]
    out = torch.zeros(rows * run + 16, dtype=torch.uint8, device="cuda")
    _launch(engine, pack_regions(view, out.data_ptr() + 7))
    assert torch.equal(out[7 : 7 + rows * run].view(rows, run), view)
    assert int(out[:7].sum()) == 0 and int(out[7 + rows * run :].sum()) == 0
    del out, view
    # (3) short runs: 32-byte runs every 48 bytes; 90 M rows -> payload 2.9 GB from a 4.3 GB span, and
    # scattered back (unpack) into a 4.3 GB span: flat offsets and strides x index products pass 2^32
    rows, run, pitch = 90_000_000, 32, 48
    assert rows * pitch > (4 << 30) and rows * pitch <= src.numel() - 1
    view = src[1 : 1 + rows * pitch].view(rows, pitch)[:, 8 : 8 + run]
    packed = torch.zeros(rows * run, dtype=torch.uint8, device="cuda")
    _launch(engine, pack_regions(view, packed.data_ptr()))
    assert torch.equal(packed.view(rows, run)[-1000:], view[-1000:])  # the tail lies beyond 2^32 on both sides
    step = 9_000_000
    for lo in range(0, rows, step):
        assert torch.equal(packed.view(rows, run)[lo : lo + step], view[lo : lo + step])
    back = torch.zeros(rows * pitch, dtype=torch.uint8, device="cuda")
    back_v = back.view(rows, pitch)[:, 8 : 8 + run]
    _launch(engine, unpack_regions(packed.data_ptr(), back_v))
    for lo in range(0, rows, step):
        assert torch.equal(back_v[lo : lo + step], view[lo : lo + step])
    assert int(back.view(rows, pitch)[:, :8].sum()) == 0 and int(back.view(rows, pitch)[:, 8 + run :].sum()) == 0


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_random_views_one_launch_fuzz(engine, seed):
    """~150 random strided views of every buffer-protocol dtype, ranks 0-5, packed back to back at arbitrary
    slab offsets by ONE launch and scattered back by ONE launch; bytes compared with the CPU oracle."""
    rs = np.random.RandomState(1000 + seed)
    dtypes = [torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int64, torch.int32, torch.int16,
              torch.int8, torch.uint8, torch.bool]
    cpu_views, gpu_views, gpu_bases, cpu_bases = [], [], [], []
    for i in range(150):
        rank = int(rs.randint(0, 6))
        budget = int(2 ** rs.uniform(4, 21))  # elements in the base
        shape = []
        for d in range(rank):
            n = max(1, int(round(budget ** (1.0 / (rank - d)) * rs.uniform(0.5, 1.5)))) if budget > 1 else 1
            n = min(n, budget)
            shape.append(n)
            budget = max(budget // n, 1)
        dtype = dtypes[int(rs.randint(0, len(dtypes)))]
        base = _cases.raw_tensor(seed * 1000 + i, shape, dtype)
        gbase = base.cuda()
        cpu_bases.append(base)
        gpu_bases.append(gbase)
        if rank == 0:
            cpu_views.append(base)
            gpu_views.append(gbase)
            continue
        state = rs.get_state()
        cpu_views.append(_cases.random_view(rs, base))
        rs.set_state(state)
        gpu_views.append(_cases.random_view(rs, gbase))
    # slab layout: members back to back with small random gaps, so offsets have every alignment
    offs, total = [], 3
    for v in cpu_views:
        offs.append(total)
        total += v.numel() * v.element_size() + int(rs.randint(0, 4))
    slab = torch.full((total + 16,), 0x5A, dtype=torch.uint8, device="cuda")
    regions = []
    for v, o in zip(gpu_views, offs):
        regions += pack_regions(v, slab.data_ptr() + o)
    _launch(engine, regions)
    want = bytearray(b"\x5a" * (total + 16))
    for v, o in zip(cpu_views, offs):
        b = oracle.serialize(v)
        want[o : o + len(b)] = b
    assert _to_host(slab) == bytes(want)
    # scatter back into marker-filled twins with the same geometry
    twins_g, twins_c, back = [], [], []
    for gb, cb, gv, cv, o in zip(gpu_bases, cpu_bases, gpu_views, cpu_views, offs):
        nbytes = cb.numel() * cb.element_size()
        tg = torch.full((nbytes,), 0xA7, dtype=torch.uint8, device="cuda").view(gb.dtype).reshape(gb.shape)
        tc = torch.full((nbytes,), 0xA7, dtype=torch.uint8).view(cb.dtype).reshape(cb.shape)
        vg = tg.as_strided(gv.shape, gv.stride(), gv.storage_offset())
        vc = tc.as_strided(cv.shape, cv.stride(), cv.storage_offset())
        back += unpack_regions(slab.data_ptr() + o, vg)
        oracle.deserialize_into(vc, bytes(want[o : o + cv.numel() * cv.element_size()]))
        twins_g.append(tg)
        twins_c.append(tc)
    _launch(engine, back)
    for i, (tg, tc) in enumerate(zip(twins_g, twins_c)):
        assert _to_host(tg.reshape(-1).view(torch.uint8)) == tc.reshape(-1).view(torch.uint8).numpy().tobytes(), i


def test_exportable_block_is_its_own_allocation(engine):
    """tsb_dev_alloc_exportable: a block a peer can map on its own (offset 0 of its own IPC handle, unlike
    arena blocks, which export their whole segment), usable as a copy target, freed through tsb_dev_free."""
    ctx = engine.ctx
    n = (3 << 20) + 123
    block = ctx.dev_alloc_exportable(n)
    arena_a = ctx.dev_alloc(1 << 20)
    arena_b = ctx.dev_alloc(1 << 20)
    try:
        handle, offset = ctx.ipc_export(block)
        assert offset == 0 and len(handle) == 64
        _h2, inner = ctx.ipc_export(block + 4096)
        assert inner == 4096 and _h2 == handle
        ha, off_a = ctx.ipc_export(arena_a)
        hb, off_b = ctx.ipc_export(arena_b)
        assert ha == hb and off_a != off_b and ha != handle  # arena blocks share their segment's handle
        src = torch.arange(n, dtype=torch.int64, device="cuda").to(torch.uint8)
        _launch(engine, [(src.data_ptr(), block, n, (), (), ())])
        back = torch.zeros(n, dtype=torch.uint8, device="cuda")
        _launch(engine, [(block, back.data_ptr(), n, (), (), ())])
        assert torch.equal(back, src)
    finally:
        ctx.dev_free(block)
        ctx.dev_free(arena_a)
        ctx.dev_free(arena_b)
    with pytest.raises(_native.TsbError):
        ctx.dev_free(block)
