"""Descriptor math of regions.py, checked on the CPU: regions built for HOST tensors are executed
by a memmove interpreter (tests/_sim.py) and compared with the oracle's serialisation."""

import ctypes

import numpy as np
import pytest
import torch

import _cases
import _sim
import oracle
from torchsnapshot_b200._native import REGION_DTYPE, TSB_MAX_DIMS
from torchsnapshot_b200.regions import (
    box_regions,
    contiguous_strides,
    pack_regions,
    regions_payload_bytes,
    regions_to_array,
    unpack_regions,
)

VIEWS = ["dense", "scalar", "transpose", "stride2", "offset", "inner", "narrow1", "empty"]


def _specs():
    out = []
    seed = 0
    for dt in _cases.DTYPES:
        for view in VIEWS:
            shape = [7, 12] if view != "scalar" else [4]
            out.append({"seed": seed, "shape": shape, "dtype": dt, "view": view})
            seed += 1
        out.append({"seed": seed, "shape": [3, 4, 5], "dtype": dt, "view": "permute3"})
    return out


@pytest.mark.parametrize("spec", _specs(), ids=lambda s: f'{s["dtype"]}-{s["view"]}')
def test_pack_and_unpack_regions_on_host(spec):
    t = _cases.build_view(spec)
    n = t.numel() * t.element_size()
    buf = np.full(n + 16, 0x5A, dtype=np.uint8)
    regions = pack_regions(t, buf.ctypes.data + 3)
    assert regions_payload_bytes(regions) == n
    assert all(len(r[3]) <= TSB_MAX_DIMS for r in regions)
    assert _sim.run_regions(regions) == n
    assert buf[3 : 3 + n].tobytes() == oracle.serialize(t)
    assert buf[:3].tobytes() == b"\x5a" * 3 and buf[3 + n :].tobytes() == b"\x5a" * 13
    # and back into a zeroed strided twin (same geometry as the source view)
    twin = _cases.build_view({**spec, "seed": spec["seed"] + 7777})
    if twin.numel():
        twin.zero_()
    _sim.run_regions(unpack_regions(buf.ctypes.data + 3, twin))
    assert oracle.serialize(twin) == oracle.serialize(t)


def test_contiguous_tensor_is_one_flat_region():
    t = torch.arange(24, dtype=torch.float32).reshape(2, 3, 4)
    (r,) = pack_regions(t, 4096)
    assert r == (t.data_ptr(), 4096, 96, (), (), ())
    # a dense slice along dim 0 stays flat; a slice along dim 1 becomes pitched rows
    (r,) = pack_regions(t[1:], 4096)
    assert r[2] == 48 and r[3] == ()
    (r,) = pack_regions(t[:, 1:], 4096)
    assert r[2] == 32 and r[3] == (2,) and r[4] == (48,) and r[5] == (32,)


def test_more_outer_dims_than_a_region_holds():
    base = _cases.raw_tensor(1, [3, 4, 5, 6, 7, 8], torch.int16)
    view = base.permute(5, 3, 1, 0, 2, 4)[:, 1:5]
    n = view.numel() * 2
    buf = np.zeros(n, dtype=np.uint8)
    regions = pack_regions(view, buf.ctypes.data)
    assert len(regions) > 1 and all(len(r[3]) <= TSB_MAX_DIMS for r in regions)
    _sim.run_regions(regions)
    assert buf.tobytes() == oracle.serialize(view)


def test_box_regions_between_two_strided_views():
    src = _cases.raw_tensor(2, [10, 12], torch.float32)
    dst = torch.zeros(9, 20)
    es = 4
    regions = box_regions(
        src.data_ptr() + (2 * 12 + 3) * es, [s * es for s in src.stride()],
        dst.data_ptr() + (1 * 20 + 5) * es, [s * es for s in dst.stride()],
        [6, 7], es,
    )
    _sim.run_regions(regions)
    want = torch.zeros(9, 20)
    want[1:7, 5:12] = src[2:8, 3:10]
    assert torch.equal(dst, want)


def test_region_array_layout():
    regions = [(1, 2, 3, (), (), ()), (10, 20, 4, (5, 6), (100, -8), (24, 4))]
    arr = regions_to_array(regions)
    assert arr.dtype == REGION_DTYPE and arr.nbytes == 256
    raw = np.frombuffer(arr.tobytes(), dtype=np.int64).reshape(2, 16)
    assert list(raw[0][:3]) == [1, 2, 3] and raw[0][3] == 0
    assert list(raw[1][:3]) == [10, 20, 4] and (raw[1][3] & 0xFFFFFFFF) == 2
    assert list(raw[1][4:6]) == [5, 6] and list(raw[1][8:10]) == [100, -8] and list(raw[1][12:14]) == [24, 4]


def test_contiguous_strides():
    assert contiguous_strides([2, 3, 4], 2) == [24, 8, 2]
    assert contiguous_strides([], 4) == []


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_views_fuzz_on_host(seed):
    """The CPU twin of tests/test_gpu_pack.py::test_random_views_one_launch_fuzz: ~200 random strided views
    (ranks 0-5, every buffer-protocol dtype, slices with steps, permutations) -> regions -> memmove
    interpreter, compared with the oracle's serialisation, both directions."""
    rs = np.random.RandomState(500 + seed)
    dtypes = list(_cases.DTYPES.values())
    for i in range(200):
        rank = int(rs.randint(0, 6))
        shape = _cases.random_base_shape(rs, rank, 13)
        dtype = dtypes[int(rs.randint(0, len(dtypes)))]
        base = _cases.raw_tensor(seed * 1000 + i, shape, dtype)
        state = rs.get_state()
        view = _cases.random_view(rs, base) if rank else base
        n = view.numel() * view.element_size()
        off = int(rs.randint(0, 16))
        buf = np.full(n + 32, 0x3C, dtype=np.uint8)
        regions = pack_regions(view, buf.ctypes.data + off)
        assert regions_payload_bytes(regions) == n and all(len(r[3]) <= TSB_MAX_DIMS for r in regions)
        _sim.run_regions(regions)
        want = oracle.serialize(view)
        assert buf[off : off + n].tobytes() == want, (i, shape, view.shape, view.stride())
        assert buf[:off].tobytes() == b"\x3c" * off and buf[off + n :].tobytes() == b"\x3c" * (32 - off)
        # scatter back into a marker-filled twin with the same geometry; bytes outside the view stay markers
        nbytes = base.numel() * base.element_size()
        twin = torch.full((nbytes,), 0xA7, dtype=torch.uint8).view(dtype).reshape(base.shape)
        ref = torch.full((nbytes,), 0xA7, dtype=torch.uint8).view(dtype).reshape(base.shape)
        rs2 = np.random.RandomState(0)
        rs2.set_state(state)
        tv = _cases.random_view(rs2, twin) if rank else twin
        rs2.set_state(state)
        rv = _cases.random_view(rs2, ref) if rank else ref
        _sim.run_regions(unpack_regions(buf.ctypes.data + off, tv))
        oracle.deserialize_into(rv, want)
        assert twin.reshape(-1).view(torch.uint8).numpy().tobytes() == ref.reshape(-1).view(torch.uint8).numpy().tobytes(), i
