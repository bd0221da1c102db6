"""tsb_copy_regions_static -- the multi-region copy for callers that are not allowed to allocate.

Some users of this library run where a cudaMalloc is off the table: the training process has given all
of HBM to its caching allocator, or the copy is issued from a place that must not trigger the
device-wide synchronisation an allocation implies.  For them the library carries a second entry point
that works out of a fixed 2 MiB array inside the library instead of the context's arenas.  This test
drives it through the C-ABI and holds its output to the output of the regular entry point
(tsb_copy_regions) on the same regions, byte for byte, on two calls in a row -- the second one with a
different, smaller job, so anything the first call left behind in the shared array would show -- and to
the bytes the CPU oracle computes from host copies of the same tensors (two CUDA paths that agree could
still both be wrong).

The file name sorts last on purpose: a run that stops at the first failure has then already run every
other test.  It also has to pass on its own in a fresh process, where its first call performs the
one-time per-device set-up.
"""

import numpy as np
import pytest
import torch

import _cases
import oracle
from torchsnapshot_b200 import _native
from torchsnapshot_b200.regions import box_regions, pack_regions, regions_to_array

pytestmark = pytest.mark.gpu


def _jobs(device):
    """Two region lists with every kind of work the kernels know: aligned and misaligned contiguous
    runs, long pitched rows (expanded into flat tiles), short pitched rows (one per warp) and
    single-element runs (one per thread)."""
    rs = np.random.RandomState(17)
    pool = torch.from_numpy(rs.randint(0, 256, size=24 << 20).astype(np.uint8)).to(device)

    def job(seed, n_flat, shift):
        r = np.random.RandomState(seed)
        tensors = []
        off = 0
        for i in range(n_flat):
            n = int(r.randint(1, 300000))
            tensors.append(pool[off + (i % 3) : off + (i % 3) + n])
            off += n + 64
        off = (off + 255) // 256 * 256  # the typed views below need an aligned storage offset
        mat = pool[off : off + (1 << 20)].view(torch.int16).reshape(512, 1024)
        views = [mat[:, 3:900], mat[:, ::2], mat[5:200, 100:132]]
        big = pool[off + (1 << 20) : off + (9 << 20)].view(torch.float32).reshape(128, 16384)
        views.append(big[:, 17:9000])  # rows of ~36 KB: long runs
        return tensors + views, shift

    return [job(1, 60, 5), job(2, 9, 1)]


def _regions_for(members, dst_base, shift):
    out = []
    off = shift
    for t in members:
        out += pack_regions(t, dst_base + off)
        off += t.numel() * t.element_size()
    return out, off


def test_static_scratch_copy_matches_regular_entry_point(engine):
    device = engine.device
    stream = torch.cuda.current_stream().cuda_stream
    for members, shift in _jobs(f"cuda:{device}"):
        total = sum(t.numel() * t.element_size() for t in members) + shift
        want = torch.full((total + 32,), 0x3C, dtype=torch.uint8, device=f"cuda:{device}")
        got = torch.full((total + 32,), 0x3C, dtype=torch.uint8, device=f"cuda:{device}")
        regions, end = _regions_for(members, want.data_ptr(), shift)
        assert end == total
        torch.cuda.synchronize()
        # the existing entry point
        engine.wait_current_stream(_native.STREAM_PACK)
        engine.copy_regions(regions, _native.STREAM_PACK)
        engine.ctx.stream_sync(_native.STREAM_PACK)
        # the entry point that allocates nothing, through the C-ABI, on torch's current stream
        regions2, _ = _regions_for(members, got.data_ptr(), shift)
        _native.copy_regions_static(device, stream, regions_to_array(regions2))
        torch.cuda.synchronize()
        assert torch.equal(got, want)  # bytes: no tolerance
        assert int(want[:shift].sum()) == 0x3C * shift and int(want[total:].sum()) == 0x3C * 32
        # and both are what the reference's serialisation of the same tensors gives, member after member
        expect = b"\x3c" * shift + b"".join(oracle.serialize(t.cpu()) for t in members) + b"\x3c" * 32
        assert got.cpu().numpy().tobytes() == expect


def test_static_scratch_copy_is_refused_on_a_capturing_stream(engine):
    """The descriptor upload reads a host buffer the next call overwrites, so the call must not end up inside a
    CUDA graph: on a stream that is being captured it is refused, nothing is enqueued, and the capture goes on."""
    device = engine.device
    src = torch.arange(4096, dtype=torch.uint8, device=f"cuda:{device}")
    dst = torch.zeros(4096, dtype=torch.uint8, device=f"cuda:{device}")
    arr = regions_to_array([(src.data_ptr(), dst.data_ptr(), 4096, (), (), ())])
    _native.copy_regions_static(device, torch.cuda.current_stream().cuda_stream, arr)  # set-up call, outside any capture
    torch.cuda.synchronize()
    assert torch.equal(dst, src)
    dst.zero_()
    side = torch.cuda.Stream(device=device)
    graph = torch.cuda.CUDAGraph()
    x = torch.zeros(8, device=f"cuda:{device}")
    torch.cuda.synchronize()
    with torch.cuda.graph(graph, stream=side):
        x += 1
        with pytest.raises(_native.TsbError, match="captured"):
            _native.copy_regions_static(device, side.cuda_stream, arr)
    graph.replay()
    torch.cuda.synchronize()
    assert int(dst.sum()) == 0 and float(x.sum()) == 8.0


def test_static_scratch_limits_are_refused_before_any_work(engine):
    device = engine.device
    stream = torch.cuda.current_stream().cuda_stream
    src = torch.zeros(1 << 20, dtype=torch.uint8, device=f"cuda:{device}")
    dst = torch.full((1 << 20,), 7, dtype=torch.uint8, device=f"cuda:{device}")
    # more payload than the tile table holds (addresses are never touched: the call is refused first)
    too_big = regions_to_array([(src.data_ptr(), dst.data_ptr(), 4 << 30, (), (), ())])
    with pytest.raises(_native.TsbError, match="static scratch holds"):
        _native.copy_regions_static(device, stream, too_big)
    # more descriptors than the upload area holds
    many = regions_to_array([(src.data_ptr() + i, dst.data_ptr() + i, 1, (), (), ()) for i in range(20000)])
    with pytest.raises(_native.TsbError, match="static scratch holds"):
        _native.copy_regions_static(device, stream, many)
    torch.cuda.synchronize()
    assert int(dst.min()) == 7  # nothing was written
    # and it still works afterwards
    ok = regions_to_array([(src.data_ptr(), dst.data_ptr() + 3, 1000, (), (), ())])
    _native.copy_regions_static(device, stream, ok)
    torch.cuda.synchronize()
    assert int(dst[3:1003].sum()) == 0 and int(dst[:3].sum()) == 21 and int(dst[1003]) == 7
