"""The C-ABI library loads and exports every symbol include/tsb200.h declares (no compute here)."""

import ctypes
import os
import re

import pytest
import torch

from torchsnapshot_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_in_header():
    text = open(os.path.join(ROOT, "include", "tsb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tsb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert _declared_in_header() == _native.declared_symbols()


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_native.LIB_PATH)
    for name in _declared_in_header():
        assert hasattr(lib, name), name


def test_version_and_struct_layout():
    lib = _native.load_library()
    assert lib.tsb_version() // 1000 == 1  # major version of the ABI
    assert _native.REGION_DTYPE.itemsize == 128
    assert _native.REGION_DTYPE.fields["size"][1] == 32
    assert _native.REGION_DTYPE.fields["src_stride"][1] == 64
    assert _native.REGION_DTYPE.fields["dst_stride"][1] == 96


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_context_creation_fails_loudly_without_gpu():
    with pytest.raises(_native.TsbError):
        _native.Context(0)


def test_missing_library_is_an_error(tmp_path):
    with pytest.raises(_native.TsbError):
        _native.load_library(str(tmp_path / "libtsb200.so"))


def test_host_helpers(tmp_path):
    # host-side helpers need no GPU: parallel memcpy and the file fast path
    import numpy as np

    src = np.random.RandomState(0).randint(0, 256, size=(9 << 20) + 13).astype(np.uint8)
    dst = np.zeros_like(src)
    _native.host_memcpy(dst.ctypes.data, src.ctypes.data, src.nbytes, 4)
    assert np.array_equal(src, dst)
    path = str(tmp_path / "f.bin")
    _native.file_write(path, src.ctypes.data, src.nbytes, 4)
    assert os.path.getsize(path) == src.nbytes
    back = np.zeros(1 << 20, dtype=np.uint8)
    _native.file_read(path, 12345, back.ctypes.data, back.nbytes, 2)
    assert np.array_equal(back, src[12345 : 12345 + (1 << 20)])
    with pytest.raises(_native.TsbError):
        _native.file_read(str(tmp_path / "missing"), 0, back.ctypes.data, 16, 1)


def test_file_direct_flag_never_changes_the_bytes(tmp_path):
    # TSB_FILE_DIRECT moves the page-aligned body with O_DIRECT where the file system takes it and falls
    # back to the buffered path otherwise (tmpfs, unaligned buffers / offsets, the sub-page tail)
    import mmap

    import numpy as np

    n = (5 << 20) + 777
    payload = np.random.RandomState(1).randint(0, 256, size=n).astype(np.uint8)
    page = mmap.mmap(-1, n + 4096)  # page-aligned, like a pinned arena block
    aligned = np.frombuffer(page, dtype=np.uint8, count=n)
    aligned[:] = payload
    for root in (str(tmp_path), "/dev/shm"):
        for name, arr in (("aligned", aligned), ("unaligned", np.frombuffer(page, dtype=np.uint8, count=n - 1, offset=1))):
            path = os.path.join(root, f"tsb_direct_{os.getpid()}_{name}.bin")
            try:
                _native.file_write(path, arr.ctypes.data, arr.nbytes, 3, _native.FILE_DIRECT)
                with open(path, "rb") as f:
                    assert f.read() == arr.tobytes()
                back_page = mmap.mmap(-1, n + 4096)
                back = np.frombuffer(back_page, dtype=np.uint8, count=1 << 20)
                for off in (0, 4096 * 3, 12345):
                    back[:] = 0
                    _native.file_read(path, off, back.ctypes.data, back.nbytes, 2, _native.FILE_DIRECT)
                    assert np.array_equal(back, arr[off : off + (1 << 20)])
                del back
                back_page.close()
            finally:
                if os.path.exists(path):
                    os.remove(path)
    with pytest.raises(_native.TsbError):
        _native.file_write(str(tmp_path / "x"), aligned.ctypes.data, 16, 1, 0x80)
    del aligned, arr
    page.close()


def test_host_gather_splits_work_by_bytes():
    # many copies of very different sizes, more threads than large copies: every byte lands once, nothing else moves
    import numpy as np

    rs = np.random.RandomState(3)
    sizes = [0, 1, 7, 4096, 100003, 5 << 20, 33 << 20, 12345, 64, 63]
    srcs = [rs.randint(0, 256, size=n).astype(np.uint8) for n in sizes]
    total = sum(sizes)
    for nthreads in (1, 3, 8):
        slab = np.full(total + 8, 0xAB, dtype=np.uint8)
        offs = np.concatenate([[0], np.cumsum(sizes)[:-1]])
        _native.host_gather([slab.ctypes.data + 4 + int(o) for o in offs],
                            [a.ctypes.data if a.nbytes else 0 for a in srcs], sizes, nthreads)
        assert slab[4 : 4 + total].tobytes() == b"".join(a.tobytes() for a in srcs)
        assert slab[:4].tobytes() == b"\xab" * 4 and slab[4 + total :].tobytes() == b"\xab" * 4
    _native.host_gather([], [], [])  # nothing to do
    with pytest.raises(_native.TsbError):
        _native.host_gather([0], [srcs[3].ctypes.data], [16])
    with pytest.raises(ValueError):
        _native.host_gather([1, 2], [1], [1])


@pytest.mark.gpu
def test_interleaved_copies_and_streamed_file_writes(engine, tmp_path):
    """tsb_memcpy_d2h_interleaved + tsb_file_write_streamed through the C-ABI: three device buffers of awkward
    sizes cross the link chunk by chunk, round-robin, one event per chunk; a writer per buffer follows its events;
    the files hold exactly the device bytes.  Also tsb_host_alloc_nogrow: a block out of the existing segments or
    TSB_EAGAIN, never new pinned memory."""
    import numpy as np

    ctx = engine.ctx
    chunk = 1 << 20
    sizes = [(5 << 20) + 12345, 1, 3 << 20, 0]
    g = torch.Generator(device="cuda").manual_seed(11)
    srcs = [torch.randint(0, 256, (max(n, 1),), dtype=torch.uint8, device="cuda", generator=g)[:n] for n in sizes]
    torch.cuda.synchronize()
    hptrs = [ctx.host_alloc(max(n, 1)) for n in sizes]
    try:
        feeds = ctx.memcpy_d2h_interleaved(_native.STREAM_D2H, hptrs, [max(t.data_ptr(), 1) for t in srcs], sizes, chunk)
        assert [len(ev) for ev, _ends in feeds] == [6, 1, 3, 0]
        assert [int(ends[-1]) for _ev, ends in feeds if len(ends)] == [sizes[0], 1, sizes[2]]
        for i, ((events, ends), n) in enumerate(zip(feeds, sizes)):
            path = str(tmp_path / f"f{i}")
            ctx.file_write_streamed(path, hptrs[i], n, events, ends)
            with open(path, "rb") as f:
                assert f.read() == srcs[i].cpu().numpy().tobytes(), i
            for ev in events:
                ctx.event_release(int(ev))
        # events that do not cover the buffer are an error, not a short file that looks complete
        events, ends = ctx.memcpy_d2h_interleaved(_native.STREAM_D2H, hptrs[:1], [srcs[0].data_ptr()], [2 << 20], chunk)[0]
        with pytest.raises(_native.TsbError):
            ctx.file_write_streamed(str(tmp_path / "short"), hptrs[0], 3 << 20, events, ends)
        for ev in events:
            ctx.event_release(int(ev))
    finally:
        for h in hptrs:
            ctx.host_free(h)
    # no-grow allocation: what the segments hold is handed out, more than that is refused without pinning
    _in_use, reserved, _bound = ctx.arena_stats(0)
    h = ctx.host_alloc_nogrow(1 << 20)
    ctx.host_free(h)
    with pytest.raises(_native.ArenaFull):
        ctx.host_alloc_nogrow(reserved + (1 << 30))
    assert ctx.arena_stats(0)[1] == reserved


def test_fs_plugin_write_streamed_follows_the_callback(tmp_path):
    """The plugin's optional `write_streamed` with a plain Python `wait_more` (what a plugin that does not write
    through libtsb200 implements against): the file is written in the order and amounts the callback releases,
    and holds exactly the buffer."""
    import asyncio

    import numpy as np

    from torchsnapshot_b200.storage_plugins.fs import FSStoragePlugin

    data = np.random.RandomState(5).randint(0, 256, size=(3 << 20) + 17).astype(np.uint8)
    released = []

    def wait_more(done):
        upto = min(done + (1 << 20) + 3, data.nbytes)  # odd-sized steps
        released.append((done, upto))
        return upto

    plugin = FSStoragePlugin(str(tmp_path))
    loop = asyncio.new_event_loop()
    try:
        loop.run_until_complete(plugin.write_streamed("a/b/file", memoryview(data), wait_more))
        loop.run_until_complete(plugin.write_streamed("empty", memoryview(data[:0]), wait_more))
        loop.run_until_complete(plugin.close())
    finally:
        loop.close()
    assert (tmp_path / "a" / "b" / "file").read_bytes() == data.tobytes()
    assert (tmp_path / "empty").read_bytes() == b""
    assert released[0] == (0, (1 << 20) + 3) and released[-1][1] == data.nbytes and len(released) == 4

    class Custom(FSStoragePlugin):
        async def write(self, write_io):
            await super().write(write_io)

    assert Custom(str(tmp_path)).write_streamed is None  # a subclass that customises write sees whole buffers there
