"""Box peaks the end-to-end numbers are judged against: pinned D2H/H2D GB/s and
tmpfs write/read GB/s through the same syscall path the FS plugin uses."""
import json, os, sys, time, tempfile, shutil, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200 import _native
from torchsnapshot_b200.engine import get_engine
eng = get_engine(0); ctx = eng.ctx
out = {"host_cores": os.cpu_count()}
N = 1 << 30
h = ctx.host_alloc(N); d = ctx.dev_alloc(N)
for name, fn, stream in (("d2h_pinned_gbps", lambda: ctx.memcpy_d2h(1, h, d, N), 1), ("h2d_pinned_gbps", lambda: ctx.memcpy_h2d(2, d, h, N), 2)):
    for _ in range(2): fn()
    ctx.stream_sync(stream)
    a = ctx.event_record(stream)
    for _ in range(8): fn()
    b = ctx.event_record(stream); ctx.event_sync(b)
    out[name] = round(8 * N / ctx.event_elapsed_ms(a, b) / 1e6, 2)
# both directions at once
a1 = ctx.event_record(1); a2 = ctx.event_record(2)
for _ in range(8):
    ctx.memcpy_d2h(1, h, d, N // 2); ctx.memcpy_h2d(2, d + N // 2, h + N // 2, N // 2)
b1 = ctx.event_record(1); b2 = ctx.event_record(2); ctx.event_sync(b1); ctx.event_sync(b2)
out["bidir_d2h_gbps"] = round(8 * (N // 2) / ctx.event_elapsed_ms(a1, b1) / 1e6, 2)
out["bidir_h2d_gbps"] = round(8 * (N // 2) / ctx.event_elapsed_ms(a2, b2) / 1e6, 2)
# storage: write/read 128 MiB pinned buffers to /dev/shm with T concurrent files x K pwrite workers each
root = tempfile.mkdtemp(dir="/dev/shm")
M = 128 << 20
def run(nfiles, workers, write=True):
    errs = []
    def one(i):
        try:
            p = os.path.join(root, f"f{i}")
            if write: _native.file_write(p, h + (i % 8) * M, M, workers)
            else: _native.file_read(p, 0, h + (i % 8) * M, M, workers)
        except Exception as e: errs.append(e)
    th = [threading.Thread(target=one, args=(i,)) for i in range(nfiles)]
    t0 = time.monotonic()
    for t in th: t.start()
    for t in th: t.join()
    dt = time.monotonic() - t0
    assert not errs, errs
    return round(nfiles * M / dt / 1e9, 2)
for nfiles, workers in ((1, 1), (1, 8), (4, 4), (16, 1), (16, 4), (32, 2)):
    for f in os.listdir(root): os.remove(os.path.join(root, f))
    out[f"tmpfs_write_gbps_{nfiles}files_x{workers}"] = run(nfiles, workers, True)
    out[f"tmpfs_read_gbps_{nfiles}files_x{workers}"] = run(nfiles, workers, False)
    # overwrite of existing pages (no page allocation)
    out[f"tmpfs_rewrite_gbps_{nfiles}files_x{workers}"] = run(nfiles, workers, True)
shutil.rmtree(root)
# plain host memcpy bandwidth (pinned -> pageable) with 8 threads
import numpy as np
dst = np.empty(N, dtype=np.uint8); dst[:] = 1
t0 = time.monotonic(); _native.host_memcpy(dst.ctypes.data, h, N, 8); out["host_memcpy_8t_gbps"] = round(N / (time.monotonic() - t0) / 1e9, 2)
t0 = time.monotonic(); _native.host_memcpy(dst.ctypes.data, h, N, 1); out["host_memcpy_1t_gbps"] = round(N / (time.monotonic() - t0) / 1e9, 2)
print(json.dumps(out, indent=1))
