"""Storage-read roofline at the bench's scale: 75 files x 124 MiB on tmpfs -> pinned buffers, K concurrent readers."""
import json, os, sys, time, tempfile, shutil, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TSB200_PINNED_BYTES", str(12 << 30))
import torch
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200 import _native
from torchsnapshot_b200.engine import get_engine
eng = get_engine(0); ctx = eng.ctx
M = 124 << 20; NF = 75
root = tempfile.mkdtemp(dir="/dev/shm")
bufs = [ctx.host_alloc(M) for _ in range(NF)]
for i in range(NF): _native.file_write(os.path.join(root, f"f{i}"), bufs[i], M, 4)
out = {}
for K, thr in ((16, 1), (32, 1), (16, 2), (8, 4), (64, 1)):
    idx = iter(range(NF)); lock = threading.Lock()
    def worker():
        while True:
            with lock:
                i = next(idx, None)
            if i is None: return
            _native.file_read(os.path.join(root, f"f{i}"), 0, bufs[i], M, thr)
    th = [threading.Thread(target=worker) for _ in range(K)]
    t0 = time.monotonic()
    for t in th: t.start()
    for t in th: t.join()
    out[f"read_{K}x{thr}_gbps"] = round(NF * M / (time.monotonic() - t0) / 1e9, 2)
# same, while an H2D stream is running at full speed (what restore does)
d = ctx.dev_alloc(1 << 30); h = ctx.host_alloc(1 << 30)
stop = False
def h2d_loop():
    n = 0
    while not stop:
        ctx.memcpy_h2d(2, d, h, 1 << 30); ctx.stream_sync(2); n += 1
    out["h2d_during_read_gbps"] = round(n * (1 << 30) / (time.monotonic() - t_h) / 1e9, 2)
idx = iter(range(NF)); lock = threading.Lock()
def worker2():
    while True:
        with lock:
            i = next(idx, None)
        if i is None: return
        _native.file_read(os.path.join(root, f"f{i}"), 0, bufs[i], M, 1)
t_h = time.monotonic(); ht = threading.Thread(target=h2d_loop); ht.start()
th = [threading.Thread(target=worker2) for _ in range(32)]
t0 = time.monotonic()
for t in th: t.start()
for t in th: t.join()
out["read_32x1_with_concurrent_h2d_gbps"] = round(NF * M / (time.monotonic() - t0) / 1e9, 2)
stop = True; ht.join()
shutil.rmtree(root)
print(json.dumps(out))
