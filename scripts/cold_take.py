"""First take in a fresh process (arenas empty: every pinned / device segment is obtained from the driver) vs. the second."""
import os, sys, json, tempfile, shutil, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200.snapshot import Snapshot, StateDict

state = bench.build_state("cuda:0")
P = bench.payload_bytes(state)
app = {"ddp": StateDict(state)}
out = {"pinned_alloc": os.environ.get("TSB200_PINNED_ALLOC", "mmap+register"), "payload_GB": round(P / 1e9, 2), "take_ms": []}
for i in range(3):
    d = tempfile.mkdtemp(dir="/dev/shm"); torch.cuda.synchronize(); t0 = time.monotonic()
    Snapshot.take(d, app, replicated=["**"]); out["take_ms"].append(round((time.monotonic() - t0) * 1e3, 1)); shutil.rmtree(d)
print(json.dumps(out))
