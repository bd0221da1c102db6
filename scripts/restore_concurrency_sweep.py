"""Restore of the bench workload under different read-side settings: requests in flight x pread workers per request."""
import os, sys, json, tempfile, shutil, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict

state = bench.build_state("cuda:0")
P = bench.payload_bytes(state)
app = {"ddp": StateDict(state)}
d = tempfile.mkdtemp(dir="/dev/shm")
snap = Snapshot.take(d, app, replicated=["**"])
outs = {"ddp": StateDict({k: torch.empty_like(v) for k, v in state.items()})}
def one(conc, tmax):
    if conc is None: os.environ.pop("TSB200_READ_CONCURRENCY", None)
    else: os.environ["TSB200_READ_CONCURRENCY"] = str(conc)
    os.environ["TSB200_IO_THREADS_MAX"] = str(tmax)
    best = []
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.monotonic()
        Snapshot(d).restore(outs); torch.cuda.synchronize()
        best.append((time.monotonic() - t0) * 1e3)
    print(json.dumps({"read_concurrency": conc, "threads_max": tmax, "ms": [round(x, 1) for x in best], "best_GBps": round(P / min(best) / 1e6, 2)}), flush=True)
one(None, 8)
for conc in (2, 3, 4, 6, 8, 12, 16):
    for tmax in (4, 8, 16):
        one(conc, tmax)
ok = all(torch.equal(outs["ddp"][k].reshape(-1).view(torch.uint8), state[k].reshape(-1).view(torch.uint8)) for k in state)
print("restored state identical:", ok)
shutil.rmtree(d)
