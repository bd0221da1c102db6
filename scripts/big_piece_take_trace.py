"""Timeline of a take of large objects (512 MiB each: what the pieces of FSDP flat parameters and embedding-table
shards are) on one GPU: when each file's bytes are in pinned memory, when its write starts and ends.
TSB200_TRACE_WRITES=1 python scripts/big_piece_take_trace.py"""
import json, os, sys, time, tempfile, shutil
os.environ.setdefault("TSB200_TRACE_WRITES", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200 import scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict
n = int(os.environ.get("PIECES", "16")); mb = int(os.environ.get("PIECE_MIB", "512"))
g = torch.Generator(device="cuda").manual_seed(0)
state = {}
for i in range(n):
    t = torch.empty(mb << 18, dtype=torch.float32, device="cuda"); t.view(torch.int32).random_(generator=g); state[f"p{i}"] = t
P = n * (mb << 20); app = {"m": StateDict(state)}
for trial in range(3):
    d = tempfile.mkdtemp(dir="/dev/shm"); torch.cuda.synchronize(); t0 = time.monotonic()
    Snapshot.take(d, app); dt = time.monotonic() - t0; shutil.rmtree(d)
    st = scheduler.last_write_stats
    rel = lambda x: round(1e3 * (x - st.begin), 1)
    print(json.dumps({"trial": trial, "take_ms": round(1e3 * dt, 1), "GBps": round(P / dt / 1e9, 1), "plan_ms": round(1e3 * (st.begin - t0), 1),
                      "staged": rel(st.staging_done), "d2h_done": rel(st.d2h_done), "writes_done": rel(st.writes_done)}))
    if trial == 2:
        for ready, start, end, nb in sorted(st.write_trace):
            print(f"  file {nb >> 20} MiB: in pinned memory {1e3 * ready:7.1f}  write {1e3 * start:7.1f} -> {1e3 * end:7.1f} ms  ({nb / (end - start) / 1e9:.2f} GB/s)")
