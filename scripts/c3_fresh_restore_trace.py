"""C3 (13 312 tensors / 8 GB): take into a new directory, then restore through the Snapshot the take returned --
what bench.py's C3 leg times -- against restoring the same snapshot again and against a Snapshot made from the
path.  Per restore: wall ms, when the first read was issued, and the per-read timeline (TSB200_TRACE_READS=1):
is a first read of freshly written tmpfs pages slower than a re-read, or is the difference host planning?"""
import json, os, shutil, statistics, sys, tempfile, time
os.environ["TSB200_TRACE_READS"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict

state = bench.build_c3("cuda:0", int(8e9), seed=1234)
app = {"c3": StateDict(state)}
outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})


def restore(snap, tag):
    torch.cuda.synchronize(); t0 = time.monotonic(); snap.restore({"c3": outs}); torch.cuda.synchronize()
    wall = 1e3 * (time.monotonic() - t0); rs = scheduler.last_read_stats
    tr = sorted(rs.trace, key=lambda x: x[1])
    read_ms = [1e3 * (r - p) for q, s, p, r, e, n in tr]; dev_ms = [1e3 * (e - r) for q, s, p, r, e, n in tr]
    print(json.dumps({"what": tag, "restore_ms": round(wall, 1), "first_read_issued_ms": round(1e3 * (rs.begin - t0), 1),
                      "reads": len(tr), "last_read_done_ms": round(1e3 * (rs.begin - t0) + 1e3 * max(r for q, s, p, r, e, n in tr), 1),
                      "last_device_phase_done_ms": round(1e3 * (rs.begin - t0) + 1e3 * max(e for q, s, p, r, e, n in tr), 1),
                      "read_ms_median": round(statistics.median(read_ms), 1), "read_ms_max": round(max(read_ms), 1),
                      "device_phase_ms_median": round(statistics.median(dev_ms), 1)}), flush=True)


for it in range(4):
    d = tempfile.mkdtemp(dir="/dev/shm")
    t0 = time.monotonic(); snap = Snapshot.take(d, app); torch.cuda.synchronize()
    print(json.dumps({"take_ms": round(1e3 * (time.monotonic() - t0), 1)}), flush=True)
    restore(snap, "the Snapshot the take returned, first restore of these files")
    restore(snap, "same Snapshot object again")
    restore(Snapshot(d), "Snapshot(path): metadata parsed, first restore of the object")
    shutil.rmtree(d)
