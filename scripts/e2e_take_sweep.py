"""End-to-end Snapshot.take of the bench state at N ranks (torchrun) against the number of storage writes each
rank keeps in flight, with every rank's timeline: where does the wall clock go when the tmpfs ceiling of the
box is far above what the take reaches (r02, N=4: 31 GB/s against 55 GB/s)?"""
import json, os, sys, time, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1: dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
def sync():
    torch.cuda.synchronize()
    if world > 1: dist.barrier(device_ids=[lr])
def gather(x):
    if world == 1: return [x]
    out = [None] * world; dist.all_gather_object(out, x); return out
state = bench.build_state(f"cuda:{lr}"); P = bench.payload_bytes(state)
app = {"ddp": StateDict(state)}
base = [tempfile.mkdtemp(prefix="tsb_sweep_", dir="/dev/shm") if rank == 0 else None]
if world > 1: dist.broadcast_object_list(base, src=0)
base = base[0]
settings = [int(x) for x in os.environ.get("SWEEP", "0,4,8,16,32").split(",")]
for conc in settings:
    if conc:
        os.environ["TSB200_IO_CONCURRENCY"] = str(conc); scheduler._MAX_PER_RANK_IO_CONCURRENCY = conc
    else:
        os.environ.pop("TSB200_IO_CONCURRENCY", None)
    walls, traces = [], None
    for i in range(4):
        d = os.path.join(base, f"c{conc}_{i}")
        sync(); t0 = time.monotonic()
        Snapshot.take(d, app, replicated=["**"])
        t_ret = time.monotonic(); sync(); w = time.monotonic() - t0
        ws = scheduler.last_write_stats
        tr = [round(1e3 * ((x if x is not None else t0) - t0), 1) for x in (ws.begin, ws.staging_done, ws.d2h_done, ws.writes_done, t_ret)] + [round(ws.bytes / 1e6)]
        traces = gather(tr)
        wmax = max(gather(w))
        if i: walls.append(wmax)
        sync()
        if rank == 0: shutil.rmtree(d, ignore_errors=True)
        sync()
    if rank == 0:
        print(json.dumps({"writes_in_flight_per_rank": conc or f"default ({scheduler.io_concurrency()})", "n_gpus": world,
                          "take_ms": [round(1e3 * x, 1) for x in walls], "GBps_best": round(P / min(walls) / 1e9, 1),
                          "per_rank_ms[plan_done,staged,d2h_done,writes_done,returned,MB]": traces}), flush=True)
if rank == 0: shutil.rmtree(base, ignore_errors=True)
if world > 1:
    dist.barrier(device_ids=[lr]); dist.destroy_process_group()
