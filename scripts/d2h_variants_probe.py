"""Device->host (and host->device) copy variants with every rank copying at once (torchrun, N ranks).

Question (VERDICT r01 "next" 3): is the per-GPU device->host rate under contention set by how the copy is
issued -- one copy-engine stream with 128 MiB copies (what the scheduler does), several streams with small
interleaved pieces, or SMs storing straight into mapped pinned memory (the pack kernel with a host
destination: pack and D2H fused, no staging block in HBM) -- or by the platform?

Also prints the wall-clock phases of async_take (what the caller is blocked for) at this rank count.
"""
import json, os, sys, time, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200.engine import get_engine
from torchsnapshot_b200.regions import regions_to_array

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
eng = get_engine(lr); ctx = eng.ctx
G = 1 << 30
TOTAL = 2 * G
h = ctx.host_alloc(TOTAL); d = ctx.dev_alloc(TOTAL)
torch.cuda.synchronize()

def sync_all():
    for s in range(4): ctx.stream_sync(s)
    torch.cuda.synchronize()
    if world > 1: dist.barrier(device_ids=[lr])

def gather(x):
    if world == 1: return [x]
    out = [None] * world; dist.all_gather_object(out, x); return out

def ce(direction, streams, piece):
    """TOTAL bytes in `piece`-sized copies round-robin over `streams`; wall clock, all ranks at once."""
    def go():
        off = 0; i = 0
        while off < TOTAL:
            n = min(piece, TOTAL - off); s = streams[i % len(streams)]
            if direction == "d2h": ctx.memcpy_d2h(s, h + off, d + off, n)
            else: ctx.memcpy_h2d(s, d + off, h + off, n)
            off += n; i += 1
        for s in streams: ctx.stream_sync(s)
    go(); best = 1e9
    for _ in range(3):
        sync_all(); t0 = time.monotonic(); go(); best = min(best, time.monotonic() - t0)
    return TOTAL / best / 1e9

def sm(direction):
    """The flat copy kernel with the host block as destination (d2h) or source (h2d): SM-driven PCIe traffic."""
    arr = regions_to_array([(d, h, TOTAL, (), (), ())] if direction == "d2h" else [(h, d, TOTAL, (), (), ())])
    eng.copy_region_array(arr); ctx.stream_sync(0); best = 1e9
    for _ in range(3):
        sync_all(); k0, k1 = eng.copy_region_array_timed(arr); ctx.event_sync(k1)
        best = min(best, ctx.event_elapsed_ms(k0, k1)); ctx.event_release(k0); ctx.event_release(k1)
    return TOTAL / best / 1e6

res = {}
M = 1 << 20
for name, fn in (
    ("d2h_ce_1stream_128MiB", lambda: ce("d2h", [1], 128 * M)),
    ("d2h_ce_2streams_32MiB", lambda: ce("d2h", [1, 3], 32 * M)),
    ("d2h_ce_4streams_32MiB", lambda: ce("d2h", [1, 3, 0, 2], 32 * M)),
    ("d2h_ce_4streams_8MiB", lambda: ce("d2h", [1, 3, 0, 2], 8 * M)),
    ("h2d_ce_1stream_128MiB", lambda: ce("h2d", [2], 128 * M)),
    ("h2d_ce_2streams_32MiB", lambda: ce("h2d", [2, 3], 32 * M)),
):
    per = gather(round(fn(), 2))
    res[name] = {"per_gpu_gbps": per, "sum_gbps": round(sum(per), 2), "min_gbps": min(per)}
sync_all()

# ---- async_take blocked time by phase, at this rank count (bench workload, replicated) ---------------------
import bench
from torchsnapshot_b200 import snapshot as snapmod
from torchsnapshot_b200.snapshot import Snapshot, StateDict
state = bench.build_state(f"cuda:{lr}")
app = {"ddp": StateDict(state)}
base = [tempfile.mkdtemp(prefix="tsb_probe_", dir="/dev/shm") if rank == 0 else None]
if world > 1: dist.broadcast_object_list(base, src=0)
base = base[0]
blocked = []
for i in range(5):
    sync_all(); t0 = time.monotonic()
    p = Snapshot.async_take(os.path.join(base, f"a{i}"), app, replicated=["**"])
    b = 1e3 * (time.monotonic() - t0)
    phases = {k: round(v, 2) for k, v in snapmod.last_take_phases.items()}
    p.wait(); sync_all()
    if rank == 0: shutil.rmtree(os.path.join(base, f"a{i}"), ignore_errors=True)
    blocked.append({"blocked_ms": round(b, 2), "phases_ms": phases})
res["async_take_blocked_by_rank_last3"] = gather(blocked[-3:])
if rank == 0:
    shutil.rmtree(base, ignore_errors=True)
    res["n_gpus"] = world; res["host_cores"] = os.cpu_count()
    print(json.dumps(res, indent=1), flush=True)  # printed before the SM variant: a fault there must not lose this
if os.environ.get("SM_COPY", "1") == "1":
    sm_res = {}
    for name, fn in (("d2h_sm_stores_to_mapped_host", lambda: sm("d2h")), ("h2d_sm_loads_from_mapped_host", lambda: sm("h2d"))):
        per = gather(round(fn(), 2))
        sm_res[name] = {"per_gpu_gbps": per, "sum_gbps": round(sum(per), 2), "min_gbps": min(per)}
    if rank == 0: print(json.dumps(sm_res, indent=1), flush=True)
sync_all()
if world > 1:
    dist.barrier(device_ids=[lr]); dist.destroy_process_group()
