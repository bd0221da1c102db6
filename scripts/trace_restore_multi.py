"""Phases of the replicated P2P restore under torchrun (wall clock per rank)."""
import json, os, sys, time, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import reshard, scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
os.environ["NCCL_DEBUG"] = "WARN"
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
state = bench.build_state(f"cuda:{lr}"); P = bench.payload_bytes(state)
outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})
base = [tempfile.mkdtemp(dir="/dev/shm") if rank == 0 else None]; dist.broadcast_object_list(base, src=0); base = base[0]
snap = Snapshot.take(os.path.join(base, "s"), {"ddp": StateDict(state)}, replicated=["**"])
def sync(): torch.cuda.synchronize(); dist.barrier(device_ids=[lr])
SWEEP = [(None, 8)] + [(c, t) for c in (2, 4, 6, 8, 16) for t in (2, 4, 8)] if os.environ.get("SWEEP") == "1" else []
for conc, tmax in SWEEP:
    if conc is None: os.environ.pop("TSB200_READ_CONCURRENCY", None)
    else: os.environ["TSB200_READ_CONCURRENCY"] = str(conc)
    os.environ["TSB200_IO_THREADS_MAX"] = str(tmax)
    ms = []
    for _ in range(3):
        sync(); t0 = time.monotonic(); snap.restore({"ddp": outs}); sync(); ms.append(round(1e3 * (time.monotonic() - t0), 1))
    rd = [None] * world; dist.all_gather_object(rd, round(reshard.last_stats.phase_ms["read"], 1))
    if rank == 0: print(json.dumps({"read_concurrency": conc, "threads_max": tmax, "restore_ms": ms, "best_GBps": round(P / min(ms) / 1e6, 2), "read_phase_ms_last": rd}), flush=True)
os.environ.pop("TSB200_READ_CONCURRENCY", None); os.environ.pop("TSB200_IO_THREADS_MAX", None)
for trial in range(4):
    sync(); t0 = time.monotonic()
    snap.restore({"ddp": outs})
    t1 = time.monotonic(); sync(); t2 = time.monotonic()
    st = reshard.last_stats
    rec = {"trial": trial, "rank": rank, "restore_ms": round(1e3 * (t1 - t0), 1), "with_barrier_ms": round(1e3 * (t2 - t0), 1),
           "phases": {k: round(v, 1) for k, v in st.phase_ms.items()}, "read_GB": round(st.read_bytes / 1e9, 2),
           "peer_GB": round(st.peer_bytes / 1e9, 2), "copy_kernel_ms": round(st.copy_ms, 2)}
    allr = [None] * world; dist.all_gather_object(allr, rec)
    if rank == 0:
        for r in allr: print(json.dumps(r), flush=True)
if rank == 0: shutil.rmtree(base, ignore_errors=True)
dist.barrier(device_ids=[lr]); dist.destroy_process_group()
