"""Hunts the intermittent ~0.85 s stall of the peer-to-peer restore.
    torchrun --nproc-per-node 2 scripts/p2p_stall_hunt.py        (ROUNDS=40 by default)
Every round does what a bench step does (take to tmpfs, restore, async take) and records, per rank, the
restore's wall time, the phase split of reshard.py (plan / ipc_open / read / exchange / copy / barrier /
cleanup), and the pinned / device arena state (in use, reserved, bound) before and after.  Rank 0 prints one
JSON line per round and a summary naming the phase that grew in the slow rounds.  NOT RUN YET in round 1 (the
GPU budget was spent when the stall was noticed)."""
import json, os, shutil, statistics, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import reshard
from torchsnapshot_b200.engine import get_engine
from torchsnapshot_b200.snapshot import Snapshot, StateDict

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
state = bench.build_state(f"cuda:{lr}")
outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})
app = {"ddp": StateDict(state)}
base = [tempfile.mkdtemp(dir="/dev/shm") if rank == 0 else None]; dist.broadcast_object_list(base, src=0); base = base[0]
eng = get_engine(lr)
def arenas():  # (in use, reserved) in MiB for the pinned host arena (0) and the device arena (1)
    out = {}
    for name, which in (("pinned", 0), ("device", 1)):
        try:
            in_use, reserved, _bound = eng.ctx.arena_stats(which)
            out[name] = [in_use >> 20, reserved >> 20]
        except Exception as e:  # the index convention is in include/tsb200.h
            out[name] = repr(e)
    return out
def sync(): torch.cuda.synchronize(); dist.barrier(device_ids=[lr])
records = []
for rnd in range(int(os.environ.get("ROUNDS", "40"))):
    path = os.path.join(base, f"r{rnd}")
    sync(); snap = Snapshot.take(path, app, replicated=["**"]); sync()
    before = arenas(); reshard.last_stats = None
    t0 = time.monotonic(); snap.restore({"ddp": outs}); t1 = time.monotonic(); sync()
    st = reshard.last_stats
    rec = {"round": rnd, "rank": rank, "restore_ms": round(1e3 * (t1 - t0), 1), "arenas_before": before, "arenas_after": arenas(),
           "phases": {k: round(v, 1) for k, v in st.phase_ms.items()} if st else None}
    pending = Snapshot.async_take(os.path.join(base, f"a{rnd}"), app, replicated=["**"]); pending.wait(); sync()
    allr = [None] * world; dist.all_gather_object(allr, rec)
    if rank == 0:
        records.append(allr); print(json.dumps(allr), flush=True)
        shutil.rmtree(path, ignore_errors=True); shutil.rmtree(os.path.join(base, f"a{rnd}"), ignore_errors=True)
    sync()
if rank == 0:
    worst = [max(r["restore_ms"] for r in allr) for allr in records]
    med = statistics.median(worst)
    slow = [i for i, w in enumerate(worst) if w > 1.5 * med]
    summary = {"median_ms": med, "slow_rounds": slow, "slow_ms": [worst[i] for i in slow]}
    if slow and records[0][0]["phases"]:
        keys = records[0][0]["phases"].keys()
        typical = {k: statistics.median(max(r["phases"].get(k, 0.0) for r in allr) for allr in records) for k in keys}
        summary["typical_phase_ms"] = typical
        summary["slow_phase_ms"] = [{k: max(r["phases"].get(k, 0.0) for r in records[i]) for k in keys} for i in slow]
    print(json.dumps(summary), flush=True); shutil.rmtree(base, ignore_errors=True)
dist.barrier(device_ids=[lr]); dist.destroy_process_group()
