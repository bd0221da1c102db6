"""Reshard-restore over NVLink: torchrun --nproc-per-node N scripts/reshard_bench.py
Saves a row-wise sharded fp32 table, then restores it (a) into a row layout shifted by half a shard
(large contiguous runs cross NVLink) and (b) into a column-wise layout (all-to-all, short runs), with
the peer-to-peer path and with the storage-only path (reference semantics)."""
import json, os, sys, time, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist
import tsb200_loader; tsb200_loader.load()
from types import SimpleNamespace as NS
import torchsnapshot_b200.io_preparer as iop
from torchsnapshot_b200 import reshard, knobs
from torchsnapshot_b200.snapshot import Snapshot, StateDict

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
GB = float(os.environ.get("TABLE_GB", "8"))
cols = int(os.environ.get("COLS", "128")); rows = int(GB * 1e9 / 4 / cols) // (2 * world) * (2 * world)

class Sharded:  # duck-typed ShardedTensor: local shards with global offsets
    def __init__(self, boxes, fill=None):
        self._shards = []
        for o, s in boxes:
            t = torch.empty(s, dtype=torch.float32, device="cuda")
            if fill is not None: t.view(torch.int32).random_(generator=fill)
            else: t.zero_()
            self._shards.append(NS(tensor=t, metadata=NS(shard_offsets=list(o), shard_sizes=list(s))))
    def local_shards(self): return self._shards
    def metadata(self): return NS(size=[rows, cols])
    def sharding_spec(self): return None
iop._is_sharded_tensor = lambda o: isinstance(o, Sharded)
import torchsnapshot_b200.snapshot as snapmod; snapmod._is_sharded_tensor = iop._is_sharded_tensor

per = rows // world
g = torch.Generator(device="cuda").manual_seed(rank)
src = Sharded([([rank * per, 0], [per, cols])], fill=g)
base = [tempfile.mkdtemp(dir="/dev/shm") if rank == 0 else None]; dist.broadcast_object_list(base, src=0); base = base[0]
snap = Snapshot.take(os.path.join(base, "s"), {"emb": StateDict(t=src)})
def sync():
    torch.cuda.synchronize(); dist.barrier(device_ids=[lr]); torch.cuda.synchronize()
def run(name, boxes, p2p):
    os.environ["TSB200_P2P_RESHARD"] = "1" if p2p else "0"
    dst = Sharded(boxes)
    best = None
    for it in range(3 if p2p else int(os.environ.get("STORAGE_ONLY_ITERS", "3"))):
        sync(); t0 = time.monotonic()
        snap.restore({"emb": StateDict(t=dst)})
        sync(); dt = time.monotonic() - t0
        best = dt if best is None else min(best, dt)
    st = reshard.last_stats if p2p else None
    res = {"case": name, "p2p": p2p, "n_gpus": world, "table_GB": round(rows * cols * 4 / 1e9, 2), "restore_s_best": round(best, 4),
           "whole_job_GBps": round(rows * cols * 4 / best / 1e9, 2)}
    if st is not None:
        t = torch.tensor([st.copy_ms, st.peer_bytes, st.local_bytes, st.read_bytes], dtype=torch.float64, device="cuda")
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX); sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        res.update({"copy_ms_max": round(mx[0].item(), 3), "peer_GB_total": round(sm[1].item() / 1e9, 3), "local_GB_total": round(sm[2].item() / 1e9, 3),
                    "storage_read_GB_total": round(sm[3].item() / 1e9, 3),
                    "nvlink_GBps_per_gpu_in": round(mx[1].item() / max(mx[0].item(), 1e-9) / 1e6, 1),
                    "rank0_phases_ms_last": {k: round(v, 1) for k, v in st.phase_ms.items()}})
    else:
        res["storage_read_GB_total"] = "every rank reads every overlapping piece"
    if rank == 0: print(json.dumps(res), flush=True)
    return dst
half = per // 2
shifted = [([(rank * per + half) % rows, 0], [per - half if rank == world - 1 else per, cols])]
if rank == world - 1: shifted.append(([0, 0], [half, cols]))
colw = cols // world
column = [([0, rank * colw], [rows, colw])]
with knobs.override_max_shard_size_bytes(512 << 20):
    for name, boxes in (("rows_shifted_half_shard", shifted), ("row_to_column", column)):
        a = run(name, boxes, True); b = run(name, boxes, False)
        for x, y in zip(a.local_shards(), b.local_shards()):
            assert torch.equal(x.tensor.view(torch.int32), y.tensor.view(torch.int32)), name
if rank == 0: shutil.rmtree(base, ignore_errors=True)
dist.barrier(device_ids=[lr]); dist.destroy_process_group()
