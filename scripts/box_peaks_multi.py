"""Concurrent pinned D2H / H2D over all ranks (torchrun): the host-link roofline when N GPUs copy at once."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200.engine import get_engine
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
eng = get_engine(lr); ctx = eng.ctx
N = 1 << 30
h = ctx.host_alloc(N); d = ctx.dev_alloc(N)
out = {"n_gpus": world}
for name, fn, stream in (("d2h", lambda: ctx.memcpy_d2h(1, h, d, N), 1), ("h2d", lambda: ctx.memcpy_h2d(2, d, h, N), 2)):
    for _ in range(2): fn()
    ctx.stream_sync(stream); torch.cuda.synchronize(); dist.barrier(device_ids=[lr])
    a = ctx.event_record(stream)
    for _ in range(6): fn()
    b = ctx.event_record(stream); ctx.event_sync(b)
    g = 6 * N / ctx.event_elapsed_ms(a, b) / 1e6
    t = torch.tensor([g], device="cuda", dtype=torch.float64); mn = t.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN); sm = t.clone(); dist.all_reduce(sm)
    out[f"{name}_concurrent_min_per_gpu_gbps"] = round(mn.item(), 2); out[f"{name}_concurrent_sum_gbps"] = round(sm.item(), 2)
# P2P read of a peer's buffer through the copy kernel (rank r reads rank r+1)
if world > 1:
    from torchsnapshot_b200.regions import regions_to_array
    handle, off = ctx.ipc_export(d); allh = [None] * world; dist.all_gather_object(allh, (handle, off))
    peer = (rank + 1) % world; base = ctx.ipc_open(allh[peer][0]) + allh[peer][1]
    dst = ctx.dev_alloc(N)
    arr = regions_to_array([(base, dst, N, (), (), ())])
    for _ in range(2): eng.copy_region_array(arr)
    ctx.stream_sync(0); dist.barrier(device_ids=[lr])
    best = 1e9
    for _ in range(5):
        k0, k1 = eng.copy_region_array_timed(arr); ctx.event_sync(k1); best = min(best, ctx.event_elapsed_ms(k0, k1))
    t = torch.tensor([N / best / 1e6], device="cuda", dtype=torch.float64); mn = t.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    out["p2p_flat_kernel_read_min_per_gpu_gbps"] = round(mn.item(), 1)
    dist.barrier(device_ids=[lr])
if rank == 0: print(json.dumps(out), flush=True)
dist.barrier(device_ids=[lr]); dist.destroy_process_group()
