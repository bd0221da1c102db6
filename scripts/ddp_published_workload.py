"""The workload behind the ONE table the reference publishes (benchmarks/ddp/README.md:9-24, main.py:18-70):
a DDP-wrapped module of 200 fp32 CUDA parameters x 100 MB (20 GB), Snapshot.take(replicated=["**"]), wall
clock around the call.  Published on 8 x A100-40GB / local FS: 13.91 s on 1 GPU, 3.38 s on 8 GPUs, torch.save 32 s.
Here: same module, this build, tmpfs, plus torch.save of the same state dict for scale.
    python scripts/ddp_published_workload.py          # 1 GPU
    torchrun --nproc-per-node N scripts/ddp_published_workload.py
"""
import json, os, shutil, sys, tempfile, time, uuid
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200 import Snapshot

lr = int(os.environ.get("LOCAL_RANK", 0)); torch.cuda.set_device(lr)
if "RANK" not in os.environ:
    os.environ.update(RANK="0", WORLD_SIZE="1", MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
rank, world = dist.get_rank(), dist.get_world_size()

class Model(torch.nn.Module):
    def __init__(self, param_size=100_000_000, num_params=200):
        super().__init__()
        for i in range(num_params):
            self.register_parameter(f"param_{i}", torch.nn.Parameter(torch.rand(param_size // 4, device="cuda")))

model = DistributedDataParallel(Model(), gradient_as_bucket_view=True)
size = sum(p.nelement() * p.element_size() for p in model.parameters())
base = [tempfile.mkdtemp(dir="/dev/shm") if rank == 0 else None]; dist.broadcast_object_list(base, src=0); base = base[0]
out = {"workload": "benchmarks/ddp: 200 x 100 MB fp32, DDP, replicated=['**']", "GB": round(size / 1e9, 2), "n_gpus": world, "sink": "tmpfs", "take_s": []}
for trial in range(4):  # the first includes the cold arenas, like the reference's single shot includes its allocations
    path = os.path.join(base, str(uuid.uuid4())); torch.cuda.synchronize(); dist.barrier(device_ids=[lr])
    t0 = time.time()
    Snapshot.take(path=path, app_state={"model": model}, replicated=["**"])
    out["take_s"].append(round(time.time() - t0, 3)); dist.barrier(device_ids=[lr])
    if rank == 0: shutil.rmtree(path, ignore_errors=True)
if rank == 0:
    t0 = time.time()
    with open(os.path.join(base, "model.pt"), "wb+") as f: torch.save(model.state_dict(), f)
    out["torch_save_s"] = round(time.time() - t0, 3)
    out["take_GBps_warm"] = round(size / min(out["take_s"]) / 1e9, 2)
    out["published_A100_local_fs_s"] = {"torch.save": 32, "Snapshot.take 1x1": 13.91, "Snapshot.take 1x8": 3.38}
    print(json.dumps(out), flush=True); shutil.rmtree(base, ignore_errors=True)
dist.barrier(device_ids=[lr]); dist.destroy_process_group()
