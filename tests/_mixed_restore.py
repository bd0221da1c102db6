"""Restore into destinations of every kind in one state dict (shared by the CPU and the GPU suite).  Test
infrastructure; the product never imports it."""
import torch

import tsb200_loader

tsb200_loader.load()
from torchsnapshot_b200 import knobs  # noqa: E402
from torchsnapshot_b200.snapshot import Snapshot, StateDict  # noqa: E402


def run(device, tmp_path):
    """Restore plans slab members over columns (snapshot._RestoreView + batcher.plan_cuda_read_columns) when the
    destination is a plain CUDA tensor of the saved dtype and shape, and entry by entry otherwise.  Destinations
    of every kind in one state dict -- matching, strided, another dtype or absent (the restored object is new), on the
    CPU, a Tensor subclass -- restore to the saved values, and the same way with batching off, where
    everything is planned entry by entry (the reference's only path, snapshot.py:694-747)."""
    g = torch.Generator(device=device).manual_seed(5)
    src = {f"t{i}": torch.randn(17 + 3 * i, 5, device=device, generator=g) for i in range(80)}
    src["ints"] = torch.arange(1000, device=device, dtype=torch.int64)
    src["half"] = torch.randn(33, 7, device=device, generator=g).to(torch.bfloat16)
    snap = Snapshot.take(str(tmp_path / "s"), {"st": StateDict(src)})

    class Sub(torch.Tensor):
        pass

    def destinations():
        out = {k: torch.zeros_like(v) for k, v in src.items()}
        out["t1"] = torch.zeros(2 * 20, 5, device=device)[::2]  # strided rows
        out["t2"] = torch.zeros(23, 5, device=device, dtype=torch.float64)  # another dtype: a new tensor comes back
        out["t3"] = torch.zeros(26, 5)  # on the CPU
        out["t4"] = None  # absent: a new tensor comes back
        out["t5"] = torch.zeros(32, 5, device=device).as_subclass(Sub)
        out["t6"] = torch.zeros(5, 35, device=device).t()  # transposed destination
        return StateDict(out)

    results = []
    for batching_off in (False, True):
        dst = destinations()
        with knobs.override_is_batching_disabled(batching_off):
            Snapshot(str(tmp_path / "s")).restore({"st": dst})
        for k, v in src.items():
            got = dst[k]
            assert got is not None and tuple(got.shape) == tuple(v.shape), k
            assert torch.equal(got.detach().cpu().to(torch.float64), v.cpu().to(torch.float64)), (k, batching_off)
        assert dst["t2"].dtype == torch.float32 and dst["t3"].device.type == "cpu" and not dst["t1"].is_contiguous()
        results.append(dst)
    # a second restore through the same Snapshot object (cached view) into fresh destinations
    dst = destinations()
    snap.restore({"st": dst})
    snap.restore({"st": dst})
    for k, v in src.items():
        assert torch.equal(dst[k].detach().cpu().to(torch.float64), v.cpu().to(torch.float64)), k
