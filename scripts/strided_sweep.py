"""Misaligned short runs: one-run-per-warp vs unit-by-unit, by run length and misalignment (TSB200_WARP_RUN_MIN)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200.engine import get_engine
from torchsnapshot_b200.regions import regions_to_array
eng = get_engine(0); ctx = eng.ctx
N = 1 << 30
src = torch.empty(N + 64, dtype=torch.uint8, device="cuda"); src.random_(0, 256)
dst = torch.empty(N // 2 + 64, dtype=torch.uint8, device="cuda")
out = {"warp_run_min": os.environ.get("TSB200_WARP_RUN_MIN", "default")}
for run in (64, 128, 256, 512, 1024, 4096):
    for shift in (0, 1, 2, 4, 8):
        rows = (N // 2) // run
        arr = regions_to_array([(src.data_ptr() + shift, dst.data_ptr(), run, (rows,), (2 * run,), (run,))])
        for _ in range(2): eng.copy_region_array(arr)
        ctx.stream_sync(0); ms = []
        for _ in range(5):
            a, b = eng.copy_region_array_timed(arr); ctx.event_sync(b)
            ms.append(ctx.event_elapsed_ms(a, b)); ctx.event_release(a); ctx.event_release(b)
        out[f"run{run}_shift{shift}"] = round(2 * rows * run / min(ms) / 1e6)
print(json.dumps(out), flush=True)
