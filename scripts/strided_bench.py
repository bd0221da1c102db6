"""The strided kernel alone (tsb_copy_strided_kernel) through the C-ABI: the run-length classes a reshard or a
non-contiguous view produces.  ITERS / WARMUP env: set both low for an ncu capture."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200.engine import get_engine
from torchsnapshot_b200.regions import pack_regions, regions_to_array

ITERS = int(os.environ.get("ITERS", "10")); WARMUP = int(os.environ.get("WARMUP", "3"))
eng = get_engine(0); ctx = eng.ctx
def timeit(arr, nbytes, label, traffic_factor):
    for _ in range(WARMUP): eng.copy_region_array(arr)
    ctx.stream_sync(0); ms = []
    for _ in range(ITERS):
        a, b = eng.copy_region_array_timed(arr); ctx.event_sync(b)
        ms.append(ctx.event_elapsed_ms(a, b)); ctx.event_release(a); ctx.event_release(b)
    best = min(ms)
    print(json.dumps({"case": label, "payload_bytes": nbytes, "kernel_ms_best": round(best, 4),
                      "algorithmic_GBps_rw": round(2 * nbytes / best / 1e6, 1),
                      "min_dram_traffic_GBps": round(traffic_factor * nbytes / best / 1e6, 1)}), flush=True)

N = 2 << 30
src = torch.empty(N, dtype=torch.uint8, device="cuda"); src.random_(0, 256)
dst = torch.empty(N // 2 + 64, dtype=torch.uint8, device="cuda")
def rows(run, pitch, n_rows, label, shift=0):
    arr = regions_to_array([(src.data_ptr() + shift, dst.data_ptr(), run, (n_rows,), (pitch,), (run,))])
    sectors = -(-run // 32) * 32 if pitch >= run + 32 else pitch  # bytes DRAM has to deliver per run (32-B sectors)
    timeit(arr, run * n_rows, label, 1 + sectors / run)
rows(1024, 2048, 1 << 20, "rows_1KB_of_2KB_pitch")
rows(512, 4096, 1 << 19, "rows_512B_of_4KB_pitch (column shard of a 1024-col fp32 table, 8 ranks)")
rows(64, 512, 1 << 22, "rows_64B_of_512B_pitch (column shard of a 128-col fp32 table, 8 ranks)")
rows(64, 512, 1 << 22, "rows_64B_of_512B_pitch_src_shift2", shift=2)
rows(4, 8, 1 << 26, "elem_4B_stride2 (t[:, ::2])")
# a real transpose: [8192, 8192] fp32 .t() -> contiguous payload (element-wise gather)
m = src[: 8192 * 8192 * 4].view(torch.float32).view(8192, 8192)
arr = regions_to_array(pack_regions(m.t(), dst.data_ptr()))
timeit(arr, 8192 * 8192 * 4, "transpose_8192x8192_fp32", 2)
# (r02) short runs whose source is misaligned by 1 / 4 bytes: 8- and 16-lane groups per run
for run, shift in ((64, 1), (64, 4), (128, 1), (128, 2), (256, 1), (512, 1)):
    rows(run, 8 * run, (1 << 28) // run, f"rows_{run}B_of_{8 * run}B_pitch_src_shift{shift}", shift=shift)
# (r02) transposes of other element sizes, and a batched one (the tiled transpose kernel)
for dt, es in ((torch.int16, 2), (torch.float64, 8)):
    m2 = src[: 8192 * 8192 * es].view(dt).view(8192, 8192)
    timeit(regions_to_array(pack_regions(m2.t(), dst.data_ptr())), 8192 * 8192 * es, f"transpose_8192x8192_{es}B_elements", 2)
m3 = src[: 64 * 1024 * 1024 * 4].view(torch.float32).view(64, 1024, 1024)
timeit(regions_to_array(pack_regions(m3.transpose(1, 2), dst.data_ptr())), 64 * 1024 * 1024 * 4, "transpose_batched_64x1024x1024_fp32", 2)
