"""Micro-benchmark of the multi-region copy kernel (CUDA events on the pack stream)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import tsb200_loader; tsb200_loader.load()
from torchsnapshot_b200 import _native
from torchsnapshot_b200.engine import get_engine
from torchsnapshot_b200.regions import regions_to_array

print('TSB200_USE_TMA =', os.environ.get('TSB200_USE_TMA'), flush=True)
eng = get_engine(0); ctx = eng.ctx
def timeit(arr, nbytes, iters=10, label=""):
    for _ in range(3): eng.copy_region_array(arr)
    ctx.stream_sync(0)
    ms=[]; tot=[]
    import time
    for _ in range(iters):
        t0=time.perf_counter()
        a,b=eng.copy_region_array_timed(arr); ctx.event_sync(b)
        tot.append((time.perf_counter()-t0)*1e3)
        ms.append(ctx.event_elapsed_ms(a,b)); ctx.event_release(a); ctx.event_release(b)
    best=min(ms); med=sorted(ms)[len(ms)//2]
    print(json.dumps({"case":label,"bytes":nbytes,"kernel_ms_best":round(best,4),"kernel_ms_med":round(med,4),
                      "GBps_rw_best":round(2*nbytes/best/1e6,1),"GBps_rw_med":round(2*nbytes/med/1e6,1),
                      "call_to_done_ms_med":round(sorted(tot)[len(tot)//2],3)}), flush=True)

# 1. one big aligned copy 4 GiB
N = 4<<30
src = torch.empty(N, dtype=torch.uint8, device="cuda"); src.random_(0,256)
dst = torch.empty(N+64, dtype=torch.uint8, device="cuda")
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr(), N, (),(),())]), N, label="single_4GiB_aligned")
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr()+4, N-16, (),(),())]), N-16, label="single_4GiB_shift4")
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr()+1, N-16, (),(),())]), N-16, label="single_4GiB_shift1")
# torch copy for comparison
a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
for _ in range(3): dst[:N].copy_(src)
torch.cuda.synchronize(); t=[]
for _ in range(10):
    a.record(); dst[:N].copy_(src); b.record(); torch.cuda.synchronize(); t.append(a.elapsed_time(b))
print(json.dumps({"case":"torch_copy_4GiB","ms_best":min(t),"GBps_rw_best":round(2*N/min(t)/1e6,1)}), flush=True)
# 2. C3-like: many tensors 4KB-4MB log-uniform, total ~4 GiB, unaligned dst after odd sizes
rs=np.random.RandomState(0); sizes=[]; tot=0
while tot < N - (8<<20):
    n=4*int(round(2**rs.uniform(10,20))); sizes.append(n); tot+=n
so=np.concatenate([[0],np.cumsum(sizes)[:-1]])
regs=[(src.data_ptr()+int(o), dst.data_ptr()+int(o), n, (),(),()) for o,n in zip(so,sizes)]
timeit(regions_to_array(regs), tot, label=f"c3_like_{len(sizes)}_tensors_aligned")
regs=[(src.data_ptr()+int(o), dst.data_ptr()+int(o)+4, n, (),(),()) for o,n in zip(so,sizes)]
timeit(regions_to_array(regs), tot, label=f"c3_like_{len(sizes)}_tensors_dst_shift4")
regs=[(src.data_ptr()+int(o), dst.data_ptr()+int(o)+2, n, (),(),()) for o,n in zip(so,sizes)]
timeit(regions_to_array(regs), tot, label=f"c3_like_{len(sizes)}_tensors_dst_shift2")
# 3. one 128 MiB slab (per-slab launch)
M=128<<20
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr(), M, (),(),())]), M, label="single_128MiB")
# 4. strided rows: 2-D pitch
rows=1<<16; rb=16384
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr(), rb, (rows,), (2*rb,), (rb,))]), rows*rb, label="rows_16KB_pitch")
rows=1<<20; rb=1024
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr(), rb, (rows,), (2*rb,), (rb,))]), rows*rb, label="rows_1KB_pitch")
rows=1<<26; rb=4
timeit(regions_to_array([(src.data_ptr(), dst.data_ptr(), rb, (rows,), (2*rb,), (rb,))]), rows*rb, label="elem_4B_stride2")
