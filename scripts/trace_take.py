"""Timeline of one Snapshot.take / restore of the bench workload (wall-clock phases)."""
import os, sys, json, tempfile, shutil, time, cProfile, pstats, io
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict
def build_c3(device, total=8e9, seed=1234):
    """BASELINE config 2: ~20k tensors of 4 KB-4 MB log-uniform, 10% bf16, 5% int64, 1% strided, 1% 0-d."""
    import numpy as np
    rs = np.random.RandomState(seed); st = {}; tot = 0; i = 0
    g = torch.Generator(device=device).manual_seed(seed)
    while tot < total:
        n = 4 * int(round(2 ** rs.uniform(10, 20))); kind = rs.uniform()
        if kind < 0.10: t = torch.empty(n // 2, dtype=torch.bfloat16, device=device); t.view(torch.int16).random_(-30000, 30000, generator=g)
        elif kind < 0.15: t = torch.empty(max(n // 8, 1), dtype=torch.int64, device=device); t.random_(generator=g)
        elif kind < 0.16:
            b = torch.empty((max(n // 512, 1), 128), dtype=torch.float32, device=device); b.view(torch.int32).random_(generator=g); t = b[:, ::2]
        elif kind < 0.17: t = torch.tensor(float(i), device=device)
        else: t = torch.empty(n // 4, dtype=torch.float32, device=device); t.view(torch.int32).random_(generator=g)
        st[f"t{i}"] = t; tot += t.numel() * t.element_size(); i += 1
    return st
if os.environ.get("WORKLOAD") == "c3": state = build_c3("cuda:0")
else: state=bench.build_state("cuda:0")
P=bench.payload_bytes(state); print("tensors", len(state), "payload GB", P/1e9)
outs=StateDict({k: torch.empty_like(v) for k,v in state.items()})
def blocked():
    d=tempfile.mkdtemp(dir="/dev/shm"); torch.cuda.synchronize(); t0=time.monotonic()
    p=Snapshot.async_take(d, app, replicated=["**"]); b=(time.monotonic()-t0)*1e3; p.wait(); shutil.rmtree(d); return round(b,1)
app={"ddp": StateDict(state)}
NT=int(os.environ.get('TRIALS','4'))
for trial in range(NT):
    d=tempfile.mkdtemp(dir="/dev/shm"); torch.cuda.synchronize()
    prof = cProfile.Profile() if (trial==3 and NT==4) else None
    t0=time.monotonic()
    if prof: prof.enable()
    snap=Snapshot.take(d, app, replicated=["**"])
    if prof: prof.disable()
    t1=time.monotonic(); st=scheduler.last_write_stats
    rel=lambda x: None if x is None else round(1e3*(x-t0),1)
    print(json.dumps({"trial":trial,"take_ms":round(1e3*(t1-t0),1),"GBps":round(P/(t1-t0)/1e9,2),"exec_begin":rel(st.begin),"pack_launched":rel(st.pack_launched),
      "staging_done":rel(st.staging_done),"d2h_done":rel(st.d2h_done),"writes_done":rel(st.writes_done),"device_ms":round(st.device_ms(),1),"kernel_ms":round(st.kernel_ms(),2)}),flush=True)
    if prof:
        s=io.StringIO(); pstats.Stats(prof,stream=s).sort_stats(os.environ.get("PROF_SORT","cumulative")).print_stats(int(os.environ.get("PROF_N","25"))); print(s.getvalue()[:int(os.environ.get("PROF_CHARS","3500"))])
    prof = cProfile.Profile() if (trial==3 and NT==4) else None
    t0=time.monotonic()
    if prof: prof.enable()
    snap.restore({"ddp": outs})
    if prof: prof.disable()
    t1=time.monotonic(); rs=scheduler.last_read_stats
    print(json.dumps({"trial":trial,"restore_ms":round(1e3*(t1-t0),1),"GBps":round(P/(t1-t0)/1e9,2),"read_begin":round(1e3*(rs.begin-t0),1),"unpack_ms":round(rs.unpack_ms,2)}),flush=True)
    if prof:
        s=io.StringIO(); pstats.Stats(prof,stream=s).sort_stats(os.environ.get("PROF_SORT","cumulative")).print_stats(int(os.environ.get("PROF_N","25"))); print(s.getvalue()[:int(os.environ.get("PROF_CHARS","3500"))])
    if rs.trace and trial==NT-1:
        tr=sorted(rs.trace, key=lambda x:x[1])
        import statistics
        rd=[(r-p)*1e3 for q,s_,p,r,e,n in tr]; dv=[(e-r)*1e3 for q,s_,p,r,e,n in tr]
        print("reads: n",len(tr),"read_ms med",round(statistics.median(rd),1),"max",round(max(rd),1),"device_phase_ms med",round(statistics.median(dv),1),"max",round(max(dv),1))
        print("first read start",round(tr[0][2]*1e3,1),"last read end",round(max(x[3] for x in tr)*1e3,1),"last done",round(max(x[4] for x in tr)*1e3,1))
        for x in tr[:3]+tr[30:33]+tr[-3:]: print("  q %.1f slot %.1f pinned %.1f read_done %.1f done %.1f MB %.0f" % tuple([1e3*v for v in x[:5]]+[x[5]/1e6]))
    shutil.rmtree(d)
    print(json.dumps({"trial":trial,"async_take_blocked_ms":blocked()}),flush=True)
