"""Timeline of one Snapshot.take / restore of the bench workload (wall-clock phases)."""
import os, sys, json, tempfile, shutil, time, cProfile, pstats, io
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import scheduler
from torchsnapshot_b200.snapshot import Snapshot, StateDict
state=bench.build_state("cuda:0"); P=bench.payload_bytes(state)
outs=StateDict({k: torch.empty_like(v) for k,v in state.items()})
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
        s=io.StringIO(); pstats.Stats(prof,stream=s).sort_stats("cumulative").print_stats(25); print(s.getvalue()[:3500])
    prof = cProfile.Profile() if (trial==3 and NT==4) else None
    t0=time.monotonic()
    if prof: prof.enable()
    snap.restore({"ddp": outs})
    if prof: prof.disable()
    t1=time.monotonic(); rs=scheduler.last_read_stats
    print(json.dumps({"trial":trial,"restore_ms":round(1e3*(t1-t0),1),"GBps":round(P/(t1-t0)/1e9,2),"read_begin":round(1e3*(rs.begin-t0),1),"unpack_ms":round(rs.unpack_ms,2)}),flush=True)
    if prof:
        s=io.StringIO(); pstats.Stats(prof,stream=s).sort_stats("cumulative").print_stats(25); print(s.getvalue()[:3500])
    shutil.rmtree(d)
