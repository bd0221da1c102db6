"""Where the blocked time of Snapshot.async_take goes for the many-small-tensors state (bench.py's C3 leg:
13 312 tensors / 8 GB on one rank): wall-clock phases of unprofiled calls, then one call under cProfile by own
time and one by cumulative time, then the same for the planning part of restore."""
import cProfile, io, json, os, pstats, shutil, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200 import scheduler, snapshot as snapmod
from torchsnapshot_b200.snapshot import Snapshot, StateDict

state = bench.build_c3("cuda:0", int(float(os.environ.get("C3_BYTES", 8e9))), seed=1234)
app = {"c3": StateDict(state)}
outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})
print("tensors", len(state), "GB", round(bench.payload_bytes(state) / 1e9, 2), flush=True)


def blocked(profile=None):
    d = tempfile.mkdtemp(dir="/dev/shm"); torch.cuda.synchronize()
    t0 = time.monotonic()
    if profile: profile.enable()
    p = Snapshot.async_take(d, app)
    if profile: profile.disable()
    b = 1e3 * (time.monotonic() - t0)
    phases = {k: round(v, 1) for k, v in snapmod.last_take_phases.items()}
    st = scheduler.last_write_stats
    phases["stage:begin,pack_launched,staged,returned (ms since the call)"] = [
        round(1e3 * (x - t0), 1) for x in (st.begin, st.pack_launched, st.staging_done, t0 + b / 1e3)]
    snap = p.wait()
    return b, phases, snap, d


def show(prof, sort, n=28, chars=4200):
    s = io.StringIO(); pstats.Stats(prof, stream=s).sort_stats(sort).print_stats(n)
    text = s.getvalue(); print(text[text.find("ncalls"):][:chars], flush=True)


for i in range(4):
    b, phases, snap, d = blocked(); shutil.rmtree(d)
    print(json.dumps({"blocked_ms": round(b, 1), "phases": phases}), flush=True)
for sort in ("tottime", "cumulative"):
    prof = cProfile.Profile(); b, phases, snap, d = blocked(prof); shutil.rmtree(d)
    print(f"--- async_take under cProfile ({sort}), {b:.1f} ms"); show(prof, sort)

b, phases, snap, d = blocked()
for i in range(3):
    torch.cuda.synchronize(); t0 = time.monotonic(); snap.restore({"c3": outs}); torch.cuda.synchronize()
    rs = scheduler.last_read_stats
    print(json.dumps({"restore_ms": round(1e3 * (time.monotonic() - t0), 1), "read_begin_ms": round(1e3 * (rs.begin - t0), 1)}), flush=True)
for sort in ("tottime", "cumulative"):
    prof = cProfile.Profile(); prof.enable(); snap.restore({"c3": outs}); prof.disable()
    print(f"--- restore under cProfile ({sort})"); show(prof, sort)
# what a job that starts up does: a Snapshot object made from the path (the metadata file is parsed), one restore
for i in range(3):
    torch.cuda.synchronize(); t0 = time.monotonic(); fresh = Snapshot(d); fresh.metadata; t1 = time.monotonic()
    fresh.restore({"c3": outs}); torch.cuda.synchronize()
    rs = scheduler.last_read_stats
    print(json.dumps({"fresh_snapshot_restore_ms": round(1e3 * (time.monotonic() - t0), 1), "metadata_ms": round(1e3 * (t1 - t0), 1),
                      "read_begin_ms": round(1e3 * (rs.begin - t0), 1)}), flush=True)
prof = cProfile.Profile(); prof.enable(); fresh = Snapshot(d); fresh.restore({"c3": outs}); prof.disable()
print("--- fresh Snapshot(path).restore under cProfile (tottime)"); show(prof, "tottime")
# a take that cannot reuse the previous plan (the tensors moved): plan caches dropped first
shutil.rmtree(d)
for i in range(3):
    snapmod.clear_plan_caches()
    b, phases, snap, d = blocked(); shutil.rmtree(d)
    print(json.dumps({"blocked_ms_planning_afresh": round(b, 1), "phases": phases}), flush=True)
