"""async_take blocked time against the device staging bound.

"Staged = packed in HBM" needs room for a second copy of the rank's state.  With less room the take runs in
waves (each wave's staging block has to drain to the host before the next is packed), and with no room at all
the pack kernel stores straight into pinned host memory.  One process per bound (the bound is read when the
engine is created): the 9.35 GB DDP state of bench.py, 1 warm-up + 3 timed async takes and 1 take per bound.
"""
import json, os, subprocess, sys, tempfile, shutil, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def one(bound):
    import torch
    import tsb200_loader; tsb200_loader.load()
    import bench
    from torchsnapshot_b200 import scheduler
    from torchsnapshot_b200.snapshot import Snapshot, StateDict
    state = bench.build_state("cuda:0"); P = bench.payload_bytes(state)
    app = {"ddp": StateDict(state)}
    blocked, walls = [], []
    for i in range(4):
        d = tempfile.mkdtemp(dir="/dev/shm"); torch.cuda.synchronize(); t0 = time.monotonic()
        p = Snapshot.async_take(d, app, replicated=["**"]); b = time.monotonic() - t0
        p.wait(); w = time.monotonic() - t0; shutil.rmtree(d)
        if i: blocked.append(1e3 * b); walls.append(1e3 * w)
    st = scheduler.last_write_stats
    print(json.dumps({"device_staging_bound_GiB": None if bound is None else round(bound / 2**30, 3),
                      "state_GB": round(P / 1e9, 2), "waves": len(st.pack_events),
                      "packed_to_host_GB": round(st.host_packed_bytes / 1e9, 2),
                      "async_take_blocked_ms": [round(x, 1) for x in blocked],
                      "async_take_total_ms": [round(x, 1) for x in walls]}), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(None if sys.argv[1] == "none" else int(sys.argv[1]))
    else:
        for bound in ("none", 12 << 30, 6 << 30, 3 << 30, 1 << 30, 256 << 20, 64 << 20):
            env = dict(os.environ)
            if bound != "none": env["TSB200_DEVICE_STAGING_BYTES"] = str(bound)
            subprocess.run([sys.executable, __file__, str(bound)], env=env, check=False)
