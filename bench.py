#!/usr/bin/env python3
"""bench.py -- take / restore throughput of the B200 checkpoint hot path.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1], SURVEY 8d "C2 ddp"): a GPT-2-XL-like list of bf16 CUDA parameters
(~1.56 B) plus Adam state (exp_avg, exp_avg_sq bf16, `step` fp32 0-d on CUDA), all replicated
(`replicated=["**"]`), so the N ranks partition the writes; synthetic random values, random-init.
Payload P = 9.3 GB for the whole job whatever N is (strong scaling).

One step =
  (a) device-timed take: plan -> ONE gather-pack launch -> D2H into pinned host memory with a null
      storage sink; timed with CUDA events from the first pack launch to the last D2H completion.
      `value` = P / max-over-ranks of that time.
  (b) end-to-end take through the public API (Snapshot.take -> FSStoragePlugin on tmpfs), wall clock,
      D2H copies inside the timed region.  `e2e.value` = P / that.
  (c) end-to-end restore (Snapshot.restore from that snapshot into CUDA tensors), wall clock, H2D inside.
  (d) async_take blocked milliseconds (time until Snapshot.async_take returns).
Inputs (9.3 GB) are far larger than the 126 MB L2, so no L2 flush is needed between iterations.
"""

import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "snapshot_take_GBps (device-timed, whole box)"
UNIT = "GB/s"
WORKLOAD = "ddp_gpt2xl_bf16_adam_replicated_9.3GB"


def param_shapes():
    """GPT-2-XL-like parameter list."""
    shapes = []
    for _ in range(48):
        shapes += [[4800, 1600], [4800], [1600, 1600], [1600], [6400, 1600], [6400], [1600, 6400], [1600],
                   [1600], [1600], [1600], [1600]]
    shapes += [[50258, 1600], [1024, 1600]]
    return shapes


def build_state(device, scale=1.0):
    """{name: tensor}: params + Adam state.  `scale` < 1 keeps only the first fraction of the layers
    (used for the bounded CPU-baseline sample)."""
    import torch

    shapes = param_shapes()
    if scale < 1.0:
        shapes = shapes[: max(int(len(shapes) * scale), 12)]
    g = torch.Generator(device=device).manual_seed(1234)
    state = {}
    for i, shp in enumerate(shapes):
        for kind in ("param", "exp_avg", "exp_avg_sq"):
            t = torch.empty(shp, dtype=torch.bfloat16, device=device)
            # random bytes defeat any compression / dedup in the sink; values are irrelevant to a copy
            t.view(torch.int16).random_(-30000, 30000, generator=g)
            state[f"{kind}/{i}"] = t
        state[f"step/{i}"] = torch.tensor(float(i), dtype=torch.float32, device=device)
    return state


def payload_bytes(state):
    return sum(t.numel() * t.element_size() for t in state.values())


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                    capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def tmp_root():
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()
    return base


# --------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference's CPU path (oracle/ref_path.py), bounded sample
# --------------------------------------------------------------------------------------------------
def run_cpu_port(sample_scale, steps, warmup):
    import torch
    from oracle import ref_path

    state = build_state("cpu", scale=sample_scale)
    nbytes = payload_bytes(state)
    outs = {k: torch.empty_like(v) for k, v in state.items()}
    take_s, restore_s = [], []
    for i in range(warmup + steps):
        root = tempfile.mkdtemp(prefix="tsb_ref_", dir=tmp_root())
        try:
            t0 = time.monotonic()
            ref_path.take(root, state)
            t1 = time.monotonic()
            ref_path.restore(root, outs)
            t2 = time.monotonic()
        finally:
            shutil.rmtree(root, ignore_errors=True)
        if i >= warmup:
            take_s.append(t1 - t0)
            restore_s.append(t2 - t1)
    return {
        "take_gbps": nbytes * len(take_s) / sum(take_s) / 1e9,
        "restore_gbps": nbytes * len(restore_s) / sum(restore_s) / 1e9,
        "ms_per_step": 1e3 * sum(take_s) / len(take_s),
        "sample_bytes": nbytes,
        "cores": ref_path.THREADS_USED,
        "host_cores": os.cpu_count(),
        "n_tensors": len(state),
    }


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    scale = 0.25  # ~2.3 GB of the 9.3 GB state: a take + restore is ~10 s of CPU work
    r = run_cpu_port(scale, args.steps, args.warmup)
    sample = (f"first {int(scale * 100)}% of the layers of the same state dict on CPU tensors "
              f"({r['sample_bytes'] / 1e9:.2f} GB, {r['n_tensors']} tensors), take+restore per step, tmpfs")
    line = {
        "impl": "reference",
        "metric": METRIC, "value": round(r["take_gbps"], 4), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(r["ms_per_step"], 3), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "storage": tmp_root(), "sample": sample},
        "cpu_baseline": {"value": round(r["take_gbps"], 4), "unit": UNIT, "cores": r["cores"], "kind": "port",
                         "sample": sample, "restore_value": round(r["restore_gbps"], 4),
                         "host_cores": r["host_cores"]},
        "e2e": {"value": round(r["take_gbps"], 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "restore": {"value": round(r["restore_gbps"], 4), "unit": UNIT},
        "note": "wall clock (no device); the reference is pure Python, timed here as the CPU port oracle/ref_path.py",
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    import tsb200_loader

    tsb200_loader.load()
    from torchsnapshot_b200 import reshard, scheduler
    from torchsnapshot_b200.batcher import batch_write_requests
    from torchsnapshot_b200.engine import get_engine
    from torchsnapshot_b200.io_preparer import prepare_write
    from torchsnapshot_b200.io_types import StoragePlugin
    from torchsnapshot_b200.partitioner import partition_write_reqs
    from torchsnapshot_b200.pg_wrapper import PGWrapper
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    # stdout carries exactly one JSON line: keep NCCL's version banner out of it
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL prints its version banner to fd 1 when the first
        # communicator comes up, so fd 1 points at stderr until that has happened
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier(device_ids=[local_rank])
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    pg = PGWrapper(None)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    class NullPlugin(StoragePlugin):
        async def write(self, write_io):
            return None

        async def read(self, read_io):
            raise NotImplementedError

        async def delete(self, path):
            return None

        async def close(self):
            return None

    state = build_state(f"cuda:{local_rank}")
    P = payload_bytes(state)  # whole-job payload: replicated state is written once
    eng = get_engine(local_rank)
    app_state = {"ddp": StateDict(state)}
    outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})

    # shared snapshot directory (rank 0 picks it)
    base = [tempfile.mkdtemp(prefix="tsb_bench_", dir=tmp_root()) if rank == 0 else None]
    if world > 1:
        dist.broadcast_object_list(base, src=0)
    base = base[0]
    import atexit

    if rank == 0:
        atexit.register(shutil.rmtree, base, ignore_errors=True)  # never leave gigabytes in tmpfs behind

    import asyncio

    def device_timed_take():
        """(a): planner + pack + D2H with a null sink; returns (device ms, kernel ms, kernel bytes, my bytes)."""
        entries, wrs_by_path = {}, {}
        for path, t in state.items():
            e, w = prepare_write(t, f"ddp/{path}", rank, replicated=True)
            entries[f"ddp/{path}"] = e
            wrs_by_path[f"ddp/{path}"] = w
        entries, wrs_by_path = partition_write_reqs(entries, wrs_by_path, pg)
        wrs = [wr for w in wrs_by_path.values() for wr in w]
        _, wrs = batch_write_requests(list(entries.values()), wrs)
        loop = asyncio.new_event_loop()
        pending = scheduler.sync_execute_write_reqs(wrs, NullPlugin(), 32 << 30, rank, loop)
        pending.sync_complete(loop)
        loop.close()
        st = scheduler.last_write_stats
        return st.device_ms(), st.kernel_ms(), st.kernel_bytes(), st.bytes

    res = {"dev_ms": [], "kern_ms": [], "kern_bytes": [], "take_wall": [], "restore_wall": [], "blocked_ms": [],
           "unpack_ms": [], "step_wall": [], "restore_phases": []}
    counters0 = None
    clocks = ClockSampler(local_rank)
    with clocks:
        for step in range(args.warmup + args.steps):
            timed = step >= args.warmup
            if step == args.warmup:
                counters0 = eng.ctx.counters()
            barrier()
            s0 = time.monotonic()
            # (a) device-timed take
            dev_ms, kern_ms, kern_bytes, _mine = device_timed_take()
            barrier()
            # (b) end-to-end take through the public API
            path = os.path.join(base, f"step{step}")
            c_b0 = eng.ctx.counters()
            t0 = time.monotonic()
            snap = Snapshot.take(path, app_state, replicated=["**"])
            barrier()
            take_wall = time.monotonic() - t0
            c_b1 = eng.ctx.counters()
            take_desc_bytes = c_b1["desc_upload_bytes"] - c_b0["desc_upload_bytes"]  # host->device in take (b)
            take_d2h_bytes = c_b1["kernel_payload_bytes"] - c_b0["kernel_payload_bytes"]  # packed, then copied out
            # (c) end-to-end restore
            reshard.last_stats = None
            t0 = time.monotonic()
            snap.restore({"ddp": outs})
            barrier()
            restore_wall = time.monotonic() - t0
            unpack_ms = scheduler.last_read_stats.unpack_ms if scheduler.last_read_stats else 0.0
            restore_h2d = scheduler.last_read_stats.device_bytes if scheduler.last_read_stats else 0
            restore_peer = 0
            restore_phases = None
            if reshard.last_stats is not None:  # N > 1: shared read + NVLink fan-out of the replicated state
                restore_phases = {k: round(v, 1) for k, v in reshard.last_stats.phase_ms.items()}
                unpack_ms += reshard.last_stats.copy_ms
                restore_h2d += reshard.last_stats.read_bytes
                restore_peer = reshard.last_stats.peer_bytes
            if rank == 0:
                shutil.rmtree(path, ignore_errors=True)
            barrier()
            # (d) async_take blocked time
            path = os.path.join(base, f"async{step}")
            t0 = time.monotonic()
            pending = Snapshot.async_take(path, app_state, replicated=["**"])
            blocked = (time.monotonic() - t0) * 1e3
            pending.wait()
            barrier()
            if rank == 0:
                shutil.rmtree(path, ignore_errors=True)
            barrier()
            if timed:
                res["dev_ms"].append(max_over_ranks(dev_ms))
                res["kern_ms"].append(kern_ms)
                res["kern_bytes"].append(kern_bytes)
                res["take_wall"].append(max_over_ranks(take_wall))
                res["restore_wall"].append(max_over_ranks(restore_wall))
                res["blocked_ms"].append(max_over_ranks(blocked))
                res["unpack_ms"].append(unpack_ms)
                res["restore_phases"].append(restore_phases)
                res["step_wall"].append(max_over_ranks(time.monotonic() - s0))
    counters1 = eng.ctx.counters()
    # verify once, outside the timed region, that the restored bytes are the saved bytes
    def same_bytes(a, b):  # bit patterns include NaNs, so compare integers, not floats
        return torch.equal(a.reshape(-1).view(torch.uint8), b.reshape(-1).view(torch.uint8))

    ok = all(same_bytes(outs[k], state[k]) for k in list(state)[:: max(len(state) // 64, 1)])
    assert ok, "restored state differs from the saved state"

    K = len(res["dev_ms"])
    value = P * K / (sum(res["dev_ms"]) / 1e3) / 1e9
    e2e_take = P * K / sum(res["take_wall"]) / 1e9
    e2e_restore = P * K / sum(res["restore_wall"]) / 1e9
    # roofline of the dominant kernel (tsb_copy_flat_kernel): algorithmic bytes = 2 x packed payload per
    # launch / CUDA-event time around the copy kernels on the pack stream (this rank)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs, burst)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    kern_total_ms = sum(res["kern_ms"])
    achieved = (sum(res["kern_bytes"]) / (kern_total_ms / 1e3) / 1e9) if kern_total_ms > 0 else 0.0
    traffic = None
    try:
        ncu = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
        if ncu.get("workload") == WORKLOAD and ncu.get("n_gpus") == world:
            traffic = ncu.get("dram_bytes_per_launch")
    except Exception:
        pass
    launches = counters1["kernel_launches"] - counters0["kernel_launches"]

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = run_cpu_port(0.25, 1, 1)
        cpu_baseline = {
            "value": round(r["take_gbps"], 4), "unit": UNIT, "cores": r["cores"], "kind": "port",
            "sample": (f"first 25% of the layers of the same state dict on CPU tensors ({r['sample_bytes'] / 1e9:.2f} GB), "
                       "oracle/ref_path.py take then restore, tmpfs, 1 warm-up + 1 timed"),
            "restore_value": round(r["restore_gbps"], 4), "host_cores": r["host_cores"],
        }

    gpu_path_port = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the reference's own GPU take path (per-tensor D2D copy_ into a ByteTensor slab, pageable .cpu() per
        # slab, 16 writers), restated in oracle/ref_path.py: the like-for-like predecessor, whole state
        from oracle import ref_path

        times = []
        for i in range(2):
            root = tempfile.mkdtemp(prefix="tsb_refgpu_", dir=tmp_root())
            try:
                torch.cuda.synchronize()
                t0 = time.monotonic()
                ref_path.take_gpu(root, state)
                times.append(time.monotonic() - t0)
            finally:
                shutil.rmtree(root, ignore_errors=True)
        gpu_path_port = {"value": round(P / times[-1] / 1e9, 3), "unit": UNIT, "kind": "port",
                         "sample": "whole state on the GPU, oracle/ref_path.py take_gpu, tmpfs, 1 warm-up + 1 timed"}

    total_launches = sum_over_ranks(launches)
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": round(sum(res["dev_ms"]) / K, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "payload_bytes": P, "tensors": len(state),
                       "storage": f"FSStoragePlugin on {tmp_root()} (tmpfs)", "l2": "inputs (9.3 GB) >> 126 MB L2, no flush",
                       "partition": f"replicated writes partitioned over {world} rank(s)"},
            "e2e": {"value": round(e2e_take, 3), "unit": UNIT, "h2d_bytes_per_step": int(take_desc_bytes),
                    "d2h_bytes_per_step": int(take_d2h_bytes),
                    "note": ("phase (b): Snapshot.take -> FSStoragePlugin, wall clock, max over ranks; bytes are this "
                             "rank's: descriptor tables up, packed payload down (pinned host memory)")},
            "restore": {"value": round(e2e_restore, 3), "unit": UNIT,
                        "note": ("Snapshot.restore from tmpfs into CUDA tensors on every rank, wall clock; N > 1: each "
                                 "replicated byte is read from storage by one rank and fans out over NVLink"),
                        "unpack_kernel_ms": round(sum(res["unpack_ms"]) / K, 3),
                        "h2d_bytes_per_step": int(restore_h2d), "nvlink_bytes_per_rank": int(restore_peer),
                        "ms_per_step": [round(1e3 * x, 1) for x in res["restore_wall"]],
                        "rank0_p2p_phases_ms_per_step": res["restore_phases"]},
            "async_take_blocked_ms": round(sum(res["blocked_ms"]) / K, 2),
            "wall_ms_per_step": round(1e3 * sum(res["step_wall"]) / K, 1),
            "gpu_launches": int(total_launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": traffic, "kernel": "tsb_copy_flat_tma_kernel",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(res["kern_bytes"][0]) if res["kern_bytes"] else 0,
                         "kernel_ms_per_launch": round(kern_total_ms / max(K, 1), 4)},
            "clocks": clocks.summary(),
        }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        if gpu_path_port is not None:
            line["reference_gpu_path_port"] = gpu_path_port
        print(json.dumps(line), flush=True)
        shutil.rmtree(base, ignore_errors=True)
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
