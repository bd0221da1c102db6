#!/usr/bin/env python3
"""bench.py -- take / restore throughput of the B200 checkpoint hot path.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Headline workload (BASELINE.json configs[1], SURVEY 8d "C2 ddp"): a GPT-2-XL-like list of bf16 CUDA parameters
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

Around the timed steps, untimed:
  * ceilings -- every rank copies 1 GiB pinned<->device at the same time, and all ranks together write and read
    tmpfs files through the FS plugin's own pwrite/pread path: the host-link and storage rooflines of THIS box,
    in THIS run, which `roofline.host_link / e2e_take / restore` are fractions of;
  * pre-flight (N > 1) -- the multi-GPU parity tests of tests/_gpu_workers.py on the NCCL group of the run:
    peer-to-peer reshard (both staging schemes) against the storage-only path and the source data, replicated
    take/restore, and the N -> N/2 -> N world-size change; a mismatch ends the run with a nonzero exit code;
  * suite -- short legs on the other BASELINE configs: C3 (many small tensors per rank), C4 shape (FSDP-style
    flat sharded parameters, take at N, reshard-restore at N/2), C5 shape (row-wise tables N -> N/2 -> N, a
    row -> column reshard over NVLink, read_object of a 1 M-row block under a 100 MiB budget).
The JSON line ends with the compact `suite` block; per-step arrays go to stderr.
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
REF_DIR = os.path.join(ROOT, "oracle", "_ref")  # the unmodified reference, installed by __graft_entry__.build()
CPU_SAMPLE_SCALE = 0.25  # share of the layers in the CPU arms' bounded sample


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


def build_c3(device, total, seed):
    """BASELINE configs[2]: tensors of 4 KB-4 MB log-uniform until `total` bytes; 10 % bf16, 5 %
    int64, 1 % non-contiguous (`t[:, ::2]`), 1 % 0-d scalars, the rest fp32."""
    import numpy as np
    import torch

    rs = np.random.RandomState(seed)
    g = torch.Generator(device=device).manual_seed(seed)
    st, tot, i = {}, 0, 0
    while tot < total:
        n = 4 * int(round(2 ** rs.uniform(10, 20)))
        kind = rs.uniform()
        if kind < 0.10:
            t = torch.empty(n // 2, dtype=torch.bfloat16, device=device)
            t.view(torch.int16).random_(-30000, 30000, generator=g)
        elif kind < 0.15:
            t = torch.empty(max(n // 8, 1), dtype=torch.int64, device=device)
            t.random_(generator=g)
        elif kind < 0.16:
            b = torch.empty((max(n // 512, 1), 128), dtype=torch.float32, device=device)
            b.view(torch.int32).random_(generator=g)
            t = b[:, ::2]
        elif kind < 0.17:
            t = torch.tensor(float(i), device=device)
        else:
            t = torch.empty(n // 4, dtype=torch.float32, device=device)
            t.view(torch.int32).random_(generator=g)
        st[f"t{i}"] = t
        tot += t.numel() * t.element_size()
        i += 1
    return st


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


def headline_config(world):
    """The `config` both arms print (the reference arm runs a bounded sample OF THIS workload; the sample is
    described in its `cpu_baseline.sample`)."""
    shapes = param_shapes()
    n_tensors = 4 * len(shapes)
    P = sum(3 * 2 * _prod(s) for s in shapes) + 4 * len(shapes)
    return {"workload": WORKLOAD, "payload_bytes": P, "tensors": n_tensors,
            "storage": f"FSStoragePlugin on {tmp_root()} (tmpfs)", "l2": "inputs (9.3 GB) >> 126 MB L2, no flush",
            "partition": f"replicated writes partitioned over {world} rank(s)"}


def _prod(xs):
    n = 1
    for x in xs:
        n *= x
    return n


def _median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2] if xs else 0.0


# --------------------------------------------------------------------------------------------------
# CPU arms: the unmodified reference from oracle/_ref when it is installed (kind "reference"), else the
# oracle port of its CPU path (oracle/ref_path.py, kind "port"); a bounded sample of the workload
# --------------------------------------------------------------------------------------------------
def reference_available():
    return os.path.isdir(os.path.join(REF_DIR, "torchsnapshot"))


def _import_reference():
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import logging

    import torchsnapshot  # noqa: F401  (the reference package, not this repository's)

    logging.getLogger("torchsnapshot").setLevel(logging.ERROR)
    return torchsnapshot


def run_cpu_arm(sample_scale, steps, warmup, with_restore_every=1):
    """Take (and restore) of the sample on CPU tensors, wall clock.  Returns a dict with take/restore GB/s."""
    import torch

    state = build_state("cpu", scale=sample_scale)
    nbytes = payload_bytes(state)
    outs = {k: torch.empty_like(v) for k, v in state.items()}
    take_s, restore_s = [], []
    if reference_available():
        ts = _import_reference()
        kind, cores = "reference", None
        app, app_out = {"ddp": ts.StateDict(**state)}, {"ddp": ts.StateDict(**outs)}

        def one(root, do_restore):
            t0 = time.monotonic()
            snap = ts.Snapshot.take(root, app, replicated=["**"])
            t1 = time.monotonic()
            if do_restore:
                snap.restore(app_out)
            return t1 - t0, time.monotonic() - t1
        # the reference's own parallelism per rank: 1 event-loop thread + 4 staging threads (scheduler.py:29, 253)
        # + up to 16 writes in flight on aiofiles' executor (scheduler.py:30)
        cores = 1 + 4 + 16
    else:
        from oracle import ref_path

        kind, cores = "port", ref_path.THREADS_USED

        def one(root, do_restore):
            t0 = time.monotonic()
            ref_path.take(root, state)
            t1 = time.monotonic()
            if do_restore:
                ref_path.restore(root, outs)
            return t1 - t0, time.monotonic() - t1
    for i in range(warmup + steps):
        root = tempfile.mkdtemp(prefix="tsb_ref_", dir=tmp_root())
        do_restore = i >= warmup and ((i - warmup) % with_restore_every) == 0
        try:
            t_take, t_restore = one(root, do_restore)
        finally:
            shutil.rmtree(root, ignore_errors=True)
        if i >= warmup:
            take_s.append(t_take)
            if do_restore:
                restore_s.append(t_restore)
    if restore_s:
        ok = all(torch.equal(outs[k].view(torch.int16) if outs[k].dtype == torch.bfloat16 else outs[k],
                             state[k].view(torch.int16) if state[k].dtype == torch.bfloat16 else state[k])
                 for k in list(state)[::16])
        assert ok, "the CPU arm restored different bytes"
    return {
        "kind": kind,
        "take_gbps": nbytes * len(take_s) / sum(take_s) / 1e9,
        "restore_gbps": (nbytes * len(restore_s) / sum(restore_s) / 1e9) if restore_s else None,
        "ms_per_step": 1e3 * sum(take_s) / len(take_s),
        "sample_bytes": nbytes,
        "cores": cores,
        "host_cores": os.cpu_count(),
        "n_tensors": len(state),
    }


def _sample_text(r):
    prog = ("the unmodified reference (oracle/_ref: Snapshot.take / restore, FS plugin)" if r["kind"] == "reference"
            else "oracle/ref_path.py (port of the reference's CPU path)")
    return (f"first {int(CPU_SAMPLE_SCALE * 100)}% of the layers of the same state dict on CPU tensors "
            f"({r['sample_bytes'] / 1e9:.2f} GB, {r['n_tensors']} tensors), {prog}, tmpfs")


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # restore is the slower half of the reference: time it on every fourth step so the run stays within minutes
    r = run_cpu_arm(CPU_SAMPLE_SCALE, args.steps, args.warmup, with_restore_every=4)
    sample = _sample_text(r) + "; take every step, restore every 4th"
    line = {
        "impl": "reference",
        "metric": METRIC, "value": round(r["take_gbps"], 4), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(r["ms_per_step"], 3), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": headline_config(args.gpus),
        "cpu_baseline": {"value": round(r["take_gbps"], 4), "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "sample": sample,
                         "restore_value": round(r["restore_gbps"], 4) if r["restore_gbps"] else None,
                         "host_cores": r["host_cores"]},
        "e2e": {"value": round(r["take_gbps"], 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "restore": {"value": round(r["restore_gbps"], 4) if r["restore_gbps"] else None, "unit": UNIT},
        "note": ("wall clock on the host cores, no device; rank 0 only; each step takes the bounded sample named in "
                 "cpu_baseline.sample of the workload named in config"),
    }
    print(json.dumps(line), flush=True)
    return 0


def reference_gpu_path(state, P):
    """The reference's own GPU path on this box (CUDA state dict, unmodified reference from oracle/_ref;
    the port oracle/ref_path.take_gpu when it is not installed): take GB/s and async_take blocked ms, measured
    the way benchmarks/torchrec/main.py:136-151 does."""
    import torch

    out = {}
    if reference_available():
        ts = _import_reference()
        app = {"ddp": ts.StateDict(**state)}
        times = []
        for _ in range(2):
            root = tempfile.mkdtemp(prefix="tsb_refgpu_", dir=tmp_root())
            try:
                torch.cuda.synchronize()
                t0 = time.monotonic()
                ts.Snapshot.take(root, app, replicated=["**"])
                times.append(time.monotonic() - t0)
            finally:
                shutil.rmtree(root, ignore_errors=True)
        root = tempfile.mkdtemp(prefix="tsb_refgpu_", dir=tmp_root())
        try:
            torch.cuda.synchronize()
            t0 = time.monotonic()
            pending = ts.Snapshot.async_take(root, app, replicated=["**"])
            blocked = time.monotonic() - t0
            pending.wait()
        finally:
            shutil.rmtree(root, ignore_errors=True)
        out = {"value": round(P / times[-1] / 1e9, 3), "unit": UNIT, "kind": "reference",
               "async_take_blocked_ms": round(1e3 * blocked, 1),
               "sample": "whole state on the GPU, unmodified reference (oracle/_ref), tmpfs, 1 warm-up + 1 timed take, 1 async_take"}
    else:
        from oracle import ref_path

        times = []
        for _ in range(2):
            root = tempfile.mkdtemp(prefix="tsb_refgpu_", dir=tmp_root())
            try:
                torch.cuda.synchronize()
                t0 = time.monotonic()
                ref_path.take_gpu(root, state)
                times.append(time.monotonic() - t0)
            finally:
                shutil.rmtree(root, ignore_errors=True)
        out = {"value": round(P / times[-1] / 1e9, 3), "unit": UNIT, "kind": "port",
               "sample": "whole state on the GPU, oracle/ref_path.py take_gpu, tmpfs, 1 warm-up + 1 timed"}
    return out


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
    ap.add_argument("--no-suite", action="store_true", help="skip the pre-flight and the C3/C4/C5 legs")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    import tsb200_loader

    tsb200_loader.load()
    from torchsnapshot_b200 import _native, reshard, scheduler
    from torchsnapshot_b200 import snapshot as snapmod
    from torchsnapshot_b200.batcher import batch_write_requests
    from torchsnapshot_b200.engine import get_engine
    from torchsnapshot_b200.io_preparer import prepare_write
    from torchsnapshot_b200.io_types import StoragePlugin
    from torchsnapshot_b200.partitioner import partition_write_reqs, set_speed_context
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
    half = max(world // 2, 1)
    sub = None
    if world > 1:
        # stdout carries exactly one JSON line: NCCL prints its version banner to fd 1 when the first
        # communicator comes up, so fd 1 points at stderr until that has happened
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier(device_ids=[local_rank])
            sub = dist.new_group(ranks=list(range(half)))  # the N/2 world of the reshard legs
            if rank < half:
                dist.barrier(group=sub, device_ids=[local_rank])
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    pg = PGWrapper(None)

    def log(*a):
        if rank == 0:
            print(*a, file=sys.stderr, flush=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def reduce_f(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(x):
        return reduce_f(x, dist.ReduceOp.MAX)

    def sum_over_ranks(x):
        return reduce_f(x, dist.ReduceOp.SUM)

    def gather_f(xs):
        """Every rank's list of floats -> list (per rank) of lists, on every rank."""
        if world == 1:
            return [list(xs)]
        t = torch.tensor(list(xs), dtype=torch.float64, device="cuda")
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [o.tolist() for o in out]

    def die(msg):
        """A rank that found wrong bytes ends the whole job: torchrun stops the other workers."""
        print(f"[bench] rank {rank}: {msg}", file=sys.stderr, flush=True)
        sys.stdout.flush()
        os._exit(17)

    def same_bytes(a, b):  # bit patterns include NaNs, so compare integers, not floats
        return torch.equal(a.contiguous().reshape(-1).view(torch.uint8), b.contiguous().reshape(-1).view(torch.uint8))

    class NullPlugin(StoragePlugin):
        async def write(self, write_io):
            return None

        async def read(self, read_io):
            raise NotImplementedError

        async def delete(self, path):
            return None

        async def close(self):
            return None

    # shared scratch directory (rank 0 picks it)
    base = [tempfile.mkdtemp(prefix="tsb_bench_", dir=tmp_root()) if rank == 0 else None]
    if world > 1:
        dist.broadcast_object_list(base, src=0)
    base = base[0]
    import atexit

    if rank == 0:
        atexit.register(shutil.rmtree, base, ignore_errors=True)  # never leave gigabytes in tmpfs behind

    eng = get_engine(local_rank)
    ctx = eng.ctx

    # ------------------------------------------------------------------------------------------
    # ceilings of this box, measured in this run (untimed)
    # ------------------------------------------------------------------------------------------
    def measure_ceilings():
        G = 1 << 30
        h = ctx.host_alloc(G)
        d = ctx.dev_alloc(G)
        rates = []
        for fn, stream in ((lambda: ctx.memcpy_d2h(_native.STREAM_D2H, h, d, G), _native.STREAM_D2H),
                           (lambda: ctx.memcpy_h2d(_native.STREAM_H2D, d, h, G), _native.STREAM_H2D)):
            fn()
            ctx.stream_sync(stream)
            barrier()
            a = ctx.event_record(stream)
            for _ in range(4):
                fn()
            b = ctx.event_record(stream)
            ctx.event_sync(b)
            rates.append(4 * G / ctx.event_elapsed_ms(a, b) / 1e6)
            ctx.event_release(a)
            ctx.event_release(b)
        per_rank = gather_f(rates)
        # storage: all ranks together, one writer thread per file like the FS plugin under load; every file has
        # source / destination memory of its own (writers that share a 1 GiB source read it out of the last-level
        # cache and overstate the ceiling: 55 GB/s against 39 GB/s for a take whose D2H had long finished)
        M = 128 << 20
        files = max(4, min(32, 2 * (os.cpu_count() or 8) // world))  # the writers a rank keeps busy (scheduler.io_concurrency)
        ctx.host_free(h)
        h = ctx.host_alloc(files * M)
        root = os.path.join(base, "ceil")
        os.makedirs(root, exist_ok=True)
        barrier()

        def run(write, tag):
            errs = []

            def one(i):
                try:
                    p = os.path.join(root, f"r{rank}_{i}_{tag}")
                    if write:
                        _native.file_write(p, h + i * M, M, 1)
                    else:
                        _native.file_read(p, 0, h + i * M, M, 1)
                except Exception as e:  # noqa: BLE001
                    errs.append(e)

            th = [threading.Thread(target=one, args=(i,)) for i in range(files)]
            barrier()
            t0 = time.monotonic()
            for t in th:
                t.start()
            for t in th:
                t.join()
            dt = max_over_ranks(time.monotonic() - t0)
            assert not errs, errs
            return world * files * M / dt / 1e9

        # best of three passes each (a single 0.1 s pass ranged from 29 to 43 GB/s from run to run); every write
        # pass makes new files, like a take does, and every read pass reads files that have not been read yet,
        # like a restore does (reading the same files again runs 10-25 % faster and is not what a restore sees)
        fs_stat = os.statvfs(root)
        free = -max_over_ranks(-float(fs_stat.f_bavail * fs_stat.f_frsize))  # the same answer on every rank
        passes = 3 if free > (64 << 30) else 1  # 3 x (4 to 8 GiB) of files for a moment
        w = max(run(True, k) for k in range(passes))
        r = max(run(False, k) for k in range(passes))
        barrier()
        if rank == 0:
            shutil.rmtree(root, ignore_errors=True)
        ctx.host_free(h)
        ctx.dev_free(d)
        d2h = [p[0] for p in per_rank]
        h2d = [p[1] for p in per_rank]
        return {"d2h_sum_gbps": round(sum(d2h), 1), "d2h_min_gbps": round(min(d2h), 1),
                "h2d_sum_gbps": round(sum(h2d), 1), "h2d_min_gbps": round(min(h2d), 1),
                "d2h_per_gpu_gbps": [round(x, 1) for x in d2h],
                "tmpfs_write_gbps": round(w, 1), "tmpfs_read_gbps": round(r, 1), "host_cores": os.cpu_count(),
                "how": (f"1 GiB pinned<->device x4 on every rank at once (CUDA events); {world * files} files of 128 MiB "
                        "written then read by one thread each from / into memory of its own, all ranks at once "
                        "(wall clock, max over ranks, best of 3 passes over files of their own)")}

    ceilings = measure_ceilings()
    log("[bench] ceilings", json.dumps(ceilings))

    # ------------------------------------------------------------------------------------------
    # pre-flight parity on the group of this run (N > 1, untimed)
    # ------------------------------------------------------------------------------------------
    suite = {}
    if world > 1 and not args.no_suite:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import _gpu_workers

        t0 = time.monotonic()
        for name in ("p2p_reshard", "ddp_replicated", "elastic_roundtrip_gpu"):
            d = os.path.join(base, f"pre_{name}")
            if rank == 0:
                os.makedirs(d, exist_ok=True)
            barrier()
            try:
                if name == "elastic_roundtrip_gpu":
                    _gpu_workers.elastic_roundtrip_gpu(rank, world, d, sub=sub)
                else:
                    getattr(_gpu_workers, name)(rank, world, d)
            except BaseException:  # noqa: B902
                import traceback

                die(f"pre-flight {name} FAILED\n{traceback.format_exc()}")
            barrier()
            if rank == 0:
                shutil.rmtree(d, ignore_errors=True)
        suite["parity_preflight"] = "ok"
        suite["preflight_s"] = round(time.monotonic() - t0, 1)
        log("[bench] pre-flight ok", suite["preflight_s"], "s")
    elif world == 1:
        suite["parity_preflight"] = "n/a (1 GPU)"

    # ------------------------------------------------------------------------------------------
    # headline workload
    # ------------------------------------------------------------------------------------------
    state = build_state(f"cuda:{local_rank}")
    P = payload_bytes(state)  # whole-job payload: replicated state is written once
    app_state = {"ddp": StateDict(state)}
    outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})

    import asyncio

    def device_timed_take():
        """(a): planner + pack + D2H with a null sink; returns (device ms, kernel ms, kernel bytes, my bytes)."""
        entries, wrs_by_path = {}, {}
        for path, t in state.items():
            e, w = prepare_write(t, f"ddp/{path}", rank, replicated=True)
            entries[f"ddp/{path}"] = e
            wrs_by_path[f"ddp/{path}"] = w
        set_speed_context("NullPlugin")  # this leg has a sink of its own kind: the links alone bound it
        entries, wrs_by_path = partition_write_reqs(entries, wrs_by_path, pg)
        wrs = [wr for w in wrs_by_path.values() for wr in w]
        _, wrs = batch_write_requests(list(entries.values()), wrs)
        loop = asyncio.new_event_loop()
        pending = scheduler.sync_execute_write_reqs(wrs, NullPlugin(), 32 << 30, rank, loop)
        pending.sync_complete(loop)
        loop.close()
        st = scheduler.last_write_stats
        return st.device_ms(), st.kernel_ms(), st.kernel_bytes(), st.kernel_bytes() // 2

    res = {"dev_ms": [], "kern_ms": [], "kern_bytes": [], "take_wall": [], "restore_wall": [], "blocked_ms": [],
           "unpack_ms": [], "step_wall": [], "restore_phases": [], "rank_dev_ms": [], "rank_bytes": [],
           "blocked_phases": []}
    counters0 = None
    first_blocked = 0.0
    take_desc_bytes = take_d2h_bytes = restore_h2d = restore_peer = 0
    clocks = ClockSampler(local_rank)
    with clocks:
        for step in range(args.warmup + args.steps):
            timed = step >= args.warmup
            if step == args.warmup:
                counters0 = ctx.counters()
            barrier()
            s0 = time.monotonic()
            # (a) device-timed take
            dev_ms, kern_ms, kern_bytes, mine = device_timed_take()
            barrier()
            # (b) end-to-end take through the public API
            path = os.path.join(base, f"step{step}")
            c_b0 = ctx.counters()
            t0 = time.monotonic()
            snap = Snapshot.take(path, app_state, replicated=["**"])
            t_ret = time.monotonic()
            barrier()
            take_wall = time.monotonic() - t0
            ws = scheduler.last_write_stats
            take_trace = [1e3 * ((x if x is not None else t0) - t0) for x in
                          (ws.begin, ws.staging_done, ws.d2h_done, ws.writes_done, t_ret)] + [ws.bytes / 1e6]
            c_b1 = ctx.counters()
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
                rs = reshard.last_stats
                restore_phases = {k: round(v, 1) for k, v in rs.phase_ms.items()}
                restore_phases["new_staging_mb"] = rs.new_staging_bytes >> 20
                restore_phases["opens"] = rs.opens
                unpack_ms += rs.copy_ms
                restore_h2d += rs.read_bytes
                restore_peer = rs.peer_bytes
            if rank == 0:
                shutil.rmtree(path, ignore_errors=True)
            barrier()
            # (d) async_take blocked time
            path = os.path.join(base, f"async{step}")
            t0 = time.monotonic()
            pending = Snapshot.async_take(path, app_state, replicated=["**"])
            blocked = (time.monotonic() - t0) * 1e3
            blocked_phases = {k: round(v, 1) for k, v in snapmod.last_take_phases.items()}
            if step == 0:
                first_blocked = max_over_ranks(blocked)
            pending.wait()
            barrier()
            if rank == 0:
                shutil.rmtree(path, ignore_errors=True)
            barrier()
            per_rank = gather_f([dev_ms, float(mine)])
            take_traces = gather_f(take_trace)
            if timed:
                res["dev_ms"].append(max_over_ranks(dev_ms))
                res["kern_ms"].append(kern_ms)
                res["kern_bytes"].append(kern_bytes)
                res["take_wall"].append(max_over_ranks(take_wall))
                res["restore_wall"].append(max_over_ranks(restore_wall))
                res["blocked_ms"].append(max_over_ranks(blocked))
                res["unpack_ms"].append(unpack_ms)
                res["restore_phases"].append(restore_phases)
                res["blocked_phases"].append(blocked_phases)
                res["step_wall"].append(max_over_ranks(time.monotonic() - s0))
                res["rank_dev_ms"].append([p[0] for p in per_rank])
                res["rank_bytes"].append([p[1] for p in per_rank])
    counters1 = ctx.counters()
    # verify, outside the timed region, that EVERY restored tensor holds the saved bytes
    bad = [k for k in state if not same_bytes(outs[k], state[k])]
    if bad:
        die(f"restored state differs from the saved state in {len(bad)} tensors, e.g. {bad[:3]}")

    K = len(res["dev_ms"])
    value = P * K / (sum(res["dev_ms"]) / 1e3) / 1e9
    e2e_take = P * K / sum(res["take_wall"]) / 1e9
    e2e_restore = P * K / sum(res["restore_wall"]) / 1e9
    # roofline of the dominant kernel (tsb_copy_flat_tma_kernel): algorithmic bytes = 2 x packed payload per
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
    total_launches = sum_over_ranks(launches)

    cpu_baseline = None
    gpu_path = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = run_cpu_arm(CPU_SAMPLE_SCALE, 1, 1)
        cpu_baseline = {
            "value": round(r["take_gbps"], 4), "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
            "sample": _sample_text(r) + ", take then restore, 1 warm-up + 1 timed",
            "restore_value": round(r["restore_gbps"], 4), "host_cores": r["host_cores"],
        }
        gpu_path = reference_gpu_path(state, P)

    # per-step detail goes to stderr; the line keeps medians
    log("[bench] per-step", json.dumps({
        "dev_ms": [round(x, 1) for x in res["dev_ms"]], "take_wall_ms": [round(1e3 * x, 1) for x in res["take_wall"]],
        "restore_wall_ms": [round(1e3 * x, 1) for x in res["restore_wall"]],
        "blocked_ms": [round(x, 1) for x in res["blocked_ms"]], "rank0_restore_phases": res["restore_phases"],
        "rank0_blocked_phases": res["blocked_phases"][-3:], "rank_dev_ms": res["rank_dev_ms"][-3:],
        "rank_bytes": res["rank_bytes"][-1:],
        "e2e_take_last_step_per_rank_ms[plan_done,staged,d2h_done,writes_done,returned,MB]":
            [[round(x, 1) for x in r] for r in take_traces]}))

    # ------------------------------------------------------------------------------------------
    # suite legs (untimed with respect to the headline; each leg times itself)
    # ------------------------------------------------------------------------------------------
    del app_state, outs, state, snap
    torch.cuda.empty_cache()
    if not args.no_suite:
        try:
            suite.update(run_suite(locals()))
        except BaseException:  # noqa: B902
            import traceback

            die(f"suite FAILED\n{traceback.format_exc()}")

    if rank == 0:
        rw = sorted(res["restore_wall"])
        n_r = len(res["rank_dev_ms"])
        rank_ms = [round(_median([res["rank_dev_ms"][i][r] for i in range(n_r)]), 1) for r in range(world)]
        rank_mb = [int(_median([res["rank_bytes"][i][r] for i in range(n_r)]) / 1e6) for r in range(world)]
        host_peak = ceilings["d2h_sum_gbps"]
        e2e_peak = min(ceilings["d2h_sum_gbps"], ceilings["tmpfs_write_gbps"])
        rst_peak = min(ceilings["h2d_sum_gbps"], ceilings["tmpfs_read_gbps"])
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": round(sum(res["dev_ms"]) / K, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": headline_config(world),
            "e2e": {"value": round(e2e_take, 3), "unit": UNIT, "h2d_bytes_per_step": int(take_desc_bytes),
                    "d2h_bytes_per_step": int(take_d2h_bytes), "restore_value": round(e2e_restore, 3),
                    "async_take_blocked_ms": round(sum(res["blocked_ms"]) / K, 2),
                    "note": ("Snapshot.take -> FSStoragePlugin, wall clock, max over ranks; bytes are rank 0's: "
                             "descriptor tables up, packed payload down (pinned host memory)")},
            "gpu_launches": int(total_launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": traffic, "kernel": "tsb_copy_flat_tma_kernel",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(res["kern_bytes"][0]) if res["kern_bytes"] else 0,
                         "kernel_ms_per_launch": round(kern_total_ms / max(K, 1), 4),
                         # the rooflines that bind the step, against this run's own ceilings
                         "host_link": {"achieved": round(value, 2), "peak": host_peak, "frac": round(value / host_peak, 3),
                                       "what": "value / sum over GPUs of concurrent pinned D2H"},
                         "e2e_take": {"achieved": round(e2e_take, 2), "peak": e2e_peak, "frac": round(e2e_take / e2e_peak, 3),
                                      "what": "e2e / min(D2H sum, tmpfs write)"},
                         "restore": {"achieved": round(e2e_restore, 2), "peak": rst_peak,
                                     "frac": round(e2e_restore / rst_peak, 3), "what": "restore / min(H2D sum, tmpfs read)"}},
            "clocks": clocks.summary(),
            "ceilings": ceilings,
            "per_rank": {"assigned_mb": rank_mb, "d2h_ms": rank_ms,
                         "d2h_gbps": [round(m / max(t, 1e-9), 1) for m, t in zip(rank_mb, rank_ms)]},
            "wall_ms_per_step": round(1e3 * sum(res["step_wall"]) / K, 1),
        }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        if gpu_path is not None:
            line["reference_gpu_path"] = gpu_path
        # the rest of BASELINE.json's metric, then the suite: the tail of the line
        line["restore"] = {"value": round(e2e_restore, 3), "unit": UNIT, "ms_median": round(1e3 * rw[len(rw) // 2], 1),
                           "ms_max": round(1e3 * rw[-1], 1), "unpack_kernel_ms": round(sum(res["unpack_ms"]) / K, 3),
                           "h2d_bytes_per_step": int(restore_h2d), "nvlink_bytes_per_rank": int(restore_peer)}
        line["async_take_blocked_ms"] = round(sum(res["blocked_ms"]) / K, 2)
        line["async_take_blocked_note"] = {
            "first_call_ms": round(first_blocked, 1), "median_ms": round(_median(res["blocked_ms"]), 2),
            "max_ms": round(max(res["blocked_ms"]), 1),
            "what": ("timed steps checkpoint the same tensors as the step before, as a training loop does, so the "
                     "previous plan is reused after an address/shape check; first_call_ms is the first async_take of "
                     "the process (plans afresh; arenas already grown by the step's take)")}
        line["suite"] = suite
        print(json.dumps(line), flush=True)
        shutil.rmtree(base, ignore_errors=True)
    if dist.is_initialized():
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()
    return 0


def run_suite(env):
    """The C3 / C4-shape / C5-shape legs.  `env` is main()'s namespace (process group, helpers, ceilings)."""
    import torch
    import torch.distributed as dist

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _elastic
    from torchsnapshot_b200 import reshard
    from torchsnapshot_b200 import snapshot as snapmod
    from torchsnapshot_b200.snapshot import Snapshot, StateDict

    rank, world, local_rank, half, sub = env["rank"], env["world"], env["local_rank"], env["half"], env["sub"]
    barrier, max_over_ranks, sum_over_ranks = env["barrier"], env["max_over_ranks"], env["sum_over_ranks"]
    base, ceilings, die, log, same_bytes = env["base"], env["ceilings"], env["die"], env["log"], env["same_bytes"]
    device = f"cuda:{local_rank}"
    out = {}
    fs = os.statvfs(base)
    tmpfs_free = fs.f_bavail * fs.f_frsize
    write_peak = min(ceilings["d2h_sum_gbps"], ceilings["tmpfs_write_gbps"])
    read_peak = min(ceilings["h2d_sum_gbps"], ceilings["tmpfs_read_gbps"])

    def timed(fn, group_ranks=None):
        """Wall clock of fn on the ranks that run it, max over all ranks, bracketed by world barriers."""
        barrier()
        t0 = time.monotonic()
        if group_ranks is None or rank in group_ranks:
            fn()
            torch.cuda.synchronize()
        dt = time.monotonic() - t0
        dt = max_over_ranks(dt)
        barrier()
        return dt

    def rm(path):
        barrier()
        if rank == 0:
            shutil.rmtree(path, ignore_errors=True)
        barrier()

    # ---- C3: many small tensors per rank (take, restore, async_take blocked) -----------------------------
    per_rank = int(min(8e9, 0.25 * tmpfs_free / world, 24e9 / world))
    state = build_c3(device, per_rank, seed=1234 + rank)
    n_t, nbytes = len(state), payload_bytes(state)
    total = sum_over_ranks(float(nbytes))
    app = {"c3": StateDict(state)}
    outs = StateDict({k: torch.empty_like(v) for k, v in state.items()})
    take_s, restore_s, blocked = [], [], []
    for it in range(3):
        p = os.path.join(base, f"c3_{it}")
        holder = {}
        t = timed(lambda: holder.__setitem__("snap", Snapshot.take(p, app)))
        r = timed(lambda: holder["snap"].restore({"c3": outs}))
        rm(p)
        pa = os.path.join(base, f"c3a_{it}")
        barrier()
        t0 = time.monotonic()
        pending = Snapshot.async_take(pa, app)
        b = max_over_ranks(1e3 * (time.monotonic() - t0))
        phases = {k: round(v, 1) for k, v in snapmod.last_take_phases.items()}
        pending.wait()
        rm(pa)
        if it > 0:
            take_s.append(t)
            restore_s.append(r)
            blocked.append(b)
    bad = [k for k in state if not same_bytes(outs[k], state[k])]
    if bad:
        die(f"C3: {len(bad)} restored tensors differ, e.g. {bad[:3]}")
    log("[bench] c3 blocked phases (rank 0)", json.dumps(phases))
    tk, rs = total / min(take_s) / 1e9, total / min(restore_s) / 1e9
    out["c3"] = {"tensors_per_rank": n_t, "gb_per_rank": round(nbytes / 1e9, 2), "take_gbps": round(tk, 1),
                 "restore_gbps": round(rs, 1), "blocked_ms": round(min(blocked), 1),
                 "take_frac": round(tk / write_peak, 2), "restore_frac": round(rs / read_peak, 2), "parity": "ok"}
    del state, outs, app
    torch.cuda.empty_cache()

    # ShardedTensor needs a process group even when the world is one rank (the legs above ran without one, like
    # a single-process user of the library)
    if world == 1 and not dist.is_initialized():
        import socket

        with socket.socket() as sock:
            sock.bind(("127.0.0.1", 0))
            port = sock.getsockname()[1]
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)  # NCCL's banner goes to stderr
        try:
            dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1,
                                    device_id=torch.device("cuda", local_rank))
            dist.barrier(device_ids=[local_rank])
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    # ---- C4 shape: FSDP-style flat sharded parameters; take at N, reshard-restore at N/2 -------------------
    total4 = min(16e9, 0.25 * tmpfs_free, 8e9 * world)
    units = 4
    n_el = int(total4 / units / 10) // 2 * 2 + 1  # odd: shard boundaries never line up across world sizes
    spec = {}
    for u in range(units):
        spec[f"p{u}"] = ([n_el], torch.bfloat16, 0)
        spec[f"m{u}"] = ([n_el], torch.float32, 0)
        spec[f"v{u}"] = ([n_el], torch.float32, 0)
    bytes4 = units * n_el * 10
    src = StateDict(_elastic.build(spec, None, "cuda", filled=True))
    p4 = os.path.join(base, "c4")
    t_take = min(timed(lambda: Snapshot.take(p4 + str(i), {"m": src}, replicated=["m/rep", "m/step"])) for i in range(2))
    del src
    torch.cuda.empty_cache()
    holder = {}
    if rank < half:
        holder["dst"] = StateDict(_elastic.build(spec, sub, "cuda", filled=False))
    reshard.last_stats = None
    t_rest = min(timed(lambda: Snapshot(p4 + "1", pg=sub).restore({"m": holder["dst"]}), range(half)) for _ in range(2))
    if rank < half:
        bad = _elastic.verify(holder["dst"], spec)
        if bad:
            die(f"C4: restored at {half} ranks, differs from the source: {bad}")
    st = reshard.last_stats
    out["c4_reshard"] = {"gb": round(bytes4 / 1e9, 1), "worlds": f"{world}->{half}", "take_gbps": round(bytes4 / t_take / 1e9, 1),
                         "restore_gbps": round(bytes4 / t_rest / 1e9, 1), "take_frac": round(bytes4 / t_take / 1e9 / write_peak, 2),
                         "restore_frac": round(bytes4 / t_rest / 1e9 / read_peak, 2),
                         "path": "p2p" if (st is not None and not st.fell_back and "sharded" not in st.skipped_groups) else "storage (no read amplification)",
                         "parity": "ok"}
    holder.clear()
    torch.cuda.empty_cache()
    rm(p4 + "0")
    rm(p4 + "1")

    # ---- C5 shape: row-wise tables N -> N/2 -> N, row -> column over NVLink, read_object ------------------
    total5 = min(12e9, 0.2 * tmpfs_free, 6e9 * world)
    cols = 128
    rows = max(int(total5 / 2 / (cols * 4)), 1_000_000 + 7)
    rows = rows // (2 * world) * (2 * world) + 1  # uneven last shard
    spec5 = {"t0": ([rows, cols], torch.float32, 0), "t1": ([rows, cols], torch.float32, 0)}
    bytes5 = 2 * rows * cols * 4
    pa, pb = os.path.join(base, "c5a"), os.path.join(base, "c5b")
    src = StateDict(_elastic.build(spec5, None, "cuda", filled=True))
    # one untimed take first: this state is larger per rank than anything before it, and the first take of a
    # larger state grows the pinned arena (huge-page first touch + cudaHostRegister, ~40 ms per GiB)
    timed(lambda: Snapshot.take(pa + "w", {"emb": src}, replicated=["emb/rep", "emb/step"]))
    rm(pa + "w")
    t1 = timed(lambda: Snapshot.take(pa, {"emb": src}, replicated=["emb/rep", "emb/step"]))
    del src
    torch.cuda.empty_cache()
    holder = {}
    if rank < half:
        holder["mid"] = StateDict(_elastic.build(spec5, sub, "cuda", filled=False))
    t2 = timed(lambda: Snapshot(pa, pg=sub).restore({"emb": holder["mid"]}), range(half))
    if rank < half and _elastic.verify(holder["mid"], spec5):
        die(f"C5: restored at {half} ranks, differs from the source")
    t3 = timed(lambda: Snapshot.take(pb, {"emb": holder["mid"]}, pg=sub, replicated=["emb/rep", "emb/step"]), range(half))
    holder.clear()
    torch.cuda.empty_cache()
    dst = StateDict(_elastic.build(spec5, None, "cuda", filled=False))
    t4 = timed(lambda: Snapshot(pb).restore({"emb": dst}))
    if _elastic.verify(dst, spec5):
        die(f"C5: restored at {world} ranks, differs from the source")
    del dst
    torch.cuda.empty_cache()
    c5 = {"gb": round(bytes5 / 1e9, 1), "worlds": f"{world}->{half}->{world}",
          "gbps": [round(bytes5 / t / 1e9, 1) for t in (t1, t2, t3, t4)], "order": "take,restore,take,restore"}
    if world > 1:
        # row -> column: every rank needs a slice of every piece; each piece is read once and fans out over NVLink
        spec_col = {k: (shape, dt, 1) for k, (shape, dt, _d) in spec5.items()}
        col = StateDict(_elastic.build(spec_col, None, "cuda", filled=False))
        reshard.last_stats = None
        t5 = min(timed(lambda: Snapshot(pb).restore({"emb": col})) for _ in range(2))
        if _elastic.verify(col, spec_col):
            die("C5: row -> column restore differs from the source")
        st = reshard.last_stats
        peer = st.peer_bytes if st is not None else 0
        nv = max_over_ranks(peer / max(st.copy_ms, 1e-9) / 1e6) if st is not None and st.copy_ms else 0.0
        c5["row_to_col"] = {"gbps": round(bytes5 / t5 / 1e9, 1), "frac": round(bytes5 / t5 / 1e9 / read_peak, 2),
                            "nvlink_gb": round(sum_over_ranks(float(peer)) / 1e9, 2), "nvlink_gbps_per_gpu_in": round(nv, 0),
                            "nvlink_frac_of_900": round(nv / 900.0, 2)}
        del col
        torch.cuda.empty_cache()
    # read_object of a 1 M-row block under a 100 MiB budget (benchmarks/load_tensor/main.py:24-25 in shape)
    if rank == 0:
        block = torch.zeros(1_000_000, cols, device="cuda")
        best = 1e9
        for _ in range(2):
            torch.cuda.synchronize()
            t0 = time.monotonic()
            Snapshot(pb).read_object("0/emb/t0", obj_out=block, memory_budget_bytes=100 << 20)
            torch.cuda.synchronize()
            best = min(best, time.monotonic() - t0)
        if not _elastic.box_equals(block, [rows, cols], [0, 0], [1_000_000, cols], 0):
            die("C5: read_object block differs from the source")
        c5["read_object_1Mrows_100MiB"] = {"ms": round(1e3 * best, 1), "gbps": round(block.numel() * 4 / best / 1e9, 1)}
        del block
    c5["parity"] = "ok"
    out["c5"] = c5
    rm(pa)
    rm(pb)
    return out


if __name__ == "__main__":
    sys.exit(main())
