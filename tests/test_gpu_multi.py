"""Multi-GPU path: one process per GPU, NCCL for metadata only, CUDA IPC + NVLink for payload.
Skipped on a box with a single GPU (run with `gpurun --gpus 2`)."""

import socket

import pytest
import torch
import torch.multiprocessing as mp

import _gpu_workers

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(fn_name, tmp_path, world=2):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    errq = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gpu_workers.run, args=(r, world, port, fn_name, str(tmp_path), errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    for p in procs:
        if p.is_alive():
            p.kill()
            errs.append("worker timed out")
    assert not errs, "\n".join(errs)
    assert all(p.exitcode == 0 for p in procs)


@pytest.mark.timeout(400)
def test_p2p_reshard_restore_two_gpus(tmp_path):
    _launch("p2p_reshard", tmp_path)


@pytest.mark.timeout(400)
def test_replicated_take_restore_two_gpus(tmp_path):
    _launch("ddp_replicated", tmp_path)


@pytest.mark.timeout(600)
@pytest.mark.parametrize("save_world,load_world", [(2, 1), (1, 2)])
def test_elastic_reshard_between_world_sizes(tmp_path, save_world, load_world):
    """Saved by `save_world` GPUs, restored by `load_world`: the flat sharded parameter is re-cut (through
    the peer-to-peer path when more than one rank loads), replicated state reaches new ranks."""
    _launch("elastic_take", tmp_path, world=save_world)
    _launch("elastic_restore", tmp_path, world=load_world)


@pytest.mark.timeout(600)
def test_world_size_change_round_trip_inside_one_job(tmp_path):
    """2 -> 1 -> 2 with a sub-group as the smaller world (bench.py's pre-flight runs the same at N = 2, 4, 8)."""
    _launch("elastic_roundtrip_gpu", tmp_path)


@pytest.mark.timeout(400)
def test_ranks_on_different_hosts_restore_through_storage(tmp_path):
    _launch("ranks_on_different_hosts", tmp_path)
