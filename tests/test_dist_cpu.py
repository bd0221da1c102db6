"""The N > 1 path on CPU: world_size-2 gloo process groups (replicated-write partitioning, manifest
gather, sharded take + reshard-restore, async commit through the store)."""

import socket

import pytest
import torch.multiprocessing as mp

import _dist_workers


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(fn_name, tmp_path, world=2):
    ctx = mp.get_context("spawn")
    errq = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dist_workers.run, args=(r, world, port, fn_name, str(tmp_path), errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    for p in procs:
        if p.is_alive():
            p.kill()
            errs.append("worker timed out")
    assert not errs, "\n".join(errs)
    assert all(p.exitcode == 0 for p in procs)


@pytest.mark.timeout(300)
def test_replicated_writes_are_partitioned(tmp_path):
    _launch("replicated_partition", tmp_path)


@pytest.mark.timeout(300)
def test_sharded_take_and_reshard_restore(tmp_path):
    _launch("sharded_reshard", tmp_path)


@pytest.mark.timeout(300)
def test_async_take_commit_across_ranks(tmp_path):
    _launch("async_commit", tmp_path)


@pytest.mark.timeout(300)
@pytest.mark.parametrize("world", [2, 4])
def test_small_objects_are_gathered_in_fixed_slots(tmp_path, world):
    _launch("slot_gather", tmp_path, world=world)


@pytest.mark.timeout(300)
def test_object_collectives_of_the_host_mirror(tmp_path):
    _launch("object_collectives", tmp_path)


@pytest.mark.timeout(600)
@pytest.mark.parametrize("new_world", [1, 3])
def test_elastic_restore_at_another_world_size(tmp_path, new_world):
    """Taken by 2 ranks, restored by 1 and by 3: sharded entries are merged over the saving ranks and
    re-cut for the loading ranks; replicated entries reach ranks that did not exist at save time."""
    _launch("elastic_take", tmp_path, world=2)
    _launch("elastic_restore", tmp_path, world=new_world)


@pytest.mark.timeout(600)
@pytest.mark.parametrize("world", [2, 4])
def test_world_size_change_round_trip_inside_one_job(tmp_path, world):
    """N -> N/2 -> N with a sub-group as the smaller world (what bench.py's pre-flight runs on the GPUs)."""
    _launch("elastic_roundtrip_cpu", tmp_path, world=world)


@pytest.mark.timeout(300)
def test_replicated_globs_over_state_dicts_that_differ_between_ranks(tmp_path):
    _launch("replicated_paths_differ", tmp_path)


@pytest.mark.timeout(300)
def test_repeated_takes_of_replicated_state_repeat_the_plan(tmp_path):
    _launch("repeated_replicated_takes", tmp_path)
