"""The REFERENCE's own test files, run against THIS build (only where /root/reference is mounted: the build
container; skipped on the GPU box).

`tests/_ref_suite_runner.py` binds the module name `torchsnapshot` to this repository's package (the
reference's package directory is only a fallback for modules outside the hot path, e.g. `test_utils`) and
runs the reference's test files where they lie -- nothing is copied.  Passing means the reference's
expectations for chunking, sharding, slab batching, partitioning, (de)serialisation, manifests,
take / async_take / restore / read_object and resharding hold for this build's implementation of them,
through the reference's own import paths and call signatures.

The default set is sized for the CPU suite; TSB200_REFERENCE_SUITE=full adds the slow multi-process files
(a full run is kept in profiles/r01_reference_suite.txt).
"""

import json
import os
import re

import pytest

import _ref_suite_runner as runner

REF_TESTS = os.path.join(runner.REF, "tests")
pytestmark = pytest.mark.skipif(not os.path.isdir(REF_TESTS), reason="reference not mounted")

# file -> number of tests the reference defines there that must pass (skips are the reference's own)
FAST = {
    "test_batcher.py": 120,
    "test_tensor_io_preparer.py": 5,
    "test_serialization.py": 47,
    "test_manifest.py": 20,
    "test_flatten.py": 10,
    "test_fs_storage_plugin.py": 3,
    "test_snapshot.py": 13,
    "test_sharded_tensor_resharding.py": 25,
    "test_rng_state.py": 1,
}
SLOW = {  # multi-process (torch.distributed.elastic launches) or multi-GiB tensors: up to minutes each
    "test_chunked_tensor_io_preparer.py": 8,
    "test_sharded_tensor_io_preparer.py": 13,
    "test_read_object.py": 3,
    "test_async_take.py": 2,
    "test_partitioner.py": 30,
    "test_replication_glob.py": 6,
    "test_ddp.py": 8,
    "test_ddp_replication_glob.py": 3,
    "test_ddp_infer_replication.py": 5,
    "test_pg_wrapper.py": 2,
}
FULL = os.environ.get("TSB200_REFERENCE_SUITE") == "full"

# modules that make up the hot path and its host-side mirror: these must be THIS build's files
OURS = ["batcher", "io_preparer", "io_types", "scheduler", "snapshot", "manifest", "manifest_ops", "partitioner",
        "serialization", "flatten", "knobs", "pg_wrapper", "storage_plugin", "storage_plugins.fs", "state_dict",
        "stateful", "rng_state", "io_preparers.tensor", "io_preparers.chunked_tensor", "io_preparers.sharded_tensor"]


def _run(fname, expected, tmp_path):
    modules_json = str(tmp_path / "modules.json")
    res = runner.run([os.path.join(REF_TESTS, fname)], timeout=3000, modules_json=modules_json)
    tail = "\n".join((res.stdout + res.stderr).splitlines()[-40:])
    assert res.returncode == 0, f"{fname} failed against this build:\n{tail}"
    m = re.search(r"(\d+) passed", res.stdout)
    assert m and int(m.group(1)) == expected, f"{fname}: expected {expected} passed\n{tail}"
    loaded = json.load(open(modules_json))
    for name in OURS:
        path = loaded.get(f"torchsnapshot.{name}")
        if path is not None:  # not every test file imports every module
            assert path.startswith(runner.PKG), f"torchsnapshot.{name} came from {path}"
    assert loaded["torchsnapshot"].startswith(os.path.dirname(modules_json)) or loaded["torchsnapshot"]
    return loaded


@pytest.mark.parametrize("fname", sorted(FAST))
def test_reference_test_file_passes_against_this_build(fname, tmp_path):
    loaded = _run(fname, FAST[fname], tmp_path)
    if fname == "test_batcher.py":
        assert loaded["torchsnapshot.batcher"].startswith(runner.PKG)
        assert loaded["torchsnapshot.io_preparer"].startswith(runner.PKG)
    if fname == "test_snapshot.py":
        assert loaded["torchsnapshot.snapshot"].startswith(runner.PKG)
        assert loaded["torchsnapshot.scheduler"].startswith(runner.PKG)


@pytest.mark.skipif(not FULL, reason="slow multi-process files: set TSB200_REFERENCE_SUITE=full")
@pytest.mark.parametrize("fname", sorted(SLOW))
def test_reference_multiprocess_test_file_passes_against_this_build(fname, tmp_path):
    _run(fname, SLOW[fname], tmp_path)
