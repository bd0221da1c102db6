"""Drop-in check against the REAL reference (only where /root/reference is mounted: the build container;
skipped on the GPU box).  The reference's Snapshot.take / restore run with this build's batcher and
scheduler patched in at the seam INTEGRATION.md names, on CPU tensors, and must produce what the
unpatched reference produces."""

import hashlib
import json
import os
import sys

import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "torchsnapshot")), reason="reference not mounted")


def _payloads(root):
    meta = json.load(open(os.path.join(root, ".snapshot_metadata")))
    out = {}
    for path, e in meta["manifest"].items():
        if e["type"] != "Tensor":
            continue
        buf = open(os.path.join(root, e["location"]), "rb").read()
        br = e["byte_range"]
        out[path] = (hashlib.sha256(buf if br is None else buf[br[0] : br[1]]).hexdigest(), br, e["dtype"], e["shape"])
    return out


def test_reference_snapshot_runs_on_this_builds_batcher_and_scheduler(tmp_path, monkeypatch):
    sys.path.insert(0, REF)
    try:
        import torchsnapshot as ref
        import torchsnapshot.snapshot as ref_snapshot
    finally:
        sys.path.remove(REF)
    from torchsnapshot_b200 import batcher, scheduler

    def app_state():
        torch.manual_seed(0)
        model = torch.nn.Sequential(torch.nn.Linear(64, 32), torch.nn.Linear(32, 8))
        optim = torch.optim.Adam(model.parameters())
        model(torch.randn(4, 64)).sum().backward()
        optim.step()
        return {"model": model, "optim": optim, "extra": ref.StateDict(bf=torch.randn(10).bfloat16(), n=3, big=torch.randn(300, 300))}

    state = app_state()
    plain = str(tmp_path / "plain")
    ref.Snapshot.take(plain, state)

    # the seam: planner + executor of the hot path
    monkeypatch.setattr(ref_snapshot, "batch_write_requests", batcher.batch_write_requests)
    monkeypatch.setattr(ref_snapshot, "batch_read_requests", batcher.batch_read_requests)
    monkeypatch.setattr(ref_snapshot, "sync_execute_write_reqs", scheduler.sync_execute_write_reqs)
    monkeypatch.setattr(ref_snapshot, "sync_execute_read_reqs", scheduler.sync_execute_read_reqs)
    patched = str(tmp_path / "patched")
    snap = ref.Snapshot.take(patched, state)
    assert _payloads(patched) == _payloads(plain)

    fresh = app_state()
    for p in fresh["model"].parameters():
        p.data.zero_()
    fresh["extra"]["big"] = torch.zeros(300, 300)
    snap.restore(fresh)
    for (k, a), (_k, b) in zip(fresh["model"].state_dict().items(), state["model"].state_dict().items()):
        assert torch.equal(a, b), k
    assert torch.equal(fresh["extra"]["big"], state["extra"]["big"]) and fresh["extra"]["n"] == 3
    # and a snapshot written through the patched path is readable by the unpatched reference
    monkeypatch.undo()
    again = app_state()
    again["extra"]["big"] = torch.zeros(300, 300)
    ref.Snapshot(patched).restore(again)
    assert torch.equal(again["extra"]["big"], state["extra"]["big"])
