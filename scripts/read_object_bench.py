"""read_object random access (BASELINE config 4 / benchmarks/load_tensor analogue): a 8.6 GB fp32 table saved
as a row-wise sharded tensor, read back (a) whole into a CUDA tensor, (b) whole with a 100 MiB memory budget,
(c) a block of 1 M rows."""
import json, os, sys, time, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from types import SimpleNamespace as NS
import tsb200_loader; tsb200_loader.load()
import torchsnapshot_b200.io_preparer as iop
import torchsnapshot_b200.snapshot as snapmod
from torchsnapshot_b200.snapshot import Snapshot, StateDict

rows, cols = 16_777_216, 128
class Sharded:
    def __init__(self, boxes):
        self._shards = []
        g = torch.Generator(device="cuda").manual_seed(0)
        for o, s in boxes:
            t = torch.empty(s, dtype=torch.float32, device="cuda"); t.view(torch.int32).random_(generator=g)
            self._shards.append(NS(tensor=t, metadata=NS(shard_offsets=list(o), shard_sizes=list(s))))
    def local_shards(self): return self._shards
    def metadata(self): return NS(size=[rows, cols])
    def sharding_spec(self): return None
iop._is_sharded_tensor = lambda o: isinstance(o, Sharded); snapmod._is_sharded_tensor = iop._is_sharded_tensor
per = rows // 4
table = Sharded([([i * per, 0], [per, cols]) for i in range(4)])
d = tempfile.mkdtemp(dir="/dev/shm")
try:
    snap = Snapshot.take(d, {"emb": StateDict(t=table)})
    total = rows * cols * 4
    def timed(label, fn, nbytes):
        best = 1e9
        for _ in range(3):
            torch.cuda.synchronize(); t0 = time.monotonic(); out = fn(); torch.cuda.synchronize(); best = min(best, time.monotonic() - t0)
        print(json.dumps({"case": label, "GB": round(nbytes / 1e9, 3), "seconds": round(best, 4), "GBps": round(nbytes / best / 1e9, 2)}), flush=True)
        return out
    dense = torch.empty(rows, cols, device="cuda")
    timed("whole table -> CUDA tensor, no budget", lambda: snap.read_object("0/emb/t", obj_out=dense), total)
    ref = torch.cat([s.tensor for s in table.local_shards()])
    assert torch.equal(dense.view(torch.int32), ref.view(torch.int32))
    dense.zero_()
    timed("whole table -> CUDA tensor, memory_budget_bytes=100 MiB", lambda: snap.read_object("0/emb/t", obj_out=dense, memory_budget_bytes=100 << 20), total)
    assert torch.equal(dense.view(torch.int32), ref.view(torch.int32))
    block = torch.empty(1_000_000, cols, device="cuda")
    timed("rows [0, 1M) -> CUDA tensor (partial-piece reads)", lambda: snap.read_object("0/emb/t", obj_out=block), block.numel() * 4)
    assert torch.equal(block.view(torch.int32), ref[:1_000_000].view(torch.int32))
finally:
    shutil.rmtree(d, ignore_errors=True)
