import os, sys, json, tempfile, shutil
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, numpy as np
import tsb200_loader; tsb200_loader.load()
import bench
from torchsnapshot_b200.snapshot import Snapshot, StateDict
scale=float(sys.argv[1]) if len(sys.argv)>1 else 0.25
state=bench.build_state("cuda:0", scale=scale)
print("tensors",len(state),"bytes",bench.payload_bytes(state))
outs=StateDict({k: torch.zeros_like(v) for k,v in state.items()})
for trial in range(3):
    d=tempfile.mkdtemp(dir="/dev/shm")
    snap=Snapshot.take(d, {"ddp": StateDict(state)}, replicated=["**"])
    # check files against device state
    meta=json.load(open(os.path.join(d,".snapshot_metadata")))["manifest"]
    cache={}
    badw=0
    for path,e in meta.items():
        if e["type"]!="Tensor": continue
        loc=e["location"]
        if loc not in cache: cache={loc: np.fromfile(os.path.join(d,loc),dtype=np.uint8)}
        buf=cache[loc]; br=e["byte_range"]; got=buf if br is None else buf[br[0]:br[1]]
        name=path[len("0/ddp/"):].replace("%2F","/")
        want=state[name].detach().cpu().contiguous().reshape(-1).view(torch.uint8).numpy()
        if not np.array_equal(got,want):
            badw+=1
            idx=np.nonzero(got!=want)[0]
            if badw<=5: print("trial",trial,"WRITE MISMATCH",path,loc,br,"nbad",len(idx),"first",idx[:3],"last",idx[-3:],"n",len(want))
    for v in outs.values(): v.zero_()
    snap.restore({"ddp": outs})
    badr=0
    for k in state:
        if not torch.equal(outs[k],state[k]):
            badr+=1
            if badr<=5:
                a=outs[k].reshape(-1).view(torch.uint8) if outs[k].ndim else outs[k].reshape(1).view(torch.uint8)
                b=state[k].reshape(-1).view(torch.uint8) if state[k].ndim else state[k].reshape(1).view(torch.uint8)
                idx=torch.nonzero(a!=b).flatten()
                print("trial",trial,"RESTORE MISMATCH",k,tuple(state[k].shape),"nbad",len(idx),"first",idx[:3].tolist(),"last",idx[-3:].tolist(), "out zero?", bool((a[idx]==0).all()))
    print("trial",trial,"bad writes",badw,"bad restores",badr,flush=True)
    shutil.rmtree(d)
