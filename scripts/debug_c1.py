import os, sys, json, tempfile, shutil
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT,"tests"))
import torch, numpy as np
import tsb200_loader; tsb200_loader.load()
from test_snapshot_cpu import _c1_app_state
from test_gpu_snapshot import _to_cuda
from torchsnapshot_b200.snapshot import Snapshot
import oracle
for trial in range(6):
    app=_to_cuda(_c1_app_state())
    d=tempfile.mkdtemp()
    Snapshot.take(d, app)
    meta=json.load(open(os.path.join(d,".snapshot_metadata")))
    bad=0
    for path,e in meta["manifest"].items():
        if e["type"]!="Tensor": continue
        buf=open(os.path.join(d,e["location"]),"rb").read()
        br=e["byte_range"]; got=buf[br[0]:br[1]]
        name=path[len("0/"):]
        if name.startswith("model/"):
            t=app["model"].state_dict()[name[len("model/"):]]
        else:
            parts=name.split("/"); t=app["optim"].state_dict()["state"][int(parts[2])][parts[3]]
        want=oracle.serialize(t.detach().cpu())
        if got!=want:
            bad+=1
            g=np.frombuffer(got,dtype=np.uint8); w=np.frombuffer(want,dtype=np.uint8)
            idx=np.nonzero(g!=w)[0]
            print("trial",trial,"MISMATCH",path,br,"nbad",len(idx),"first",idx[:5],"last",idx[-5:], "got zeros?", int((g[idx]==0).all()))
    print("trial",trial,"bad",bad, flush=True)
    shutil.rmtree(d)
