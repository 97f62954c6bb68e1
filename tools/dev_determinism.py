"""Developer script (GPU box): the same rollout twice in one process; are the final states bit-identical?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from myosuite_mjx_b200 import mjx, partition
m = bench.load_model("hand")
model = mjx.put_model(m)
nw, nsteps = 8192, int(sys.argv[1]) if len(sys.argv) > 1 else 400
outs = []
for rep in range(2):
    d = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields="all")
    first = None
    for i in range(nsteps):
        if i % 10 == 0: d.storage("ctrl").copy_(partition.ctrl_hash_torch(m.nu, 0, nw, 1234, i // 10, "cuda"))
        mjx.step(model, d, 1)
        if rep == 1 and i % 20 == 19 and first is None:
            pass
    torch.cuda.synchronize()
    outs.append((d.qpos.clone(), d.qacc.clone(), d.status.clone()))
    del d
dq = (outs[0][0] != outs[1][0]).any(dim=1)
print("steps", nsteps, "worlds whose final qpos differs between two identical runs:", int(dq.sum()), "of", nw, "; flagged:", int((outs[0][2] != 0).sum()), int((outs[1][2] != 0).sum()))
