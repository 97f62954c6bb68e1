"""Developer script (GPU box): status bits of the bench rollout over time (which bit, how many worlds, from which step on)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from myosuite_mjx_b200 import mjx, partition
m = bench.load_model("hand")
model = mjx.put_model(m)
nw = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 1400
d = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields=os.environ.get("FIELDS", "all"))
names = {1: "BADQACC", 2: "NCON_OVERFLOW", 4: "NEFC_OVERFLOW", 8: "SOLVER_MAXITER", 16: "BADQPOS", 32: "BADQVEL"}
last = None
for i in range(nsteps):
    if i % 10 == 0: d.storage("ctrl").copy_(partition.ctrl_hash_torch(m.nu, 0, nw, 1234, i // 10, "cuda"))
    mjx.step(model, d, 1)
    if i % 50 == 49:
        st = d.status.cpu().numpy().ravel()
        cur = {names[b]: int(((st & b) != 0).sum()) for b in names if ((st & b) != 0).any()}
        if cur != last:
            print("step", i + 1, cur, "ncon max", int(d.ncon.max()) if "ncon" in d.fields() else None, "nefc max", int(d.nefc.max()) if "nefc" in d.fields() else None, "niter max", int(d.solver_niter.max()) if "solver_niter" in d.fields() else None, flush=True)
            last = cur
print("done", last)
