"""Developer script (GPU box): workload statistics of the bench rollout (contacts, rows, solver iterations)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from myosuite_mjx_b200 import mjx
m = bench.load_model(os.environ.get("MODEL", "hand"))
model = mjx.put_model(m)
nw = 8192
q0 = bench.initial_qpos(m, nw, 1234)
d = mjx.put_data(model, dict(qpos=q0), nworld=nw, fields=("ncon", "nefc", "solver_niter", "qacc"), nconmax=int(os.environ.get("NCONMAX", 0)), njmax=int(os.environ.get("NJMAX", 0)))
gen = torch.Generator(device="cuda"); gen.manual_seed(1234)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 60):
    if i % 10 == 0: d.storage("ctrl").uniform_(0, 1, generator=gen)
    mjx.step(model, d, 1)
    if i % 10 == 9:
        torch.cuda.synchronize()
        nc, ne, ni = d.ncon.cpu().numpy().ravel(), d.nefc.cpu().numpy().ravel(), d.solver_niter.cpu().numpy().ravel()
        print(f"step {i+1}: ncon mean {nc.mean():.2f} p50 {np.median(nc)} p99 {np.percentile(nc,99)} max {nc.max()} | nefc mean {ne.mean():.1f} p99 {np.percentile(ne,99)} max {ne.max()} | niter mean {ni.mean():.2f} hist {np.bincount(ni)[:12]} | flagged {(d.status.cpu().numpy()!=0).sum()} bits {[(b, int(((d.status.cpu().numpy() & b)!=0).sum())) for b in (1,2,4,8,16,32)]}")
