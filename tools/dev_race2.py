"""Developer script (GPU box): two batches stepped side by side from the same start; when and where do they first differ?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from myosuite_mjx_b200 import mjx, partition
m = bench.load_model(os.environ.get("MODEL", "hand"))
model = mjx.put_model(m)
nw, nsteps = 8192, int(sys.argv[1]) if len(sys.argv) > 1 else 300
A = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields="all")
B = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields="all")
seen = torch.zeros(nw, dtype=torch.bool, device="cuda")
nrep = 0
for i in range(nsteps):
    if i % 10 == 0:
        c = partition.ctrl_hash_torch(m.nu, 0, nw, 1234, i // 10, "cuda")
        A.storage("ctrl").copy_(c); B.storage("ctrl").copy_(c)
    mjx.step(model, A, 1); mjx.step(model, B, 1)
    dq = (A.qacc != B.qacc).any(dim=1) & ~seen
    if dq.any():
        idx = dq.nonzero().flatten()
        for w in idx[:4].tolist():
            if nrep < int(os.environ.get("NREP", 24)):
                print("step %d world %d (lane %d): niter %d/%d nefc %d/%d ncon %d/%d  max|dqacc| %.3e  |qacc| %.3e  fields differing: %s" % (
                    i, w, w % 32, int(A.solver_niter[w, 0]), int(B.solver_niter[w, 0]), int(A.nefc[w, 0]), int(B.nefc[w, 0]), int(A.ncon[w, 0]), int(B.ncon[w, 0]),
                    float((A.qacc[w] - B.qacc[w]).abs().max()), float(A.qacc[w].abs().max()),
                    [k for k in ("xpos", "ten_length", "qM", "qfrc_bias", "actuator_force", "qacc_smooth", "contact_dist", "efc_aref", "efc_j", "efc_d", "efc_force", "qfrc_constraint") if (getattr(A, k)[w] != getattr(B, k)[w]).any()]), flush=True)
                nrep += 1
        seen |= dq
print("worlds that ever differed:", int(seen.sum()))
