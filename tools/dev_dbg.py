"""Developer script (GPU box): elbow, small batch: state after forward / each single step, against the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, torch, parity, bench, oracle
from myosuite_mjx_b200 import mjx
name = sys.argv[1] if len(sys.argv) > 1 else "elbow"
nw = 32
m = bench.load_model(name)
qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
model = mjx.put_model(m)
d = mjx.put_data(model, dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl), nworld=nw, fields="all")
o = oracle.Oracle(m); o.reset(); o.set_state(qpos[0], qvel[0], act[0], ctrl[0])
def show(tag):
    torch.cuda.synchronize()
    print(tag, "qpos", d.qpos[0, :4].cpu().numpy(), "qvel", d.qvel[0, :4].cpu().numpy(), "act", d.act[0, :3].cpu().numpy(), "qacc", d.qacc[0, :4].cpu().numpy(),
          "time", float(d.time[0, 0]), "status", int(d.status[0, 0]), "nefc", int(d.nefc[0, 0]))
    print("   oracle qpos", np.array(o.qpos)[:4], "qvel", np.array(o.qvel)[:4], "qacc", np.array(o.qacc)[:4])
show("init")
if "fwd" in sys.argv:
    mjx.forward(model, d); o.forward(); show("forward")
for i in range(3):
    mjx.step(model, d, 1); o.step(1); show(f"step{i}")
