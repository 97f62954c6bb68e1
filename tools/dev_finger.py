import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, parity, oracle
from myosuite_mjx_b200 import mjcf
m = mjcf.Model.load(os.path.join(ROOT, "myosuite_mjx_b200/models/finger.npz"))
nw = 8
qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
ref, ex = parity.oracle_forward(m, qpos, qvel, act, ctrl, ["ten_length", "ten_J", "site_xpos", "geom_xpos", "geom_xmat"])
model, d, got = parity.gpu_forward(m, qpos, qvel, act, ctrl)
np.set_printoptions(precision=5, suppress=True, linewidth=200)
print("site err", np.abs(got.site_xpos.reshape(nw,-1) - ref["site_xpos"]).max(), "geom err", np.abs(got.geom_xpos.reshape(nw,-1)-ref["geom_xpos"]).max(), np.abs(got.geom_xmat.reshape(nw,-1)-ref["geom_xmat"]).max())
for w in range(nw):
    print("w", w, "q", qpos[w], "\n  len gpu", got.ten_length[w], "\n  len ref", ref["ten_length"][w])
