"""Developer script (GPU box): per-world constrained-solve comparison CUDA vs oracle (double and float builds)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import parity, oracle
from myosuite_mjx_b200 import mjcf
name = sys.argv[1] if len(sys.argv) > 1 else "hand_sar"
m = mjcf.Model.load(os.path.join(ROOT, "myosuite_mjx_b200/models", name + ".npz"))
nw = 32
qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
F = ["qacc", "qacc_smooth", "qfrc_constraint", "qfrc_smooth"]
ref, ex = parity.oracle_forward(m, qpos, qvel, act, ctrl, F)
reff, exf = parity.oracle_forward(m, qpos, qvel, act, ctrl, F, use_float=True)
model, d, got = parity.gpu_forward(m, qpos, qvel, act, ctrl, nconmax=32, njmax=160)
sc = np.abs(ref["qacc"]).max()
print("qacc scale", sc, "meaninertia", m.stat_meaninertia)
for w in range(nw):
    e = np.abs(got.qacc[w] - ref["qacc"][w]); ef = np.abs(reff["qacc"][w] - ref["qacc"][w])
    k = int(e.argmax())
    print(f"w{w:2d} ncon {ex['ncon'][w]:2d} nefc {ex['nefc'][w]:3d} niter ref {ex['niter'][w]} f32cpu {exf['niter'][w]} gpu {int(got.solver_niter[w,0])} "
          f"status {int(got.status[w,0])} | gpu err {e.max():9.3g} (dof {k}, qacc {ref['qacc'][w][k]:9.3g}) f32cpu err {ef.max():9.3g} | smooth err {np.abs(got.qacc_smooth[w]-ref['qacc_smooth'][w]).max():.3g}")
