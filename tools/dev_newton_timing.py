"""Developer script (GPU box): per-phase clock sums of the solver kernel (MYO_NEWTON_TIMING), steady state of the bench rollout."""
import ctypes, os, sys
os.environ["MYO_NEWTON_TIMING"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from myosuite_mjx_b200 import mjx, lib, partition
m = bench.load_model("hand")
model = mjx.put_model(m)
nw = 8192
d = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields="all")
L = lib.lib(); L.myo_debug_timing.restype = ctypes.c_int64
def run(n0, n1):
    for i in range(n0, n1):
        if i % 10 == 0: d.storage("ctrl").copy_(partition.ctrl_hash_torch(m.nu, 0, nw, 1234, i // 10, "cuda"))
        mjx.step(model, d, 1)
A, B = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (300, 400)
run(0, A)
n = L.myo_debug_timing(d._h, None)
a = np.zeros(n, dtype=np.int64); L.myo_debug_timing(d._h, a.ctypes.data_as(ctypes.c_void_p)); a0 = a.reshape(-1, 32).copy()
run(A, B)
L.myo_debug_timing(d._h, a.ctypes.data_as(ctypes.c_void_p)); a1 = a.reshape(-1, 32) - a0
names = ["makeConstraint", "warm start / residuals", "force + gradient", "hessian assemble", "hessian update", "cholesky", "chol solve", "M s, J s", "line search", "update + cost", "final force", "passes"]
per = a1 / float(B - A)
print("steps", A, B, "ncon mean %.2f nefc mean %.2f max %d" % (float(d.ncon.float().mean()), float(d.nefc.float().mean()), int(d.nefc.max())))
tot = per[:, :11].sum(axis=1) + per[:, 12:16].sum(axis=1)
print("tiles", len(per), "cycles per tile per step: mean %.0f  p50 %.0f  max %.0f  (%.1f us at 1.965 GHz max)" % (tot.mean(), np.median(tot), tot.max(), tot.max() / 1965))
print("passes per tile per step: mean %.2f max %.2f" % (per[:, 11].mean(), per[:, 11].max()))
for k, nme in enumerate(names[:11]):
    print("%-24s mean %8.0f cycles (%5.1f %%)   per pass %8.0f" % (nme, per[:, k].mean(), 100 * per[:, k].mean() / tot.mean(), per[:, k].mean() / max(per[:, 11].mean(), 1e-9) if k >= 2 and k <= 9 else per[:, k].mean()))
for k, nme in ((12, "  (M s)"), (13, "  (Jb s + barrier)"), (15, "  (constraintForce)"), (16, "    (cf: row forces)"), (17, "    (cf: base forces)"), (18, "    (cf: J^T f)")):
    print("%-24s mean %8.0f cycles (%5.1f %%)   per pass %8.0f" % (nme, per[:, k].mean(), 100 * per[:, k].mean() / tot.mean(), per[:, k].mean() / max(per[:, 11].mean(), 1e-9)))
