"""Developer script (GPU box): start / end of every kernel of one myo_step on the device clock (MYO_NEWTON_TIMING), averaged over steps."""
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
d = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields=os.environ.get("FIELDS", "all"))
L = lib.lib()
A, N = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (300, 20)
def run(n0, n1):
    for i in range(n0, n1):
        if i % 10 == 0: d.storage("ctrl").copy_(partition.ctrl_hash_torch(m.nu, 0, nw, 1234, i // 10, "cuda"))
        mjx.step(model, d, 1)
run(0, A)
names = ["kin", "tendon", "inertia", "collide", "velact", "finish", "conout", "newton", "collideMpr", "collideCompact", "makecon", "posout", "tenout"]
buf = (ctypes.c_ulonglong * 64)()
L.myo_debug_timeline(d._h, buf)
acc = np.zeros((len(names), 2)); tot = 0.0
for s in range(N):
    run(A + s, A + s + 1)
    L.myo_debug_timeline(d._h, buf)
    a = np.array(buf[:2 * len(names)], dtype=np.float64).reshape(-1, 2)
    t0 = a[0, 0]
    acc += (a - t0) / 1e3; tot += (a[:, 1].max() - t0) / 1e3
    mpr = np.array(buf[48:50], dtype=np.float64) if s == 0 else mpr
acc /= N
print("step %d..%d: kernel start / end / duration on the device clock, us from the start of kin (mean of %d steps, each synchronised)" % (A, A + N, N))
for k in np.argsort(acc[:, 0]):
    print("%-16s %8.1f %8.1f %8.1f" % (names[k], acc[k, 0], acc[k, 1], acc[k, 1] - acc[k, 0]))
print("step span %.1f us" % (tot / N))
print("convex narrow phase, one step: pairs listed %d, with a portal %d (per world-step %.2f / %.2f)" % (mpr[0], mpr[1], mpr[0] / nw, mpr[1] / nw))
