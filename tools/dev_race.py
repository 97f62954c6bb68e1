"""Developer script (GPU box): the same forward pass on two batches with identical clenched-hand states; which scratch regions differ?"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from myosuite_mjx_b200 import mjx, lib, partition
NAMES = ("total qpos qvel act ctrl frames xanchor xaxis rootcom cinert crb cdof cdof_dot cvel cacc qM qLD qLDinv qH qHinv tenlen tenvel tenJ actlen actvel "
         "actforce actdot bias passive qfrc_act smooth qfrc_con qacc_smooth qacc warm tmpv tforce counts wstate slots ulen uwrap uwpt gpre upts tmpbody "
         "cgpose cgshape con conJ conU conW efc H Hsave nvec limrow basev cand ctmp nv hshared wantw").split()
m = bench.load_model("hand")
model = mjx.put_model(m)
nw, nwarm = int(os.environ.get("NW", 8192)), int(sys.argv[1]) if len(sys.argv) > 1 else 300
d = mjx.put_data(model, dict(qpos=bench.initial_qpos(m, nw, 1234)), nworld=nw, fields="all")
for i in range(nwarm):
    if i % 10 == 0: d.storage("ctrl").copy_(partition.ctrl_hash_torch(m.nu, 0, nw, 1234, i // 10, "cuda"))
    mjx.step(model, d, 1)
torch.cuda.synchronize()
state = {k: getattr(d, k).clone() for k in ("qpos", "qvel", "act", "ctrl", "qacc_warmstart")}
L = lib.lib(); L.myo_debug_scratch.restype = ctypes.c_int64
def run():
    e = mjx.put_data(model, {k: v.cpu().numpy() for k, v in state.items()}, nworld=nw, fields="all")
    for _ in range(int(os.environ.get("STEP", 0))): mjx.step(model, e, 1)
    if not os.environ.get("STEP"): mjx.forward(model, e)
    torch.cuda.synchronize()
    lay = (ctypes.c_int * 80)(); n = L.myo_debug_layout(e._h, lay, 80)
    lay = dict(zip(NAMES, list(lay)[:n]))
    tot = L.myo_debug_scratch(e._h, None)
    buf = np.zeros(tot, dtype=np.float32)
    L.myo_debug_scratch(e._h, buf.ctypes.data_as(ctypes.c_void_p))
    return lay, buf.view(np.uint32).reshape(-1, lay["total"], 32), e.qacc.clone(), int(e.nefc.max())
layA, A, qa, nefc = run()
layB, B, qb, _ = run()
print("layout fields", len(layA), "total", layA["total"], "tiles", A.shape[0], "nefc max", nefc, "| qacc differs in", int((qa != qb).any(dim=1).sum()), "worlds")
offs = sorted((v, k) for k, v in layA.items() if k not in ("total", "nv", "hshared", "wantw") and (v > 0 or k == "qpos"))
for i, (o, k) in enumerate(offs):
    end = offs[i + 1][0] if i + 1 < len(offs) else layA["total"]
    diff = (A[:, o:end, :] != B[:, o:end, :])
    if diff.any():
        w = np.argwhere(diff)
        print("%-10s [%5d, %5d): %7d differing words in %4d worlds; first at tile %d elem %d lane %d" % (k, o, end, int(diff.sum()), len(set((a, c) for a, b, c in w)), w[0][0], w[0][1], w[0][2]))
