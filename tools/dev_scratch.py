"""Developer script (GPU box): run the phase pipeline of one step kernel by kernel (MYO_DEBUG_STOP) and print chosen scratch
slots of world 0 after each kernel.  usage: dev_scratch.py <model> <slot names...>"""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ("total qpos qvel act ctrl frames xanchor xaxis rootcom cinert crb cdof cdof_dot cvel cacc qM qLD qLDinv qH qHinv tenlen tenvel tenJ actlen actvel "
         "actforce actdot bias passive qfrc_act smooth qfrc_con qacc_smooth qacc warm tmpv tforce counts wstate slots ulen uwrap uwpt gpre upts tmpbody "
         "cgpose con conJ conU conW efc H Hsave nvec limrow basev cand ctmp nv hshared wantw").split()
if len(sys.argv) > 1 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np, torch, parity, bench
    from myosuite_mjx_b200 import mjx, lib
    name, slots = sys.argv[2], sys.argv[3:]
    m = bench.load_model(name)
    nw = 32
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl), nworld=nw, fields="all")
    mjx.step(model, d, 1); torch.cuda.synchronize()
    L = lib.lib()
    lay = (ctypes.c_int * 80)(); n = L.myo_debug_layout(d._h, lay, 80)
    lay = dict(zip(NAMES, list(lay)[:n]))
    L.myo_debug_scratch.restype = ctypes.c_int64
    tot = L.myo_debug_scratch(d._h, None)
    buf = np.zeros(tot, dtype=np.float32)
    L.myo_debug_scratch(d._h, buf.ctypes.data_as(ctypes.c_void_p))
    stride = tot // lay["total"]
    S = buf.reshape(lay["total"], stride)
    print("after", os.environ.get("MYO_DEBUG_STOP"), "kernels:", " ".join(f"{s}@{lay[s]}={S[lay[s]:lay[s] + 3, 0]}" for s in slots), "| nan rows", int(np.isnan(S[:, 0]).sum()))
else:
    name = sys.argv[1]
    for k in range(1, 9):
        r = subprocess.run([sys.executable, __file__, "one", name] + sys.argv[2:], capture_output=True, text=True, env=dict(os.environ, MYO_DEBUG_STOP=str(k)))
        print((r.stdout + "\n".join(l for l in r.stderr.split("\n") if "Error" in l or "myostep" in l)).strip()[-700:])
