"""Developer script (GPU box): CUDA path vs oracle, per-field error report for every model."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import parity
from myosuite_mjx_b200 import mjcf

names = sys.argv[1:] or ["elbow", "hand", "hand_pose", "hand_sar", "legs", "finger"]
for name in names:
    m = mjcf.Model.load(os.path.join(ROOT, "myosuite_mjx_b200/models", name + ".npz"))
    nw = 48
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
    t = time.time()
    ref, ex = parity.oracle_forward(m, qpos, qvel, act, ctrl, parity.FLOAT_FIELDS)
    print(f"== {name}: oracle {time.time()-t:.2f}s  ncon {ex['ncon'][:12]} nefc {ex['nefc'][:12]} niter {ex['niter'][:12]}")
    try:
        model, d, got = parity.gpu_forward(m, qpos, qvel, act, ctrl)
    except Exception as e:
        print("   GPU FAILED:", repr(e)); continue
    rep = parity.field_errors(ref, got, parity.FLOAT_FIELDS)
    bad = {k: v for k, v in rep.items() if v[1] > 1e-4}
    print("   worst:", sorted(((v[1], k) for k, v in rep.items()), reverse=True)[:6])
    print("   bad (rel>1e-4):", {k: (float(f"{v[0]:.3g}"), float(f"{v[1]:.3g}")) for k, v in bad.items()})
    print("   gpu ncon", got.ncon[:12, 0], "nefc", got.nefc[:12, 0], "niter", got.solver_niter[:12, 0], "status", np.unique(got.status))
    ncon_ok = np.array_equal(got.ncon[:, 0], np.array(ex["ncon"])); nefc_ok = np.array_equal(got.nefc[:, 0], np.array(ex["nefc"]))
    geom_ok = (not hasattr(got, "contact_geom")) or all(np.array_equal(got.contact_geom.reshape(nw, -1, 2)[w, :ex["ncon"][w]], ex["contact_geom"][w]) for w in range(nw) if ex["ncon"][w] <= d.nconmax)
    print("   ncon match", ncon_ok, "nefc match", nefc_ok, "contact geoms match", geom_ok)
    # multi-step drift
    import torch
    from myosuite_mjx_b200 import mjx
    import oracle
    nsteps = 100
    d2 = mjx.put_data(model, dict(qpos=qpos, qvel=qvel * 0, act=act * 0, ctrl=ctrl), nworld=nw)
    t = time.time(); mjx.step(model, d2, nsteps); torch.cuda.synchronize(); tg = time.time() - t
    o = oracle.Oracle(m)
    errs = []
    for w in range(8):
        o.reset(); o.set_state(qpos[w], qvel[w] * 0, act[w] * 0, ctrl[w]); o.step(nsteps)
        errs.append((np.abs(d2.qpos[w].cpu().numpy() - o.qpos).max(), np.abs(d2.qvel[w].cpu().numpy() - o.qvel).max(), np.abs(o.qvel).max()))
    print(f"   {nsteps}-step drift (qpos err, qvel err, |qvel|): {[tuple(float(f'{x:.2g}') for x in e) for e in errs[:4]]}  gpu {tg*1e3:.1f} ms  status {np.unique(d2.status.cpu().numpy())}")
