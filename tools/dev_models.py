"""Developer script (GPU box): every model through forward + 3 steps in its own process, with per-kernel synchronisation,
so a faulting kernel is named and does not poison the other models."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np, torch, parity, bench
    from myosuite_mjx_b200 import mjx
    name, nw = sys.argv[2], int(sys.argv[3])
    m = bench.load_model(name)
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl), nworld=nw, fields="all")
    mjx.forward(model, d); torch.cuda.synchronize()
    print(name, nw, os.environ.get("MYO_SOLVER", "tile"), "forward ok: ncon", int(d.ncon.sum()), "nefc", int(d.nefc.sum()), "niter", int(d.solver_niter.sum()), "qacc finite", bool(torch.isfinite(d.qacc).all()),
          "qacc sum", float(d.qacc.double().abs().sum()), "qfrc_con sum", float(d.qfrc_constraint.double().abs().sum()))
    mjx.step(model, d, 3); torch.cuda.synchronize()
    st = d.status.cpu().numpy().ravel()
    print(name, nw, "step ok: qpos finite", bool(torch.isfinite(d.qpos).all()), "status bits", {b: int(((st & b) != 0).sum()) for b in (1, 2, 4, 8, 16, 32)},
          "qacc abs max", float(d.qacc.abs().max()), "time", float(d.time.max()))
else:
    for name in sys.argv[1:] or ["elbow", "finger", "hand", "hand_sar", "legs"]:
        for nw, solver in ((32, "team"), (32, "tile"), (48, "tile")):
            r = subprocess.run([sys.executable, __file__, "one", name, str(nw)], capture_output=True, text=True, env=dict(os.environ, MYO_DEBUG_SYNC="1", MYO_SOLVER=solver))
            err = "\n".join(l for l in r.stderr.split("\n") if "myostep" in l or "Error" in l)
            print((r.stdout + err).strip()[-900:])
