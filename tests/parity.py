"""Shared helpers of the parity tests: seeded random states, CUDA path vs CPU oracle field by field."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

# fields compared after a forward pass: name -> (atol scale reference field or None, rtol)
FLOAT_FIELDS = ["xpos", "xquat", "xmat", "xipos", "ximat", "xanchor", "xaxis", "geom_xpos", "geom_xmat", "site_xpos", "site_xmat",
                "subtree_com", "cinert", "crb", "cdof", "cdof_dot", "cvel", "ten_length", "ten_velocity", "ten_J",
                "actuator_length", "actuator_velocity", "actuator_moment", "actuator_force", "act_dot", "qM", "qLDiagInv",
                "qfrc_bias", "qfrc_passive", "qfrc_spring", "qfrc_damper", "qfrc_actuator", "qfrc_smooth", "qacc_smooth",
                "qfrc_constraint", "qacc"]


def random_state(m, nworld, seed, qpos_spread=0.3, qvel_scale=1.0, free_height=None):
    """Seeded states: hinge/slide joints uniform inside the middle `qpos_spread` of their range around qpos0 (or
    +-qpos_spread rad when unlimited), free joints at qpos0, qvel ~ N(0, qvel_scale), act, ctrl ~ U[0,1]."""
    rng = np.random.default_rng(seed)
    qpos = np.tile(np.asarray(m.qpos0, dtype=np.float64), (nworld, 1))
    for j in range(m.njnt):
        if m.jnt_type[j] in (2, 3):
            a = m.jnt_qposadr[j]
            if m.jnt_limited[j]:
                lo, hi = m.jnt_range[j]
                c = np.clip(m.qpos0[a], lo, hi)
                half = 0.5 * qpos_spread * (hi - lo)
                qpos[:, a] = np.clip(c + rng.uniform(-half, half, nworld), lo, hi)
            else:
                qpos[:, a] = m.qpos0[a] + rng.uniform(-qpos_spread, qpos_spread, nworld) * (0.02 if m.jnt_type[j] == 2 else 1.0)
    qvel = rng.normal(0, qvel_scale, (nworld, m.nv))
    act = rng.uniform(0, 1, (nworld, m.na))
    ctrl = rng.uniform(0, 1, (nworld, m.nu))
    return qpos, qvel, act, ctrl


def oracle_forward(m, qpos, qvel, act, ctrl, fields, warm=None, use_float=False):
    import oracle
    o = oracle.Oracle(m, use_float=use_float)
    out = {k: [] for k in fields}
    extra = {"ncon": [], "nefc": [], "contact_geom": [], "contact_dist": [], "contact_pos": [], "contact_frame": [], "niter": [],
             "ten_wrapadr": [], "ten_wrapnum": [], "wrap_obj": [], "wrap_xpos": []}
    for w in range(qpos.shape[0]):
        o.set_state(qpos[w], qvel[w], act[w], ctrl[w], warm[w] if warm is not None else np.zeros(m.nv))
        o.forward()
        for k in fields:
            src = {"qM": "qMdense"}.get(k, k)
            out[k].append(np.array(o.field(src), dtype=np.float64))
        extra["ncon"].append(o.ncon); extra["nefc"].append(o.nefc)
        extra["niter"].append(int(o.solver_niter[0]))
        for k in ("ten_wrapadr", "ten_wrapnum", "wrap_obj", "wrap_xpos"):   # tendon path points (mj_tendon)
            extra[k].append(np.array(o.field(k)))
        extra["contact_geom"].append(np.array(o.contact_geom).reshape(-1, 2))
        extra["contact_dist"].append(np.array(o.contact_dist))
        extra["contact_pos"].append(np.array(o.contact_pos).reshape(-1, 3))
        extra["contact_frame"].append(np.array(o.contact_frame).reshape(-1, 9))
    return {k: np.stack(v) for k, v in out.items()}, extra


def gpu_forward(m, qpos, qvel, act, ctrl, warm=None, nconmax=0, njmax=0, step=0):
    import torch
    from myosuite_mjx_b200 import mjx
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl, qacc_warmstart=warm), nworld=qpos.shape[0],
                     nconmax=nconmax, njmax=njmax)
    if step:
        mjx.step(model, d, step)
    else:
        mjx.forward(model, d)
    torch.cuda.synchronize()
    return model, d, mjx.get_data(model, d)


def field_errors(ref, got_ns, fields):
    """max abs error and error relative to the field's own scale (max |ref| over the batch), per field"""
    rep = {}
    for k in fields:
        r = ref[k]
        g = np.asarray(getattr(got_ns, k.lower()), dtype=np.float64).reshape(r.shape)
        scale = max(np.abs(r).max(), 1e-12)
        err = np.abs(g - r).max() if r.size else 0.0
        rep[k] = (err, err / scale)
    return rep
