"""Compile-time muscle length-range estimation (MuJoCo's mj_setLengthRange), run on the GPU with the step kernel itself.

Models that give no `lengthrange=` on their muscles (config 2, myosuite/simhive/myo_sim/finger/myofinger_v0.xml:11-15)
get it from a short simulation at model-compile time: for every such actuator and for both directions, push the
mechanism along the actuator's moment arm with a force that would produce `accel` of acceleration, with velocity
damping and only joint / tendon limits active, and record the actuator length the mechanism settles at
(MuJoCo 3.2.8 `mjLROpt` defaults: mode muscle, useexisting, accel 20, timeconst 1, timestep 0.01, inttotal 10,
interval 2, tolrange 0.05).  The 2 * nu (actuator, direction) simulations run as 2 * nu worlds of one batch.

Difference from MuJoCo's evalAct, which damps qvel between mj_step1 and mj_step2: here the damping is applied before
the whole step, so the velocity-dependent bias force sees the damped velocity.  The recorded quantity is the length at
the static equilibrium the process converges to, which does not depend on that detail.  tests/test_lengthrange.py
repeats the procedure with the float64 CPU oracle and checks the committed fixture against it.
"""
from __future__ import annotations

import copy
import math

import numpy as np
import torch

from . import mjx

# mjtDisableBit values (reference: myosuite/physics/sim_scene.py:120-142 lists the same bit positions)
_DSBL = {"frictionloss": 1 << 2, "contact": 1 << 4, "passive": 1 << 5, "gravity": 1 << 6, "actuation": 1 << 10}
LROPT = dict(useexisting=True, accel=20.0, maxforce=0.0, timeconst=1.0, timestep=0.01, inttotal=10.0, interval=2.0, tolrange=0.05)


def needs_length_range(m) -> np.ndarray:
    """muscle actuators without a valid lengthrange (lengthrange[0] >= lengthrange[1])"""
    lr = np.asarray(m.actuator_lengthrange, dtype=np.float64).reshape(-1, 2)
    muscle = (np.asarray(m.actuator_gaintype) == 2) | (np.asarray(m.actuator_biastype) == 2)
    return np.nonzero(muscle & ~(lr[:, 0] < lr[:, 1]))[0]


def set_length_range(m, device: int = 0, **opt):
    """Fills `m.actuator_lengthrange` for the actuators that need it; returns the (ids, ranges, spread) it computed."""
    o = dict(LROPT); o.update(opt)
    ids = needs_length_range(m)
    if len(ids) == 0:
        return ids, np.zeros((0, 2)), np.zeros((0, 2))
    sim = copy.copy(m)
    sim.opt_disableflags = int(m.opt_disableflags) | _DSBL["frictionloss"] | _DSBL["contact"] | _DSBL["passive"] | _DSBL["gravity"] | _DSBL["actuation"]
    sim.opt_timestep = float(o["timestep"])
    lr = np.asarray(m.actuator_lengthrange, dtype=np.float64).reshape(-1, 2).copy()
    lr[ids] = [0.0, 1.0]                       # any valid range: actuation is disabled during the estimation
    sim.actuator_lengthrange = lr
    model = mjx.put_model(sim, device=device)
    nw = 2 * len(ids)
    d = mjx.make_data(model, nw, fields=("actuator_length", "actuator_moment", "qM", "qfrc_applied"))
    dev = d.qpos.device
    act_of_world = torch.as_tensor(np.repeat(ids, 2), device=dev)
    sign = torch.as_tensor(np.tile([-1.0, 1.0], len(ids)), dtype=torch.float32, device=dev)     # side 0: shorten, side 1: lengthen
    damp = math.exp(-o["timestep"] / max(0.01, o["timeconst"]))
    nstep = int(round(o["inttotal"] / o["timestep"]))
    first_rec = nstep - int(round(o["interval"] / o["timestep"]))
    lo = torch.full((nw,), float("inf"), device=dev); hi = torch.full((nw,), -float("inf"), device=dev)
    rows = torch.arange(nw, device=dev)
    for k in range(nstep):
        mjx.forward(model, d)                                            # position-dependent moment arm and inertia
        moment = d.actuator_moment[rows, act_of_world]                   # [nw, nv]
        x = torch.linalg.solve(d.qm.double(), moment.double().unsqueeze(-1)).squeeze(-1)
        nrm = x.norm(dim=1).clamp_min(1e-15)
        frc = moment.double() * (sign.double() * o["accel"] / nrm).unsqueeze(-1)
        if o["maxforce"] > 0:
            fn = frc.norm(dim=1).clamp_min(1e-15)
            frc = frc * torch.clamp(o["maxforce"] / fn, max=1.0).unsqueeze(-1)
        d.qfrc_applied.copy_(frc.float())
        d.qvel.mul_(damp)
        mjx.step(model, d, 1)
        if k >= first_rec:
            length = d.actuator_length[rows, act_of_world]
            lo = torch.minimum(lo, length); hi = torch.maximum(hi, length)
    torch.cuda.synchronize(dev)
    if int(d.status.ne(0).sum()) or not bool(torch.isfinite(lo).all() and torch.isfinite(hi).all()):
        raise RuntimeError("length-range simulation diverged")
    lo, hi = lo.double().cpu().numpy().reshape(-1, 2), hi.double().cpu().numpy().reshape(-1, 2)
    rng = np.stack([lo[:, 0], hi[:, 1]], axis=1)                         # side 0 gives the minimum, side 1 the maximum
    spread = hi - lo                                                     # residual motion inside the recording interval
    if np.any(rng[:, 0] >= rng[:, 1]):
        raise RuntimeError("invalid length range (min >= max)")
    if np.any(spread > o["tolrange"] * (rng[:, 1] - rng[:, 0])[:, None]):
        raise RuntimeError("length-range simulation did not converge (mjLROpt.tolrange)")
    out = np.asarray(m.actuator_lengthrange, dtype=np.float64).reshape(-1, 2).copy()
    out[ids] = rng
    m.actuator_lengthrange = out
    return ids, rng, spread
