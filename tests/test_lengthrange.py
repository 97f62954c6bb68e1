"""Compile-time muscle length ranges of the finger model (config 2): the committed fixture was filled by
myosuite_mjx_b200/lengthrange.py running the CUDA step (tools/make_finger_fixture.py on the GPU box); here the same
procedure (MuJoCo's mj_setLengthRange recipe, mjLROpt defaults) is repeated with the float64 CPU oracle."""
import copy
import math

import numpy as np
import pytest

import oracle


def oracle_length_range(m, ids, accel=20.0, timeconst=1.0, timestep=0.01, inttotal=10.0, interval=2.0):
    sim = copy.copy(m)
    sim.opt_disableflags = int(m.opt_disableflags) | (1 << 2) | (1 << 4) | (1 << 5) | (1 << 6) | (1 << 10)
    sim.opt_timestep = timestep
    lr = np.asarray(m.actuator_lengthrange, dtype=np.float64).reshape(-1, 2).copy(); lr[ids] = [0.0, 1.0]
    sim.actuator_lengthrange = lr
    damp = math.exp(-timestep / max(0.01, timeconst))
    nstep = int(round(inttotal / timestep)); first = nstep - int(round(interval / timestep))
    out = np.zeros((len(ids), 2))
    for n, i in enumerate(ids):
        for side in (0, 1):
            o = oracle.Oracle(sim)
            o.reset()
            lo, hi = np.inf, -np.inf
            for k in range(nstep):
                o.forward()
                moment = o.actuator_moment.reshape(m.nu, m.nv)[i].copy()
                M = o.qMdense.reshape(m.nv, m.nv)
                nrm = max(np.linalg.norm(np.linalg.solve(M, moment)), 1e-15)
                o.qfrc_applied[:] = moment * ((2 * side - 1) * accel / nrm)
                o.qvel[:] *= damp
                o.step(1)
                if k >= first:
                    lo, hi = min(lo, o.actuator_length[i]), max(hi, o.actuator_length[i])
            out[n, side] = lo if side == 0 else hi
    return out


def test_finger_fixture_has_valid_ranges(models):
    m = models("finger")
    lr = np.asarray(m.actuator_lengthrange).reshape(-1, 2)
    assert np.all(lr[:, 0] < lr[:, 1])
    o = oracle.Oracle(m); o.reset(); o.forward()
    L0 = np.array(o.actuator_length)
    assert np.all(lr[:, 0] - 1e-6 <= L0) and np.all(L0 <= lr[:, 1] + 1e-6)      # the rest pose lies inside every range
    # the simulated range cannot exceed what the joint limits allow by more than the soft-limit penetration
    from myosuite_mjx_b200 import setconst
    rng = np.random.default_rng(0)
    lo, hi = np.asarray(m.jnt_range)[:, 0], np.asarray(m.jnt_range)[:, 1]
    Ls = np.array([setconst.forward_pos(m, lo + rng.uniform(0, 1, m.nq) * (hi - lo))["ten_length"][np.asarray(m.actuator_trnid).reshape(m.nu, -1)[:, 0]]
                   for _ in range(200)])
    assert np.all(Ls.min(axis=0) > lr[:, 0] - 0.01) and np.all(Ls.max(axis=0) < lr[:, 1] + 0.01)


def test_finger_length_range_matches_oracle_procedure(models):
    m = models("finger")
    ids = np.arange(m.nu)
    ref = oracle_length_range(m, ids)
    got = np.asarray(m.actuator_lengthrange).reshape(-1, 2)
    np.testing.assert_allclose(got, ref, rtol=2e-4, atol=2e-6)


def test_finger_steps_stably(models):
    m = models("finger")
    o = oracle.Oracle(m)
    o.reset(); o.set_state(ctrl=np.full(m.nu, 0.5))
    o.step(500)
    assert np.isfinite(o.qpos).all() and np.abs(o.qpos).max() < 3.0 and np.abs(o.actuator_force).max() < 1e5
