"""Pins the oracle (and the model compiler) to REAL MuJoCo whenever one is importable -- the only route by which parity of
the physics step can leave "unpinned" (VERDICT r1 item 4a).

`mujoco` 3.2.8 is third-party code that is neither vendored in /root/reference nor installed in this image, so on this
image every test below SKIPS, and DESIGN.md keeps saying PARITY UNPINNED.  With `pip install mujoco==3.2.8` (or a
driver-populated baseline/_ref) and the reference's MJCF files reachable (MYO_MJCF_ROOT), they run:
  * compiled constants of our MJCF compiler (mjcf.py + setconst.py) against mj_loadXML's;
  * one mj_forward and one mj_step of the oracle against mj_forward / mj_step, field by field."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import mujoco_probe  # noqa: E402

MODELS = ["elbow", "finger", "hand", "hand_pose", "hand_sar", "legs"]


def _load(name):
    got = mujoco_probe.load(name)
    if got is None:
        pytest.skip("no importable mujoco (or no MJCF files): parity of the physics step stays UNPINNED on this image")
    return got


def test_probe_reports_absence_cleanly():
    """The probe itself never raises: it returns a module or None, and a path or None."""
    m = mujoco_probe.mujoco_module()
    assert m is None or hasattr(m, "mj_step")
    p = mujoco_probe.mjcf_path("hand")
    assert p is None or os.path.exists(p)
    assert mujoco_probe.mjcf_path("no-such-model") is None


@pytest.mark.parametrize("name", MODELS)
def test_compiled_constants_match_mujoco(models, name):
    mj, mm = _load(name)
    m = models(name)
    for f, tol in (("body_mass", 1e-9), ("body_inertia", 1e-7), ("body_ipos", 1e-8), ("body_iquat", 1e-7), ("dof_invweight0", 1e-6),
                   ("body_invweight0", 1e-6), ("tendon_length0", 1e-8), ("tendon_invweight0", 1e-6), ("actuator_acc0", 1e-6),
                   ("actuator_lengthrange", 1e-4), ("qpos0", 0), ("jnt_range", 0), ("geom_size", 0), ("geom_rbound", 1e-9)):
        a, b = np.asarray(getattr(m, f), dtype=np.float64), np.asarray(getattr(mm, f), dtype=np.float64)
        assert a.shape == b.shape or a.size == b.size, f
        np.testing.assert_allclose(a.ravel(), b.ravel(), rtol=max(tol, 1e-12), atol=max(tol, 1e-12) * max(1.0, np.abs(b).max()), err_msg=f"{name}: {f}")
    assert (m.nq, m.nv, m.nu, m.nbody, m.ngeom, m.nsite, m.ntendon, m.nM) == (mm.nq, mm.nv, mm.nu, mm.nbody, mm.ngeom, mm.nsite, mm.ntendon, mm.nM)


@pytest.mark.parametrize("name", MODELS)
def test_oracle_forward_and_step_match_mujoco(models, name):
    import oracle
    import parity
    mj, mm = _load(name)
    m = models(name)
    qpos, qvel, act, ctrl = parity.random_state(m, 4, 99, qpos_spread=0.8)
    o = oracle.Oracle(m)
    d = mj.MjData(mm)
    for w in range(4):
        mj.mj_resetData(mm, d)
        d.qpos[:] = qpos[w]; d.qvel[:] = qvel[w]; d.act[:] = act[w]; d.ctrl[:] = ctrl[w]
        mj.mj_forward(mm, d)
        o.reset(); o.set_state(qpos[w], qvel[w], act[w], ctrl[w]); o.forward()
        for f in ("xpos", "xquat", "site_xpos", "ten_length", "actuator_length", "actuator_velocity", "actuator_force", "qfrc_bias",
                  "qfrc_passive", "qfrc_actuator", "qacc_smooth", "qacc", "qfrc_constraint", "subtree_com", "cinert", "cvel"):
            ref = np.asarray(getattr(d, f), dtype=np.float64).ravel()
            got = np.asarray(o.field(f), dtype=np.float64).ravel()
            np.testing.assert_allclose(got, ref, rtol=1e-7, atol=1e-7 * max(1.0, np.abs(ref).max()), err_msg=f"{name} world {w}: {f}")
        assert o.ncon == d.ncon and o.nefc == d.nefc
        n = d.ncon
        assert np.array_equal(np.asarray(o.contact_geom).reshape(-1, 2)[:n], np.stack([c.geom for c in d.contact[:n]]) if n else np.zeros((0, 2), dtype=int))
        mj.mj_step(mm, d); o.step(1)
        np.testing.assert_allclose(np.asarray(o.qpos), d.qpos, rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(np.asarray(o.qvel), d.qvel, rtol=1e-7, atol=1e-7)
        np.testing.assert_allclose(np.asarray(o.act), d.act, rtol=1e-9, atol=1e-9)
