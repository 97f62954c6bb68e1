"""Analytic known-answer tests for the parts of the oracle (and of the model compiler) that nothing else anchors
(VERDICT r1 "next round" 4b, ADVICE): constraint parameters, contact forces, the narrow phase, compiled constants.

Every expected value below is derived by hand from MuJoCo's published definitions (documentation chapters "Computation /
Constraint model", "Modeling / Solver parameters", "XML reference"), not from the code under test.  They do NOT pin the
oracle to MuJoCo (that needs MuJoCo: tests/test_mujoco_pin.py) -- they pin it to the physics it restates.
CPU only: the CUDA path is compared with this oracle in the -m gpu tests."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "oracle"))

G = 9.81


def _compile(tmp_path, xml, name="m.xml"):
    from myosuite_mjx_b200 import mjcf
    p = tmp_path / name
    p.write_text(xml)
    return mjcf.from_xml_path(str(p))


BALL = """<mujoco><option timestep="0.002"/><worldbody>
  <geom name="floor" type="plane" size="1 1 0.1"/>
  <body name="ball" pos="0 0 0.1"><freejoint/><geom name="ball" type="{type}" size="{size}" density="1000"/></body>
</worldbody></mujoco>"""


# ------------------------------------------------------------------ compiled constants of primitive geoms (mj_setConst rules)
@pytest.mark.parametrize("gtype,size,mass,inertia", [
    ("sphere", "0.05", 1000 * 4 / 3 * np.pi * 0.05**3, None),
    ("box", "0.03 0.04 0.05", 1000 * 8 * 0.03 * 0.04 * 0.05, "box"),
    ("cylinder", "0.03 0.06", 1000 * np.pi * 0.03**2 * 0.12, "cyl"),
    ("ellipsoid", "0.03 0.04 0.05", 1000 * 4 / 3 * np.pi * 0.03 * 0.04 * 0.05, "ell"),
    ("capsule", "0.03 0.06", 1000 * (np.pi * 0.03**2 * 0.12 + 4 / 3 * np.pi * 0.03**3), "cap")])
def test_primitive_mass_inertia_invweight(tmp_path, gtype, size, mass, inertia):
    m = _compile(tmp_path, BALL.format(type=gtype, size=size))
    assert m.body_mass[1] == pytest.approx(mass, rel=1e-12)
    if inertia is None:
        I = np.full(3, 0.4 * mass * 0.05**2)
    elif inertia == "box":
        a, b, c = 0.03, 0.04, 0.05
        I = mass / 3 * np.array([b * b + c * c, a * a + c * c, a * a + b * b])
    elif inertia == "cyl":
        r, h = 0.03, 0.12
        I = np.array([mass * (3 * r * r + h * h) / 12] * 2 + [mass * r * r / 2])
    elif inertia == "ell":
        a, b, c = 0.03, 0.04, 0.05
        I = mass / 5 * np.array([b * b + c * c, a * a + c * c, a * a + b * b])
    else:   # capsule = cylinder + two hemispheres (parallel axis from the hemisphere's own centroid, 3r/8 above its base)
        r, hh = 0.03, 0.06
        mc, ms = 1000 * np.pi * r * r * 2 * hh, 1000 * 4 / 3 * np.pi * r**3
        Iz = mc * r * r / 2 + ms * 0.4 * r * r
        Ix = mc * (3 * r * r + 4 * hh * hh) / 12 + ms * (0.4 * r * r + hh * hh + 0.75 * hh * r)
        I = np.array([Ix, Ix, Iz])
    # the compiler may permute / realign principal axes: compare the sorted principal moments
    np.testing.assert_allclose(np.sort(m.body_inertia[1]), np.sort(I), rtol=1e-9)
    # a single free body: dof_invweight0 = 1/m for the translations; the rotational entries are the mean of 1/I
    np.testing.assert_allclose(m.dof_invweight0[:3], 1 / mass, rtol=1e-9)
    np.testing.assert_allclose(m.dof_invweight0[3:], np.mean(1 / m.body_inertia[1]), rtol=1e-9)
    np.testing.assert_allclose(m.body_invweight0[1], [1 / mass, np.mean(1 / m.body_inertia[1])], rtol=1e-9)


def test_elbow_acc0_is_norm_of_minv_moment(models):
    """actuator_acc0 = || M^-1 moment^T || at qpos0 (mj_setConst); for the 1-dof elbow that is |moment| / M."""
    import oracle
    m = models("elbow")
    o = oracle.Oracle(m)
    o.reset(); o.forward()
    M = float(np.asarray(o.qMdense)[0])
    mom = np.asarray(o.actuator_moment).reshape(m.nu, m.nv)[:, 0]
    np.testing.assert_allclose(np.asarray(m.actuator_acc0), np.abs(mom) / M, rtol=1e-6)
    assert M == pytest.approx(1 / float(m.dof_invweight0[0]), rel=1e-6)          # dof_invweight0 = (M^-1)_ii at qpos0
    lr = np.asarray(m.actuator_lengthrange)   # given in the XML for this model (the finger's are simulated: tests/test_lengthrange.py)
    assert np.all(lr[:, 0] < lr[:, 1])


# ------------------------------------------------------------------ constraint parameters: solref / solimp -> K, B, impedance, R, aref
def _impedance(solimp, x):
    d0, dw, width, mid, power = solimp
    y = min(1.0, abs(x) / width)
    if power == 1:
        s = y
    elif y <= mid:
        s = y**power / mid**(power - 1)
    else:
        s = 1 - (1 - y)**power / (1 - mid)**(power - 1)
    return d0 + s * (dw - d0)


def test_joint_limit_row_closed_forms(models):
    """One violated joint limit of the elbow: efc_pos, efc_D = 1/R with R = (1 - d)/d * diagApprox, efc_aref = -B v - K d (pos - margin)
    with K = 1/(dmax^2 tc^2 zeta^2), B = 2/(dmax tc) (MuJoCo "Solver parameters", positive solref)."""
    import oracle
    m = models("elbow")
    lo, hi = m.jnt_range[0]
    o = oracle.Oracle(m)
    for q, v in ((hi + 0.05, 0.3), (lo - 0.02, -0.7)):
        o.reset(); o.set_state(qpos=[q], qvel=[v]); o.forward()
        assert o.nefc == 1 and int(o.efc_type[0]) == 3
        dist = hi - q if q > hi else q - lo
        sign = -1.0 if q > hi else 1.0
        tc, zeta = m.jnt_solref[0]
        tc = max(tc, 2 * m.opt_timestep)                    # refsafe
        solimp = np.asarray(m.jnt_solimp[0], dtype=np.float64)
        d = _impedance(solimp, dist - m.jnt_margin[0])
        dmax = solimp[1]
        K, B = 1 / (dmax**2 * tc**2 * zeta**2), 2 / (dmax * tc)
        R = (1 - d) / d * float(m.dof_invweight0[0])
        assert float(o.efc_pos[0]) == pytest.approx(dist, abs=1e-12)
        assert float(o.efc_D[0]) == pytest.approx(1 / R, rel=1e-9)
        assert float(o.efc_aref[0]) == pytest.approx(-B * sign * v - K * d * (dist - m.jnt_margin[0]), rel=1e-9)
        assert float(np.asarray(o.efc_J)[0]) == sign


# ------------------------------------------------------------------ contact force at rest = weight; limit force balances gravity
def test_sphere_on_plane_rest_force_equals_weight(tmp_path):
    import oracle
    m = _compile(tmp_path, BALL.format(type="sphere", size="0.05"))
    o = oracle.Oracle(m)
    o.reset(); o.step(3000)
    w = float(m.body_mass[1]) * G
    assert o.ncon == 1 and o.nefc == 4 and np.array_equal(np.asarray(o.contact_geom), [0, 1]) and int(o.contact_dim[0]) == 3
    f = np.asarray(o.efc_force)
    assert f.sum() == pytest.approx(w, rel=1e-6) and np.ptp(f) < 1e-9 * w, "pyramid: normal force = sum of the row forces, symmetric at rest"
    assert np.abs(np.asarray(o.qvel)).max() < 1e-7 and abs(float(o.qpos[2]) - 0.05) < 1e-3 and float(o.qpos[2]) < 0.05
    np.testing.assert_allclose(np.asarray(o.contact_frame)[:3], [0, 0, 1], atol=1e-12)


def test_pendulum_rests_on_its_joint_limit(tmp_path):
    import oracle
    m = _compile(tmp_path, """<mujoco><compiler angle="radian"/><option timestep="0.002"/><worldbody>
      <body pos="0 0 1"><joint name="h" type="hinge" axis="0 1 0" limited="true" range="-0.5 0.5" damping="0.05"/>
        <geom type="capsule" fromto="0 0 0 0.3 0 0" size="0.02" density="1000"/></body>
    </worldbody></mujoco>""")
    o = oracle.Oracle(m)
    o.reset(); o.step(6000)           # the arm falls (rotation about +y lowers +x) until the limit holds it
    q = float(o.qpos[0])
    assert abs(q) > 0.5 and abs(q) - 0.5 < 5e-3 and abs(float(o.qvel[0])) < 1e-6
    assert o.nefc == 1
    np.testing.assert_allclose(np.asarray(o.qfrc_constraint), -np.asarray(o.qfrc_smooth), rtol=1e-6)
    # gravity torque about the hinge = m g * (lever of the centre of mass)
    mass = float(m.body_mass[1])
    assert abs(float(o.qfrc_smooth[0])) == pytest.approx(mass * G * 0.15 * np.cos(q), rel=1e-3)


# ------------------------------------------------------------------ narrow phase: closed-form and convex (MPR) against analytic answers
PAIR = """<mujoco><option timestep="0.002" gravity="0 0 0"/><worldbody>
  <body pos="0 0 0"><freejoint/><geom name="a" type="{t1}" size="{s1}"/></body>
  <body pos="0 0 1"><freejoint/><geom name="b" type="{t2}" size="{s2}" margin="0.5"/></body>
</worldbody></mujoco>"""


def _quat(axis, ang):
    axis = np.asarray(axis, dtype=float) / np.linalg.norm(axis)
    return np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * axis])


def test_capsule_capsule_distance_vs_brute_force(tmp_path):
    import oracle
    m = _compile(tmp_path, PAIR.format(t1="capsule", s1="0.02 0.1", t2="capsule", s2="0.03 0.15"))
    o = oracle.Oracle(m)
    rng = np.random.default_rng(0)
    s = np.linspace(-1, 1, 801)
    hits = 0
    for _ in range(40):
        p1, p2 = rng.uniform(-0.1, 0.1, 3), rng.uniform(-0.1, 0.1, 3) + [0, 0, 0.12]
        q1, q2 = _quat(rng.normal(size=3), rng.uniform(0, np.pi)), _quat(rng.normal(size=3), rng.uniform(0, np.pi))
        o.reset(); o.set_state(qpos=np.concatenate([p1, q1, p2, q2])); o.forward()
        R = np.asarray(o.geom_xmat).reshape(2, 3, 3)
        a = p1[None] + s[:, None] * 0.1 * R[0][:, 2][None]
        b = p2[None] + s[:, None] * 0.15 * R[1][:, 2][None]
        dist = np.sqrt(((a[:, None, :] - b[None, :, :])**2).sum(-1)).min() - 0.05
        if dist < 0.5:
            assert o.ncon == 1
            assert float(o.contact_dist[0]) == pytest.approx(dist, abs=2e-5)     # sampling resolution of the brute force
            hits += 1
    assert hits >= 30


def test_mpr_sphere_ellipsoid_on_axis(tmp_path):
    """Convex narrow phase: a sphere pushed into an ellipsoid along a principal axis has depth c + r - h, normal along the axis."""
    import oracle
    m = _compile(tmp_path, PAIR.format(t1="ellipsoid", s1="0.03 0.04 0.05", t2="sphere", s2="0.02").replace('margin="0.5"', ""))
    o = oracle.Oracle(m)
    for axis, semi in ((0, 0.03), (1, 0.04), (2, 0.05)):
        for h in (semi + 0.02 - 0.004, semi + 0.02 - 0.012):
            p2 = np.zeros(3); p2[axis] = h
            o.reset(); o.set_state(qpos=np.concatenate([[0, 0, 0, 1, 0, 0, 0], p2, [1, 0, 0, 0]])); o.forward()
            assert o.ncon == 1
            assert float(o.contact_dist[0]) == pytest.approx(h - semi - 0.02, abs=2e-6)
            n = np.asarray(o.contact_frame)[:3]
            assert abs(n[axis]) == pytest.approx(1.0, abs=1e-5)
            expected = np.zeros(3); expected[axis] = 0.5 * (semi + (h - 0.02))     # midway between the two surfaces
            np.testing.assert_allclose(np.asarray(o.contact_pos)[:3], expected, atol=2e-5)
