"""The analytic known-answer cases of tests/test_oracle_kat2.py on the CUDA path itself (tiny MJCF models compiled at test
time by the package's own compiler): contact force at rest = weight, convex narrow phase against the closed form,
capsule-capsule distances against brute force, a joint limit holding a pendulum.  float32 tolerances stated per case."""
import numpy as np
import pytest

import test_oracle_kat2 as K

pytestmark = pytest.mark.gpu


def _batch(m, qpos, fields="all"):
    from myosuite_mjx_b200 import mjx
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=qpos), nworld=len(qpos), fields=fields)
    return model, d


def test_sphere_rests_with_its_weight_on_the_plane(tmp_path):
    import torch
    from myosuite_mjx_b200 import mjx
    m = K._compile(tmp_path, K.BALL.format(type="sphere", size="0.05"))
    nw = 40   # a partial tile; each world dropped from its own height
    q = np.tile(np.asarray(m.qpos0, dtype=np.float64), (nw, 1)); q[:, 2] = 0.05 + np.linspace(0.0, 0.05, nw)
    model, d = _batch(m, q)
    for _ in range(8):
        mjx.step(model, d, 500)
    torch.cuda.synchronize()
    w = float(m.body_mass[1]) * K.G
    f = d.efc_force[:, :4].double().cpu().numpy()
    assert (d.ncon[:, 0] == 1).all() and (d.nefc[:, 0] == 4).all() and int(d.status.abs().max()) == 0
    np.testing.assert_allclose(f.sum(axis=1), w, rtol=2e-4)          # float32 solve
    assert np.ptp(f, axis=1).max() < 1e-3 * w
    assert float(d.qvel.abs().max()) < 1e-4 and (d.qpos[:, 2] < 0.05).all() and (d.qpos[:, 2] > 0.049).all()
    assert torch.equal(d.contact_geom[:, 0], torch.tensor([[0, 1]], dtype=torch.int32, device="cuda").expand(nw, -1))


def test_convex_narrow_phase_matches_closed_form(tmp_path):
    import torch
    from myosuite_mjx_b200 import mjx
    m = K._compile(tmp_path, K.PAIR.format(t1="ellipsoid", s1="0.03 0.04 0.05", t2="sphere", s2="0.02").replace('margin="0.5"', ""))
    cases = [(axis, semi, semi + 0.02 - pen) for axis, semi in ((0, 0.03), (1, 0.04), (2, 0.05)) for pen in (0.004, 0.012)]
    q = np.zeros((len(cases), 14)); q[:, 3] = 1; q[:, 10] = 1
    for i, (axis, semi, h) in enumerate(cases):
        q[i, 7 + axis] = h
    model, d = _batch(m, q)
    mjx.forward(model, d)
    torch.cuda.synchronize()
    assert (d.ncon[:, 0] == 1).all()
    for i, (axis, semi, h) in enumerate(cases):
        assert float(d.contact_dist[i, 0]) == pytest.approx(h - semi - 0.02, abs=5e-6)      # float32 poses into a float64 MPR
        assert abs(float(d.contact_frame[i, 0, axis])) == pytest.approx(1.0, abs=1e-5)
        assert float(d.contact_pos[i, 0, axis]) == pytest.approx(0.5 * (semi + h - 0.02), abs=2e-5)


def test_capsule_capsule_distance_vs_brute_force(tmp_path):
    import torch
    from myosuite_mjx_b200 import mjx
    m = K._compile(tmp_path, K.PAIR.format(t1="capsule", s1="0.02 0.1", t2="capsule", s2="0.03 0.15"))
    rng = np.random.default_rng(0)
    nw = 64
    q = np.zeros((nw, 14))
    for w in range(nw):
        q[w, :3] = rng.uniform(-0.1, 0.1, 3); q[w, 7:10] = rng.uniform(-0.1, 0.1, 3) + [0, 0, 0.12]
        q[w, 3:7] = K._quat(rng.normal(size=3), rng.uniform(0, np.pi)); q[w, 10:14] = K._quat(rng.normal(size=3), rng.uniform(0, np.pi))
    model, d = _batch(m, q)
    mjx.forward(model, d)
    torch.cuda.synchronize()
    R = d.geom_xmat.double().cpu().numpy().reshape(nw, 2, 3, 3)
    s = np.linspace(-1, 1, 801)
    for w in range(nw):
        a = q[w, :3][None] + s[:, None] * 0.1 * R[w, 0][:, 2][None]
        b = q[w, 7:10][None] + s[:, None] * 0.15 * R[w, 1][:, 2][None]
        dist = np.sqrt(((a[:, None, :] - b[None, :, :])**2).sum(-1)).min() - 0.05
        assert int(d.ncon[w, 0]) == 1 and float(d.contact_dist[w, 0]) == pytest.approx(dist, abs=3e-5)


def test_pendulum_rests_on_its_joint_limit(tmp_path):
    import torch
    from myosuite_mjx_b200 import mjx
    m = K._compile(tmp_path, """<mujoco><compiler angle="radian"/><option timestep="0.002"/><worldbody>
      <body pos="0 0 1"><joint name="h" type="hinge" axis="0 1 0" limited="true" range="-0.5 0.5" damping="0.05"/>
        <geom type="capsule" fromto="0 0 0 0.3 0 0" size="0.02" density="1000"/></body>
    </worldbody></mujoco>""")
    model, d = _batch(m, np.zeros((33, 1)))
    for _ in range(12):
        mjx.step(model, d, 500)
    torch.cuda.synchronize()
    q = d.qpos[:, 0].double().cpu().numpy()
    assert (np.abs(q) > 0.5).all() and (np.abs(q) - 0.5 < 5e-3).all() and float(d.qvel.abs().max()) < 1e-5
    np.testing.assert_allclose(d.qfrc_constraint.cpu().numpy(), -d.qfrc_smooth.cpu().numpy(), rtol=1e-4)
    assert float(d.qfrc_smooth[0, 0].abs()) == pytest.approx(float(m.body_mass[1]) * K.G * 0.15 * np.cos(q[0]), rel=1e-3)
