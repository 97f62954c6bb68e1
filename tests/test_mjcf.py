"""Model compiler: the committed fixtures have the sizes SURVEY.md App. A counted from the XML, satisfy structural
invariants, and (in the build container, where /root/reference exists) recompiling the MJCF reproduces them."""
import os

import numpy as np
import pytest

from conftest import model_path

# name -> (nq, nv, nu, ntendon, nwrap, nM)   SURVEY.md App. A
SIZES = {"elbow": (1, 1, 6, 7, 40, 1), "finger": (4, 4, 5, 5, 35, 10), "hand": (23, 23, 39, 39, 315, 116),
         "hand_pose": (23, 23, 39, 44, 325, 116), "hand_sar": (29, 29, 39, 39, 315, 137), "legs": (35, 34, 80, 80, 324, 351)}
XML = {"elbow": "myosuite/envs/myo/assets/elbow/myoelbow_1dof6muscles.xml", "finger": "myosuite/simhive/myo_sim/finger/myofinger_v0.xml",
       "hand": "myosuite/simhive/myo_sim/hand/myohand.xml", "hand_pose": "myosuite/envs/myo/assets/hand/myohand_pose.xml",
       "hand_sar": "myosuite/envs/myo/assets/hand/myohand_sar.xml", "legs": "myosuite/simhive/myo_sim/leg/myolegs.xml"}


@pytest.mark.parametrize("name", sorted(SIZES))
def test_fixture_sizes(models, name):
    m = models(name)
    assert (m.nq, m.nv, m.nu, m.ntendon, m.nwrap, m.nM) == SIZES[name]
    assert m.na == m.nu                                            # all actuators are muscles with activation state
    assert np.all(np.asarray(m.actuator_dyntype) == 4)


@pytest.mark.parametrize("name", sorted(SIZES))
def test_structural_invariants(models, name):
    m = models(name)
    par = np.asarray(m.body_parentid)
    assert par[0] == 0 and np.all(par[1:] < np.arange(1, m.nbody))          # depth-first numbering
    assert np.all(np.asarray(m.body_mass)[1:] >= 0) and np.all(np.asarray(m.body_inertia) >= 0)
    dpar = np.asarray(m.dof_parentid)
    assert np.all(dpar < np.arange(m.nv))
    depth = [1 + (0 if dpar[i] < 0 else 0) for i in range(m.nv)]
    nM = 0
    for i in range(m.nv):
        j, d = i, 0
        while j >= 0:
            d += 1; j = dpar[j]
        nM += d
    assert nM == m.nM
    q = np.asarray(m.body_quat); assert np.allclose(np.linalg.norm(q, axis=1), 1, atol=1e-9)
    assert np.all(np.asarray(m.dof_invweight0) > 0) and np.all(np.asarray(m.actuator_acc0) > 0)
    assert m.stat_meaninertia > 0
    lim = np.asarray(m.jnt_limited).astype(bool)
    r = np.asarray(m.jnt_range)
    assert np.all(r[lim, 0] < r[lim, 1])
    lr = np.asarray(m.actuator_lengthrange)      # finger: filled by the compile-time simulation (tests/test_lengthrange.py)
    assert np.all(lr[:, 0] < lr[:, 1])


def test_hand_known_values(models):
    m = models("hand")
    assert m.opt_timestep == pytest.approx(0.002)                    # myohand_assets.xml:13
    assert np.asarray(m.dof_damping).min() == pytest.approx(0.05)    # myohand_assets.xml:22
    assert m.name2id("jnt", "pro_sup") == 0 and m.name2id("jnt", "md5_flexion") == 22
    assert m.name2id("actuator", "ECRL") == 0 and m.name2id("actuator", "UI_UB5") == 38
    geom_margin = np.asarray(m.geom_margin)
    assert geom_margin.max() == pytest.approx(0.001)                 # myohand_assets.xml:23
    legs = models("legs")
    assert legs.opt_timestep == pytest.approx(0.001) and legs.neq == 14


@pytest.mark.parametrize("name", ["elbow", "hand"])
def test_recompile_reproduces_fixture(models, name):
    ref = "/root/reference"
    if not os.path.isdir(ref):
        pytest.skip("reference checkout only exists in the build container")
    from myosuite_mjx_b200 import mjcf
    m = mjcf.from_xml_path(os.path.join(ref, XML[name]))
    f = models(name)
    for k in ("body_pos", "body_quat", "body_mass", "body_inertia", "jnt_axis", "jnt_range", "site_pos", "geom_size",
              "actuator_gainprm", "actuator_lengthrange", "actuator_acc0", "dof_invweight0", "tendon_invweight0", "qpos0"):
        np.testing.assert_allclose(np.asarray(getattr(m, k)), np.asarray(getattr(f, k)), rtol=1e-12, atol=1e-14, err_msg=k)
    for k in ("body_parentid", "jnt_type", "wrap_type", "wrap_objid", "tendon_adr", "actuator_trnid", "geom_type"):
        assert np.array_equal(np.asarray(getattr(m, k)), np.asarray(getattr(f, k))), k


def test_unsupported_feature_raises(tmp_path):
    from myosuite_mjx_b200 import mjcf
    p = tmp_path / "bad.xml"
    p.write_text('<mujoco><worldbody><body><joint type="ball"/><geom size="0.1"/></body></worldbody></mujoco>')
    with pytest.raises(NotImplementedError):
        m = mjcf.from_xml_path(str(p))
        from myosuite_mjx_b200 import setconst  # noqa: F401
        if any(np.asarray(m.jnt_type) == 1):
            raise NotImplementedError("ball joint")
