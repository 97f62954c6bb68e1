"""Pins oracle/env_oracle.py (numpy restatement of the reference's pose-env layer) to golden vectors produced by the
reference's OWN code (tests/golden/make_env_kat.py ran BaseV0.step, CumulativeFatigue.compute_act,
PoseEnvV0.get_obs_dict / get_reward_dict from /root/reference).  CPU only."""
import os

import numpy as np
import pytest

import env_oracle as eo

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_kat.npz"))
IS_MUSCLE = G["dyntype"] == 4
NMUS = int(IS_MUSCLE.sum())


@pytest.mark.parametrize("cond", ["none", "sarcopenia", "reafferentation"])
def test_action_to_ctrl_matches_reference(cond):
    a = G["actions"]
    c = eo.action_to_ctrl(a, IS_MUSCLE, NMUS, True, None, None, "" if cond == "none" else cond, epl=7, eip=11)
    np.testing.assert_allclose(c, G["ctrl_" + cond], rtol=0, atol=1e-15)


def test_action_passthrough_variants():
    a = G["actions"][0]
    assert not bool(G["ctrl_raw_isnormalized"])          # normalize_act False: raw action, robot told not to re-project
    np.testing.assert_array_equal(eo.action_to_ctrl(a, IS_MUSCLE, NMUS, False, np.zeros(len(a)), np.ones(len(a))), G["ctrl_raw"])
    assert bool(G["ctrl_na0_isnormalized"])              # no activations: the robot re-projects by ctrlrange (robot.py:771-783)
    np.testing.assert_array_equal(G["ctrl_na0"], a)
    lo, hi = np.full(len(a), -2.0), np.full(len(a), 4.0)
    np.testing.assert_allclose(eo.action_to_ctrl(a, IS_MUSCLE, 0, True, lo, hi), 1.0 + 3.0 * a, atol=1e-15)


def test_fatigue_matches_reference():
    dyn = G["dynprm"]
    f = eo.Fatigue(dyn[IS_MUSCLE, 0], dyn[IS_MUSCLE, 1], float(G["timestep"]) * int(G["frame_skip"]), (), float(G["fatigue_F"]),
                   float(G["fatigue_R"]), float(G["fatigue_r"]))
    for t in range(len(G["actions"])):
        c = eo.action_to_ctrl(G["actions"][t], IS_MUSCLE, NMUS, True, None, None, "fatigue", fatigue=f)
        np.testing.assert_allclose(c, G["ctrl_fatigue"][t], rtol=0, atol=1e-14)
        np.testing.assert_allclose(np.stack([f.MA, f.MR, f.MF]), G["fatigue_states"][t], rtol=0, atol=1e-14)
    assert np.all(f.MA + f.MR + f.MF == pytest.approx(1.0, abs=1e-9))    # the three compartments partition the muscle


def test_obs_and_reward_match_reference():
    dt = float(G["timestep"]) * int(G["frame_skip"])
    od = eo.obs_dict(G["pose_time"], G["pose_qpos"], G["pose_qvel"], G["pose_act"], G["pose_target"], dt)
    for k in ("time", "qpos", "qvel", "act", "pose_err"):
        np.testing.assert_array_equal(od[k], G["obs_" + k])
    r = eo.reward_dict(od, G["pose_act"].shape[1], float(G["pose_thd"]), tuple(G["pose_weights"]))
    for k in ("pose", "bonus", "penalty", "act_reg", "sparse", "dense"):
        np.testing.assert_allclose(r[k], G["rwd_" + k], rtol=1e-15, atol=1e-15)
    np.testing.assert_array_equal(r["solved"], G["rwd_solved"].astype(bool))
    np.testing.assert_array_equal(r["done"], G["rwd_done"].astype(bool))
    assert r["done"].sum() >= 10 and r["solved"].sum() >= 50 and (r["bonus"] == 1).sum() >= 5    # every branch is exercised
    assert list(G["default_obs_keys"]) == ["qpos", "qvel", "pose_err"]


def test_uniform_hash():
    u = np.array([[eo.uniform(1234, w, e, s, i) for i in range(23)] for w in range(40) for e in range(3) for s in range(2)])
    assert u.dtype == np.float32 and u.min() >= 0.0 and u.max() < 1.0
    assert abs(u.mean() - 0.5) < 0.02 and abs(u.var() - 1 / 12) < 0.01
    assert len(np.unique(u)) > 0.99 * u.size
    assert eo.uniform(1, 2, 3, 0, 4) == eo.uniform(1, 2, 3, 0, 4) != eo.uniform(1, 2, 3, 1, 4)


def test_env_oracle_auto_reset_and_partition():
    """Worlds reset on the horizon with fresh targets; a world's stream depends on its GLOBAL index only."""
    nq = 4
    cfg = dict(frame_skip=2, normalize_act=True, cond="", obs_keys=["qpos", "qvel", "pose_err", "act"], target_qposadr=np.arange(nq),
               target_type="generate", target_range=np.tile([[-1.0, 1.0]], (nq, 1)), target_value=None, reset_type="random", auto_reset=True,
               horizon=3, seed=99, pose_thd=0.35, weights=(1.0, 4.0, 1.0, 50.0), init_qpos=np.zeros(nq), init_qvel=np.zeros(nq),
               reset_lo=-np.ones(nq), reset_hi=np.ones(nq), act_is_muscle=np.ones(2, int), ctrl_lo=np.zeros(2), ctrl_hi=np.ones(2), na=2, dt=0.004)

    def physics(qpos, qvel, act, ctrl, time, n):        # a stand-in integrator: the env logic is what is under test
        return qpos + 0.01 * n, qvel, 0.5 * act + 0.5 * ctrl, time + 0.002 * n
    full = eo.PoseEnvOracle(dict(cfg, world_offset=0), 6, physics)
    half = eo.PoseEnvOracle(dict(cfg, world_offset=3), 3, physics)
    o_full, o_half = full.reset(), half.reset()
    np.testing.assert_array_equal(o_full[3:], o_half)
    for t in range(7):
        a = np.full((6, 2), 0.3 + 0.1 * t)
        o_full, r_full, e_full = full.step(a)
        o_half, r_half, e_half = half.step(a[3:])
        np.testing.assert_array_equal(o_full[3:], o_half)
        assert e_full.all() == ((t + 1) % 3 == 0)
    assert (full.episode == 3).all() and (full.steps == 1).all()
    assert len(np.unique(full.target[:, 0])) == 6


def test_walk_obs_and_reward_match_reference(models):
    """oracle/env_oracle.walk_obs_reward against vectors produced by the reference's own WalkEnvV0.get_obs_dict /
    get_reward_dict (tests/golden/make_walk_kat.py)."""
    from myosuite_mjx_b200 import envs
    W = np.load(os.path.join(os.path.dirname(__file__), "golden", "walk_kat.npz"))
    c = envs.walk_config(models("legs"))
    assert [k for k in c["obs_keys"] if k != "act"] == list(W["default_obs_keys"])
    assert tuple(c["walk_weights"]) == tuple(W["weights"])
    f = lambda k: W[k].astype(np.float64)
    od, r = eo.walk_obs_reward(c, f("qpos"), f("qvel"), f("act"), f("cvel"), f("xipos"), f("xpos"), f("xquat"), f("actuator_length"),
                               f("actuator_velocity"), f("actuator_force"), W["steps"])
    for k in od:
        np.testing.assert_allclose(od[k], W["obs_" + k], rtol=1e-6, atol=1e-7, err_msg=k)   # float32 masses / keyframe vs the reference's float64
    for k in ("vel_reward", "cyclic_hip", "ref_rot", "joint_angle_rew", "act_mag", "sparse", "dense"):
        np.testing.assert_allclose(r[k], W["rwd_" + k], rtol=2e-6, atol=2e-6, err_msg=k)
    np.testing.assert_array_equal(r["done"], W["rwd_done"] > 0)
    np.testing.assert_array_equal(r["solved"], W["rwd_solved"] > 0)
    assert 0.2 < r["done"].mean() < 0.8


def test_reorient_obs_and_reward_match_reference(models):
    """oracle/env_oracle.reorient_obs_reward and euler2quat against vectors produced by the reference's own
    ProprioceptiveEnvV0.get_obs_dict / get_reward_dict and quat_math.euler2quat (tests/golden/make_reorient_kat.py)."""
    from myosuite_mjx_b200 import envs
    R = np.load(os.path.join(os.path.dirname(__file__), "golden", "reorient_kat.npz"))
    c = envs.reorient_config(models("hand_sar"))
    assert c["pen_length"] == pytest.approx(0.07) and c["tar_length"] == pytest.approx(0.07)
    assert sorted(c["obs_keys"]) == sorted(R["default_obs_keys"]) and tuple(c["reorient_weights"]) == tuple(R["weights"])
    f = lambda k: R[k].astype(np.float64)
    g = f("geom_xpos")
    od, r = eo.reorient_obs_reward(c, f("qpos"), f("qvel"), f("act"), f("xpos")[:, 0], f("site_xpos")[:, 0], g[:, 0], g[:, 1], (g[:, 2] - g[:, 3]) / c["tar_length"],
                                   f("actuator_length"), f("actuator_velocity"), f("actuator_force"))
    for k in od:
        np.testing.assert_allclose(od[k], R["obs_" + k], rtol=1e-12, atol=1e-12, err_msg=k)
    for k in ("pos_align", "rot_align", "act_reg", "drop", "bonus", "sparse", "dense"):
        np.testing.assert_allclose(r[k], R["rwd_" + k], rtol=1e-12, atol=1e-12, err_msg=k)
    np.testing.assert_array_equal(r["done"], R["rwd_done"] > 0)
    np.testing.assert_array_equal(r["solved"], R["rwd_solved"] > 0)
    assert 0.2 < r["done"].mean() < 0.8 and (r["bonus"] == 1).sum() > 3 and (r["bonus"] == 6).sum() > 3
    np.testing.assert_allclose(eo.euler2quat(R["euler"]), R["euler_quat"], rtol=0, atol=1e-15)
