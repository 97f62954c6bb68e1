"""The MJX-style env surface (reset(rng) -> State, step(State, action) -> State, pipeline_init / pipeline_step) on the
GPU, and size-independent properties at BASELINE.json's full batch size (8192 worlds)."""
import os

import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu


def test_pipeline_env_matches_oracle(models):
    """pipeline_init = reset + state write + forward; pipeline_step = n_frames mj_steps with ctrl = action."""
    import oracle
    import torch
    from myosuite_mjx_b200 import pipeline
    m = models("hand")
    nw = 8
    env = pipeline.PipelineEnv(m, n_frames=5, nworld=nw, fields="core")
    assert env.dt == pytest.approx(0.01) and env.action_size == 39
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 31, qpos_spread=0.4)
    s = env.pipeline_init(qpos, 0 * qvel)
    assert float(s.time.abs().max()) == 0 and float(s.act.abs().max()) == 0
    s = env.pipeline_step(s, ctrl)
    torch.cuda.synchronize()
    o = oracle.Oracle(m)
    for w in range(nw):
        o.reset(); o.set_state(qpos[w], 0 * qvel[w], 0 * act[w], ctrl[w]); o.step(5)
        assert np.abs(s.q[w].cpu().numpy() - o.qpos).max() < 2e-4
        assert np.abs(s.qd[w].cpu().numpy() - o.qvel).max() < 2e-2 * max(1.0, np.abs(o.qvel).max())
        assert float(s.time[w, 0]) == pytest.approx(0.01, abs=1e-6)


def test_pose_env_reset_step_shapes(models):
    import torch
    from myosuite_mjx_b200 import pipeline
    m = models("hand_pose")
    env = pipeline.PoseEnv(m, n_frames=10, nworld=64)
    st = env.reset(1234)
    # obs = qpos, qvel * dt, pose_err, act (pose_v0.py:15 + base_v0.py:34-38)
    assert st.obs.shape == (64, 2 * m.nq + m.nv + m.na) and st.reward.shape == (64,) and st.done.shape == (64,)
    assert set(st.metrics) == {"pose", "bonus", "penalty", "act_reg", "sparse", "solved", "done"}
    lo = torch.as_tensor(np.asarray(m.jnt_range)[:, 0], device="cuda", dtype=torch.float32)
    hi = torch.as_tensor(np.asarray(m.jnt_range)[:, 1], device="cuda", dtype=torch.float32)
    assert bool(((st.pipeline_state.qpos >= lo - 1e-6) & (st.pipeline_state.qpos <= hi + 1e-6)).all())
    assert bool(((env.target >= lo - 1e-6) & (env.target <= hi + 1e-6)).all())
    st2 = env.reset(1234)
    q_first = st2.pipeline_state.qpos.clone()
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    for _ in range(5):
        a = torch.rand((64, m.nu), generator=g, device="cuda") * 2 - 1
        st2 = env.step(st2, a)
    torch.cuda.synchronize()
    assert torch.isfinite(st2.obs).all() and torch.isfinite(st2.reward).all()
    assert float(st2.pipeline_state.time.min()) == pytest.approx(5 * 10 * 0.002, abs=1e-5)
    assert float((st2.pipeline_state.qpos - q_first).abs().max()) > 1e-3
    assert float(st2.pipeline_state.act.min()) >= 0 and float(st2.pipeline_state.act.max()) <= 1
    obs2 = st2.obs.clone()
    # same seed, same actions -> same trajectory (determinism, the property tests/test_envs.py:110-120 checks)
    st3 = env.reset(1234)
    g.manual_seed(7)
    for _ in range(5):
        st3 = env.step(st3, torch.rand((64, m.nu), generator=g, device="cuda") * 2 - 1)
    assert torch.equal(st3.obs, obs2)


GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_kat.npz"))


@pytest.mark.parametrize("cond", ["", "fatigue", "reafferentation"])
def test_env_pre_kernel_matches_reference_golden(models, cond):
    """action -> ctrl on the device against vectors produced by the reference's own BaseV0.step / CumulativeFatigue
    (tests/golden/make_env_kat.py).  float32 tolerance: 2e-6 absolute on ctrl in [0, 1]; fatigue compartments, which
    integrate 300 steps in float32 against the reference's float64, 2e-5."""
    import torch
    from myosuite_mjx_b200 import envs, mjx
    m = models("hand_pose")
    nu, nw = m.nu, 32
    assert nu == 39 and (np.asarray(m.actuator_dyntype) == 4).all()
    model = mjx.put_model(m)
    d = mjx.make_data(model, nw, fields="core")
    cfg = envs.pose_config(m, frame_skip=int(GOLD["frame_skip"]), muscle_condition=cond, epl_id=7, eip_id=11, auto_reset=False, horizon=0)
    cfg["tau_act"], cfg["tau_deact"] = GOLD["dynprm"][:nu, 0].astype(np.float32), GOLD["dynprm"][:nu, 1].astype(np.float32)
    env = envs.FusedEnv(d, cfg)
    env.reset()
    T = 300 if cond == "fatigue" else 40
    for t in range(T):
        a = torch.as_tensor(GOLD["actions"][t, :nu], dtype=torch.float32, device="cuda").expand(nw, nu)
        env.step(a)
        ctrl = d.ctrl.cpu().numpy()
        assert np.abs(ctrl - GOLD["ctrl_" + (cond or "none")][t, :nu]).max() < (2e-5 if cond == "fatigue" else 2e-6), (cond, t)
        if cond == "fatigue" and t % 50 == 49:
            f = env.storage("fatigue").cpu().numpy().reshape(3, nu, nw)
            assert np.abs(f - GOLD["fatigue_states"][t][:, :, None]).max() < 2e-5
    assert int(env.steps.min()) == T and int(env.episode.max()) == 1


def test_env_forward_matches_reference_golden(models):
    """obs / reward terms / done of given states against the reference's own PoseEnvV0.get_obs_dict / get_reward_dict.
    float32: obs exact to rounding of the inputs (1e-6 rel), reward terms rel 2e-6; the thresholded terms (bonus, solved,
    done, penalty) must agree exactly except where the float64 distance is within 1e-6 of a threshold."""
    import torch
    from myosuite_mjx_b200 import envs, mjx
    m = models("hand_pose")
    n = GOLD["pose_qpos"].shape[0]
    model = mjx.put_model(m)
    d = mjx.make_data(model, n, fields="core")
    cfg = envs.pose_config(m, obs_keys=("time", "qpos", "qvel", "act", "pose_err"), auto_reset=False, horizon=0)
    env = envs.FusedEnv(d, cfg)
    env.reset()
    dev = d.qpos.device
    d.qpos.copy_(torch.as_tensor(GOLD["pose_qpos"], dtype=torch.float32, device=dev)); d.qvel.copy_(torch.as_tensor(GOLD["pose_qvel"], dtype=torch.float32, device=dev))
    d.act.copy_(torch.as_tensor(GOLD["pose_act"], dtype=torch.float32, device=dev)); d.time.copy_(torch.as_tensor(GOLD["pose_time"][:, None], dtype=torch.float32, device=dev))
    env.target.copy_(torch.as_tensor(GOLD["pose_target"], dtype=torch.float32, device=dev))
    env.forward()
    obs = env.obs.cpu().numpy()
    ref = np.concatenate([GOLD["obs_" + k] for k in ("time", "qpos", "qvel", "act", "pose_err")], axis=1)
    assert obs.shape == ref.shape
    assert np.abs(obs - ref).max() < 2e-6 * max(1.0, np.abs(ref).max())
    terms = env.rwd_terms.cpu().numpy()
    dist = -GOLD["rwd_pose"]
    thd, far = float(GOLD["pose_thd"]), 2 * np.pi
    safe = (np.abs(dist - thd) > 1e-5) & (np.abs(dist - 1.5 * thd) > 1e-5) & (np.abs(dist - far) > 1e-5)
    assert safe.sum() > 0.95 * n
    for i, k in enumerate(envs.RWD_TERMS):
        g = GOLD["rwd_" + k].astype(np.float64)
        sel = safe if k in ("bonus", "penalty", "solved", "done") else slice(None)
        assert np.abs(terms[:, i][sel] - g[sel]).max() <= 2e-6 * max(1.0, np.abs(g).max()), k
    assert np.abs(env.reward.cpu().numpy()[:, 0][safe] - GOLD["rwd_dense"][safe]).max() < 1e-4      # weights up to 50
    assert np.array_equal(env.done.cpu().numpy()[:, 0][safe] > 0, GOLD["rwd_done"][safe] > 0)
    assert int(env.steps.max()) == 0                                                              # forward does not count as a step


def test_env_rollout_matches_oracle(models):
    """Full env loop (reset -> step x N with auto-reset on a short horizon) against oracle/env_oracle.PoseEnvOracle driving
    the mj_step oracle: same targets and reset poses bit for bit (shared counter-based hash), obs within the physics
    tolerance of tests/test_parity_gpu.py accumulated over frame_skip steps."""
    import env_oracle as eo
    import oracle
    import torch
    from myosuite_mjx_b200 import envs, mjx
    m = models("finger")
    nw, horizon, fs = 6, 4, 5
    model = mjx.put_model(m)
    d = mjx.make_data(model, nw, fields="core")
    base = envs.pose_config(m)
    tr = np.stack([base["reset_lo"], base["reset_hi"]], axis=1) * 0.6
    cfg = envs.pose_config(m, frame_skip=fs, target_jnt_range=tr, reset_type="random", horizon=horizon, auto_reset=True, seed=77, world_offset=10)
    cfg["reset_lo"], cfg["reset_hi"] = 0.5 * cfg["reset_lo"], 0.5 * cfg["reset_hi"]
    env = envs.FusedEnv(d, cfg)
    orcs = [oracle.Oracle(m) for _ in range(nw)]

    def physics(qpos, qvel, act, ctrl, time, n):
        q2, v2, a2, t2 = qpos.copy(), qvel.copy(), act.copy(), time.copy()
        for w, o in enumerate(orcs):
            if time[w] == 0:
                o.reset()
            o.set_state(qpos[w], qvel[w], act[w], ctrl[w]); o.field("time")[:] = time[w]
            o.step(n)
            q2[w], v2[w], a2[w], t2[w] = o.qpos, o.qvel, o.act, o.field("time")[0]
        return q2, v2, a2, t2
    ref = eo.PoseEnvOracle(cfg, nw, physics)
    env.reset(); obs_ref = ref.reset()
    torch.cuda.synchronize()
    assert np.array_equal(env.target.cpu().numpy(), ref.target.astype(np.float32))
    assert np.array_equal(d.qpos.cpu().numpy(), ref.qpos.astype(np.float32))
    assert np.abs(env.obs.cpu().numpy() - obs_ref).max() < 1e-6
    rng = np.random.default_rng(5)
    nreset = 0
    for t in range(2 * horizon + 1):
        a = rng.uniform(-1, 1, size=(nw, m.nu))
        env.step(torch.as_tensor(a, dtype=torch.float32, device="cuda"))
        obs_ref, r, ended = ref.step(a)
        got = env.obs.cpu().numpy()
        assert np.array_equal(env.done.cpu().numpy()[:, 0] > 0, ended), t
        assert np.array_equal(env.target.cpu().numpy(), ref.target.astype(np.float32))
        assert np.array_equal(env.episode.cpu().numpy()[:, 0], ref.episode) and np.array_equal(env.steps.cpu().numpy()[:, 0], ref.steps)
        assert np.abs(got - obs_ref).max() < 5e-3, (t, np.abs(got - obs_ref).max())
        assert np.abs(env.reward.cpu().numpy()[:, 0] - r["dense"]).max() < 2e-2
        nreset += int(ended.sum())
        # the oracle restarts each env step from the library's state so that float32 drift does not accumulate across steps
        ref.qpos, ref.qvel, ref.act = d.qpos.cpu().numpy().astype(np.float64), d.qvel.cpu().numpy().astype(np.float64), d.act.cpu().numpy().astype(np.float64)
    assert nreset == 2 * nw


def test_env_world_offset_partition(models):
    """A world's episode stream depends on its global index only: ranks that own [0, 4) and [4, 8) reproduce a single
    8-world batch (the multi-GPU partition of DESIGN.md section 4 applied to the env layer)."""
    import torch
    from myosuite_mjx_b200 import pipeline
    m = models("finger")
    full = pipeline.PoseEnv(m, n_frames=5, nworld=8, horizon=3)
    lo, hi = pipeline.PoseEnv(m, n_frames=5, nworld=4, horizon=3), pipeline.PoseEnv(m, n_frames=5, nworld=4, horizon=3, world_offset=4)
    sf, sl, sh = full.reset(3), lo.reset(3), hi.reset(3)
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    for _ in range(7):
        a = torch.rand((8, m.nu), generator=g, device="cuda") * 2 - 1
        sf, sl, sh = full.step(sf, a), lo.step(sl, a[:4]), hi.step(sh, a[4:])
        assert torch.equal(sf.obs[:4], sl.obs) and torch.equal(sf.obs[4:], sh.obs)
        assert torch.equal(sf.reward[4:], sh.reward) and torch.equal(sf.done[:4], sl.done)
    assert int(full.env.episode.min()) == 3


def test_walk_forward_matches_reference_golden(models):
    """Walk obs / reward terms / done of given mjData fields against the reference's own WalkEnvV0.get_obs_dict /
    get_reward_dict (tests/golden/make_walk_kat.py).  float32 tolerance: obs 2e-6 relative to the largest entry of each key
    (inputs are float32-exact), reward terms 2e-5 (exp of float32 sums), dense 2e-3 (weights up to 100)."""
    import torch
    from myosuite_mjx_b200 import envs, pipeline
    W = np.load(os.path.join(os.path.dirname(__file__), "golden", "walk_kat.npz"))
    m = models("legs")
    n = W["qpos"].shape[0]
    env = pipeline.WalkEnv(m, nworld=n, auto_reset=False, horizon=0)
    env.reset(0)
    d, e = env._data, env.env
    put = lambda t, a: t.copy_(torch.as_tensor(np.asarray(a, dtype=np.float32).reshape(n, -1), device=t.device).reshape(t.shape))
    put(d.qpos, W["qpos"]); put(d.qvel, W["qvel"]); put(d.act, W["act"]); put(d.cvel, W["cvel"]); put(d.xipos, W["xipos"]); put(d.xpos, W["xpos"])
    put(d.xquat, W["xquat"]); put(d.actuator_length, W["actuator_length"]); put(d.actuator_velocity, W["actuator_velocity"])
    put(d.actuator_force, W["actuator_force"])
    e.steps.copy_(torch.as_tensor(W["steps"][:, None] + 1, dtype=torch.int32, device="cuda"))     # the reference's self.steps lags by one
    e.forward()
    obs = e.obs.cpu().numpy()
    o = 0
    for k in env.cfg["obs_keys"]:
        ref = W["obs_" + k]
        got = obs[:, o:o + ref.shape[1]]; o += ref.shape[1]
        assert np.abs(got - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max()), k
    assert o == obs.shape[1]
    terms = e.rwd_terms.cpu().numpy()
    height = W["obs_height"][:, 0]
    safe = np.abs(height - 0.8) > 1e-5
    for i, k in enumerate(envs.WALK_RWD_TERMS):
        g = W["rwd_" + k].astype(np.float64)
        sel = safe if k in ("done", "solved") else slice(None)
        assert np.abs(terms[:, i][sel] - g[sel]).max() <= 2e-5 * max(1.0, np.abs(g).max()), k
    assert np.abs(e.reward.cpu().numpy()[:, 0][safe] - W["rwd_dense"][safe]).max() < 2e-3
    assert np.array_equal(e.done.cpu().numpy()[:, 0][safe] > 0, W["rwd_done"][safe] > 0)


def test_walk_env_step_semantics(models):
    """myoLegWalk-v0 through the registry: (1) after a step the observation's position / velocity entries describe the
    STEPPED state (classic-path semantics, dm_control's trailing mj_step1) -- checked against a full mj_forward of a copy
    of the state; (2) actuator_force keeps the last step's value; (3) worlds that fall are reset to the keyframe and
    their observation is the keyframe's; (4) the whole observation equals the oracle's walk_obs_reward of the bound fields."""
    import env_oracle as eo
    import torch
    from myosuite_mjx_b200 import mjx, registry
    m = models("legs")
    nw = 16
    env = registry.make("myoLegWalk-v0", nworld=nw, model=m, horizon=12)
    st = env.reset(5)
    d, e, c = env._data, env.env, env.cfg
    obs0 = st.obs.clone()
    assert obs0.shape == (nw, sum({"qpos_without_xy": 33, "qvel": 34, "com_vel": 2, "torso_angle": 4, "feet_heights": 2, "height": 1,
                                   "feet_rel_positions": 6, "phase_var": 1, "muscle_length": 80, "muscle_velocity": 80, "muscle_force": 80, "act": 80}[k] for k in c["obs_keys"]))
    key2 = np.asarray(m.key_qpos, dtype=np.float32)[2]
    assert np.array_equal(d.qpos.cpu().numpy(), np.tile(key2, (nw, 1)))
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    saw_reset = False
    for t in range(14):
        a = torch.rand((nw, m.nu), generator=g, device="cuda") * 2 - 1
        frc_before = None
        st = env.step(st, a)
        ended = st.done.cpu().numpy() > 0
        f = lambda name: getattr(d, name).cpu().numpy().astype(np.float64)
        steps_ref = np.maximum(e.steps.cpu().numpy()[:, 0] - 1, 0)
        od, r = eo.walk_obs_reward(c, f("qpos"), f("qvel"), f("act"), f("cvel").reshape(nw, -1, 6), f("xipos").reshape(nw, -1, 3),
                                   f("xpos").reshape(nw, -1, 3), f("xquat").reshape(nw, -1, 4), f("actuator_length"), f("actuator_velocity"),
                                   f("actuator_force"), steps_ref)
        ref = np.concatenate([od[k] for k in c["obs_keys"]], axis=-1)
        assert np.abs(st.obs.cpu().numpy() - ref).max() < 1e-4 * max(1.0, np.abs(ref).max()), t
        # (1) refreshed fields == full forward of the same state
        chk = d.clone()
        mjx.forward(env.sys, chk)
        for name in ("xpos", "xipos", "xquat", "cvel", "actuator_length", "actuator_velocity"):
            a_, b_ = getattr(d, name).cpu().numpy(), getattr(chk, name).cpu().numpy()
            assert np.abs(a_ - b_).max() <= 1e-5 * max(1.0, np.abs(b_).max()), (t, name)
        if ended.any():
            saw_reset = True
            assert np.array_equal(d.qpos.cpu().numpy()[ended], np.tile(key2, (int(ended.sum()), 1)))
            assert np.abs(st.obs.cpu().numpy()[ended] - obs0.cpu().numpy()[ended]).max() < 1e-6
            assert (e.steps.cpu().numpy()[ended, 0] == 0).all()
        else:
            # (2) the force entries are the last mj_step's, not the refreshed state's
            assert np.abs(d.actuator_force.cpu().numpy() - chk.actuator_force.cpu().numpy()).max() > 1e-3
    assert saw_reset and torch.isfinite(st.obs).all() and torch.isfinite(st.reward).all()


def test_reorient_forward_matches_reference_golden(models):
    """Reorient obs / reward terms / done of given mjData fields against the reference's own ProprioceptiveEnvV0 code
    (tests/golden/make_reorient_kat.py).  float32: obs 2e-6 relative per key, reward terms 1e-5 (rot_align is a ratio of
    float32 sums), thresholded terms exact away from the 0.075 / 0.9 / 0.95 thresholds."""
    import torch
    from myosuite_mjx_b200 import envs, pipeline
    R = np.load(os.path.join(os.path.dirname(__file__), "golden", "reorient_kat.npz"))
    m = models("hand_sar")
    n = R["qpos"].shape[0]
    env = pipeline.ReorientEnv(m, nworld=n, auto_reset=False, horizon=0, randomize_target=False)
    env.reset(0)
    d, e, c = env._data, env.env, env.cfg
    dev = d.qpos.device
    T = lambda a: torch.as_tensor(np.asarray(a, dtype=np.float32), device=dev)
    d.qpos.copy_(T(R["qpos"])); d.qvel.copy_(T(R["qvel"])); d.act.copy_(T(R["act"]))
    d.actuator_length.copy_(T(R["actuator_length"])); d.actuator_velocity.copy_(T(R["actuator_velocity"])); d.actuator_force.copy_(T(R["actuator_force"]))
    d.xpos[:, c["obj_bid"]] = T(R["xpos"][:, 0]); d.site_xpos[:, c["eps_ball_sid"]] = T(R["site_xpos"][:, 0])
    d.geom_xpos[:, c["top_gid"]] = T(R["geom_xpos"][:, 0]); d.geom_xpos[:, c["bot_gid"]] = T(R["geom_xpos"][:, 1])
    e.target.copy_(T((R["geom_xpos"][:, 2].astype(np.float64) - R["geom_xpos"][:, 3]) / c["tar_length"]))
    e.forward()
    obs = e.obs.cpu().numpy()
    o = 0
    for k in c["obs_keys"]:
        ref = R["obs_" + k]
        got = obs[:, o:o + ref.shape[1]]; o += ref.shape[1]
        assert np.abs(got - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max()), k
    assert o == obs.shape[1]
    terms = e.rwd_terms.cpu().numpy()
    pos, rot = -R["rwd_pos_align"], R["rwd_rot_align"]
    safe = (np.abs(pos - 0.075) > 1e-5) & (np.abs(rot - 0.9) > 1e-5) & (np.abs(rot - 0.95) > 1e-5)
    assert safe.sum() > 0.95 * n
    for i, k in enumerate(envs.REORIENT_RWD_TERMS):
        g = R["rwd_" + k].astype(np.float64)
        sel = safe if k in ("drop", "bonus", "solved", "done") else slice(None)
        assert np.abs(terms[:, i][sel] - g[sel]).max() <= 1e-5 * max(1.0, np.abs(g).max()), k
    assert np.abs(e.reward.cpu().numpy()[:, 0][safe] - R["rwd_dense"][safe]).max() < 2e-4


def test_reorient_env_step_semantics(models):
    """myoHandReorient8-v0 through the registry: desired orientations per episode equal the oracle's (shared counter-based
    draws + the reference's euler2quat), the observation equals the oracle's reorient_obs_reward of the bound fields (which
    describe the stepped state), dropped worlds restart from the open-hand pose with a new target."""
    import env_oracle as eo
    import torch
    from myosuite_mjx_b200 import registry
    m = models("hand_sar")
    nw = 16
    env = registry.make("myoHandReorient8-v0", nworld=nw, model=m, horizon=6, seed=0)
    st = env.reset(21)
    d, e, c = env._data, env.env, env.cfg
    assert c["frame_skip"] == 5 and c["dt"] == pytest.approx(0.01)
    want = np.stack([eo.reorient_des_rot(c, 21, w, 0) for w in range(nw)])
    assert np.abs(e.target.cpu().numpy() - want).max() < 2e-6 and np.abs(np.linalg.norm(want, axis=1) - 1).max() < 1e-6
    assert np.array_equal(d.qpos.cpu().numpy(), np.tile(c["init_qpos"], (nw, 1)))
    g = torch.Generator(device="cuda"); g.manual_seed(9)
    nreset = 0
    for t in range(8):
        st = env.step(st, torch.rand((nw, m.nu), generator=g, device="cuda") * 2 - 1)
        f = lambda name: getattr(d, name).cpu().numpy().astype(np.float64)
        xp, sx, gx = f("xpos").reshape(nw, -1, 3), f("site_xpos").reshape(nw, -1, 3), f("geom_xpos").reshape(nw, -1, 3)
        od, r = eo.reorient_obs_reward(c, f("qpos"), f("qvel"), f("act"), xp[:, c["obj_bid"]], sx[:, c["eps_ball_sid"]], gx[:, c["top_gid"]], gx[:, c["bot_gid"]],
                                       e.target.cpu().numpy().astype(np.float64), f("actuator_length"), f("actuator_velocity"), f("actuator_force"))
        ref = np.concatenate([od[k] for k in c["obs_keys"]], axis=-1)
        assert np.abs(st.obs.cpu().numpy() - ref).max() < 1e-4 * max(1.0, np.abs(ref).max()), t
        ended = st.done.cpu().numpy() > 0
        if ended.any():
            nreset += int(ended.sum())
            assert np.array_equal(d.qpos.cpu().numpy()[ended], np.tile(c["init_qpos"], (int(ended.sum()), 1)))
            ep = e.episode.cpu().numpy()[:, 0]
            want = np.stack([eo.reorient_des_rot(c, 21, w, ep[w] - 1) for w in np.nonzero(ended)[0]])
            assert np.abs(e.target.cpu().numpy()[ended] - want).max() < 2e-6
    assert nreset >= nw and torch.isfinite(st.obs).all() and torch.isfinite(st.reward).all()


def test_registry_pose_envs(models):
    """The registered kwargs of the BASELINE pose configs (myobase/__init__.py) reach the device env."""
    import torch
    from myosuite_mjx_b200 import registry
    for env_id, fixture, thd in (("myoElbowPose1D6MRandom-v0", "elbow", 0.175), ("myoFingerPoseRandom-v0", "finger", 0.35), ("myoHandPoseRandom-v0", "hand_pose", 0.7)):
        m = models(fixture)
        env = registry.make(env_id, nworld=32, model=m)
        assert env.cfg["pose_thd"] == pytest.approx(thd) and env.cfg["horizon"] == 100 and env.cfg["frame_skip"] == 10
        st = env.reset(11)
        tr = torch.as_tensor(env.cfg["target_range"], device="cuda")
        assert bool(((env.target >= tr[:, 0] - 1e-6) & (env.target <= tr[:, 1] + 1e-6)).all())
        for _ in range(3):
            st = env.step(st, torch.zeros((32, m.nu), device="cuda"))
        assert torch.isfinite(st.obs).all() and st.obs.shape[1] == 2 * m.nq + m.nv + m.na
    env = registry.make("myoHandPose3Fixed-v0", nworld=4, model=models("hand_pose"))
    env.reset(0)
    assert np.allclose(env.target.cpu().numpy(), np.tile(registry.ASL_QPOS[3].astype(np.float32), (4, 1)))


def test_full_size_properties(models):
    """8192 worlds (BASELINE metric size): replicated worlds stay bit-identical, states stay finite and inside
    a sane envelope, zero-ctrl rest pose of the elbow is a fixed point of the damping-free statics check."""
    import torch
    from myosuite_mjx_b200 import mjx
    m = models("hand")
    nw = 8192
    qpos, qvel, act, ctrl = parity.random_state(m, 64, 2, qpos_spread=0.3)
    rep = nw // 64
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=np.tile(qpos, (rep, 1)), qvel=np.tile(0 * qvel, (rep, 1)), act=np.tile(act, (rep, 1)),
                                 ctrl=np.tile(ctrl, (rep, 1))), nworld=nw, fields="core")
    for _ in range(3):
        mjx.step(model, d, 10)
    torch.cuda.synchronize()
    q = d.qpos.reshape(rep, 64, -1)
    assert torch.equal(q, q[:1].expand_as(q)), "replicated worlds diverged: the kernel is not world-independent"
    assert torch.isfinite(d.qpos).all() and float(d.qvel.abs().max()) < 1e3
    assert float(d.time.min()) == pytest.approx(30 * 0.002, abs=1e-5)
    lo = torch.as_tensor(np.asarray(m.jnt_range)[:, 0], device="cuda", dtype=torch.float32)
    hi = torch.as_tensor(np.asarray(m.jnt_range)[:, 1], device="cuda", dtype=torch.float32)
    assert float((d.qpos - hi).max()) < 0.3 and float((lo - d.qpos).max()) < 0.3      # soft limits hold


@pytest.mark.parametrize("fields", ["core", "all"])   # "all": the state copies run on the side stream under the wide-output kernel
def test_step_host_round_trip(models, fields):
    import ctypes
    import torch
    from myosuite_mjx_b200 import lib, mjx
    m = models("elbow")
    model = mjx.put_model(m)
    nw = 128
    d = mjx.make_data(model, nw, fields=fields)
    e = mjx.make_data(model, nw, fields=fields)
    # host buffers in the ABI layout [k][nworld]
    hq = torch.full((1, nw), 0.4).pin_memory(); hv = torch.zeros((1, nw)).pin_memory()
    ha = torch.zeros((m.na, nw)).pin_memory(); hc = torch.rand((m.nu, nw)).pin_memory()
    oq = torch.empty_like(hq).pin_memory(); ov = torch.empty_like(hv).pin_memory(); oa = torch.empty_like(ha).pin_memory()
    vp = ctypes.c_void_p
    for rep in range(3):   # repeated calls reuse the side stream and its events
        lib.check(lib.lib().myo_step_host(d._h, 3, vp(hq.data_ptr()) if rep == 0 else None, vp(hv.data_ptr()) if rep == 0 else None,
                                          vp(ha.data_ptr()) if rep == 0 else None, vp(hc.data_ptr()),
                                          vp(oq.data_ptr()), vp(ov.data_ptr()), vp(oa.data_ptr()), None))
    e.storage("qpos").copy_(hq.cuda()); e.storage("ctrl").copy_(hc.cuda())
    mjx.step(model, e, 9)
    torch.cuda.synchronize()
    assert torch.equal(oq, e.storage("qpos").cpu()) and torch.equal(ov, e.storage("qvel").cpu()) and torch.equal(oa, e.storage("act").cpu())


def test_length_range_simulation_on_device(models):
    """mj_setLengthRange on the GPU (put_model does it for models without lengthrange): reproduces the fixture."""
    import copy
    from myosuite_mjx_b200 import lengthrange
    m = copy.copy(models("finger"))
    want = np.asarray(m.actuator_lengthrange, dtype=np.float64).reshape(-1, 2).copy()
    m.actuator_lengthrange = np.zeros_like(want)
    assert len(lengthrange.needs_length_range(m)) == m.nu
    ids, rng, spread = lengthrange.set_length_range(m)
    np.testing.assert_allclose(rng, want, rtol=1e-5, atol=1e-7)
    assert spread.max() < 1e-4


def test_reorient8_draws_the_object_shape_per_world():
    """myoHandReorient8-v0 (reference: reorient_sar_v0.py:176-260 Geometries8EnvV0.reset): every episode of every world draws
    the object's collision geom -- capsule / ellipsoid / cylinder / box (geom_type 3..6), one of the two sizes of the reference's
    tables, top and bottom markers moved along -- into the batch's per-world geom override buffers; physics then runs each
    world with its own shape.  The reference draws from np_random: the distribution is reproduced, not the stream."""
    import torch
    from myosuite_mjx_b200 import registry
    nw = 512
    env = registry.make("myoHandReorient8-v0", nworld=nw)
    d = env._data
    g = [str(x) for x in env.sys.mj.names_geom]
    obj, top, bot = g.index("obj"), g.index("top"), g.index("bot")
    table = {3: ([0.013, 0.025, 0.025], [0.019, 0.040, 0.040]), 4: ([0.011, 0.025, 0.025], [0.019, 0.040, 0.040]),
             5: ([0.013, 0.025, 0.025], [0.019, 0.040, 0.040]), 6: ([0.017, 0.017, 0.017], [0.023, 0.023, 0.023])}

    def check():
        t = d.geom_type_ovr[:, obj].cpu().numpy(); s = d.geom_size_ovr[:, obj].cpu().numpy()
        tp = d.geom_pos_ovr[:, top].cpu().numpy(); bp = d.geom_pos_ovr[:, bot].cpu().numpy()
        assert set(np.unique(t)) == {3, 4, 5, 6} and np.bincount(t)[3:].min() > nw // 8, np.bincount(t)
        for w in range(nw):
            opts = table[int(t[w])]
            k = 0 if np.allclose(s[w], opts[0], atol=1e-7) else 1
            assert np.allclose(s[w], opts[k], atol=1e-7), (w, t[w], s[w])
            z = 1.3 * s[w][1] if t[w] == 3 else (s[w][1] if t[w] == 5 else s[w][2])
            assert np.allclose(tp[w], [0, 0, z], atol=1e-7) and np.allclose(bp[w], [0, 0, -z], atol=1e-7)
        others = torch.ones(len(g), dtype=torch.bool); others[obj] = False
        assert (d.geom_type_ovr[:, others.to(d.geom_type_ovr.device)] == -1).all()      # only the object changes shape
        return t.copy(), s.copy()
    st = env.reset(7)
    t0, s0 = check()
    assert torch.isfinite(st.obs).all()
    for _ in range(55):      # past the 50-step horizon: every world starts a new episode with a new draw
        st = env.step(st, torch.rand((nw, env.action_size), device="cuda") * 2 - 1)
    t1, s1 = check()
    assert (t0 != t1).mean() > 0.5 and torch.isfinite(st.obs).all() and torch.isfinite(st.reward).all()
    # the marker geoms follow the drawn shape in the kinematics: |top - bottom| = 2 z per world
    span = (d.geom_xpos[:, top] - d.geom_xpos[:, bot]).norm(dim=1).cpu().numpy()
    z = np.where(t1 == 3, 1.3 * s1[:, 1], np.where(t1 == 5, s1[:, 1], s1[:, 2]))
    assert np.allclose(span, 2 * z, atol=2e-6)
