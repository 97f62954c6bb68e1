"""The MJX-style env surface (reset(rng) -> State, step(State, action) -> State, pipeline_init / pipeline_step) on the
GPU, and size-independent properties at BASELINE.json's full batch size (8192 worlds)."""
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
    assert st.obs.shape == (64, m.nq + m.nv) and st.reward.shape == (64,) and st.done.shape == (64,)
    assert set(st.metrics) == {"pose", "bonus", "act_reg", "penalty"}
    lo = torch.as_tensor(np.asarray(m.jnt_range)[:, 0], device="cuda", dtype=torch.float32)
    hi = torch.as_tensor(np.asarray(m.jnt_range)[:, 1], device="cuda", dtype=torch.float32)
    assert bool(((st.pipeline_state.qpos >= lo - 1e-6) & (st.pipeline_state.qpos <= hi + 1e-6)).all())
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
    # same seed, same actions -> same trajectory (determinism, the property tests/test_envs.py:110-120 checks)
    st3 = env.reset(1234)
    g.manual_seed(7)
    for _ in range(5):
        st3 = env.step(st3, torch.rand((64, m.nu), generator=g, device="cuda") * 2 - 1)
    assert torch.equal(st3.obs, st2.obs)


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


def test_step_host_round_trip(models):
    import ctypes
    import torch
    from myosuite_mjx_b200 import lib, mjx
    m = models("elbow")
    model = mjx.put_model(m)
    nw = 128
    d = mjx.make_data(model, nw, fields="core")
    e = mjx.make_data(model, nw, fields="core")
    # host buffers in the ABI layout [k][nworld]
    hq = torch.full((1, nw), 0.4).pin_memory(); hv = torch.zeros((1, nw)).pin_memory()
    ha = torch.zeros((m.na, nw)).pin_memory(); hc = torch.rand((m.nu, nw)).pin_memory()
    oq = torch.empty_like(hq).pin_memory(); ov = torch.empty_like(hv).pin_memory(); oa = torch.empty_like(ha).pin_memory()
    vp = ctypes.c_void_p
    lib.check(lib.lib().myo_step_host(d._h, 3, vp(hq.data_ptr()), vp(hv.data_ptr()), vp(ha.data_ptr()), vp(hc.data_ptr()),
                                      vp(oq.data_ptr()), vp(ov.data_ptr()), vp(oa.data_ptr()), None))
    e.storage("qpos").copy_(hq.cuda()); e.storage("ctrl").copy_(hc.cuda())
    mjx.step(model, e, 3)
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
