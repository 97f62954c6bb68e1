"""Generates tests/golden/env_kat.npz by RUNNING the reference's own env code on seeded inputs:

    BaseV0.step                      /root/reference/myosuite/envs/myo/base_v0.py:81-120   (action -> ctrl_desired)
    CumulativeFatigue.compute_act    /root/reference/myosuite/envs/myo/fatigue.py:61-108   (3CC-r muscle fatigue)
    PoseEnvV0.get_obs_dict           /root/reference/myosuite/envs/myo/myobase/pose_v0.py:98-116
    PoseEnvV0.get_reward_dict        /root/reference/myosuite/envs/myo/myobase/pose_v0.py:118-138

The three files are loaded from the reference tree by path; their third-party imports (mujoco, gym, dm_control, the
MujocoEnv base class) are absent from this image and are replaced by inert stand-ins that the functions above never
compute with (mjtDyn.mjDYN_MUSCLE = 4 is MuJoCo 3.2.8's enum value).  `self` is a plain namespace holding the
attributes the methods read.  Run in the build container only:
    python tests/golden/make_env_kat.py [/root/reference]
The GPU box has no /root/reference: tests load the committed .npz."""
import importlib.util
import os
import sys
import types
from types import SimpleNamespace

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


_mod("mujoco", mjtDyn=SimpleNamespace(mjDYN_MUSCLE=4))
gym = SimpleNamespace(utils=SimpleNamespace(seeding=SimpleNamespace(np_random=lambda seed=None: (np.random.default_rng(seed), seed))))
_mod("myosuite"); _mod("myosuite.utils", gym=gym)
_mod("myosuite.envs", env_base=None); _mod("myosuite.envs.myo")
sys.modules["myosuite.envs"].env_base = _mod("myosuite.envs.env_base", MujocoEnv=type("MujocoEnv", (), {}))
fatigue = _load("myosuite.envs.myo.fatigue", "myosuite/envs/myo/fatigue.py")
base_v0 = _load("myosuite.envs.myo.base_v0", "myosuite/envs/myo/base_v0.py")
_mod("myosuite.envs.myo.myobase")
pose_v0 = _load("myosuite.envs.myo.myobase.pose_v0", "myosuite/envs/myo/myobase/pose_v0.py")

rng = np.random.default_rng(20261104)
out = {}

# ---------------------------------------------------------------- model stand-in: 39 muscles + 3 non-muscle actuators
nu, nmus = 42, 39
dyntype = np.array([4] * nmus + [0] * (nu - nmus))
dynprm = np.zeros((nu, 10)); dynprm[:nmus, 0] = rng.uniform(0.008, 0.02, nmus); dynprm[:nmus, 1] = rng.uniform(0.03, 0.06, nmus)
timestep, frame_skip = 0.002, 10
mj_model = SimpleNamespace(actuator_dyntype=dyntype, actuator_dynprm=dynprm, opt=SimpleNamespace(timestep=timestep), na=nmus)
out.update(dyntype=dyntype, dynprm=dynprm, timestep=timestep, frame_skip=frame_skip)


class _Robot:
    def step(self, ctrl_desired, ctrl_normalized, step_duration, realTimeSim, render_cbk=None):
        self.ctrl_desired, self.ctrl_normalized = np.array(ctrl_desired, dtype=np.float64), ctrl_normalized
        return self.ctrl_desired


def make_self(cond, normalize_act=True, na=nmus):
    s = SimpleNamespace(sim=SimpleNamespace(model=SimpleNamespace(actuator_dyntype=dyntype, na=na)), normalize_act=normalize_act,
                        muscle_condition=cond, robot=_Robot(), dt=timestep * frame_skip, mujoco_render_frames=False, mj_render=None,
                        forward=lambda **kw: None)
    if cond == "fatigue":
        s.muscle_fatigue = fatigue.CumulativeFatigue(mj_model, frame_skip, seed=0)
    if cond == "reafferentation":
        s.EPLpos, s.EIPpos = 7, 11
    return s


# ---- BaseV0.step: action -> ctrl_desired for every muscle condition, T consecutive env steps
T = 300
actions = rng.uniform(-1.5, 1.5, size=(T, nu))
actions[100:200] = np.clip(actions[100] + 0.02 * rng.normal(size=(100, nu)).cumsum(0), -2, 2)   # slowly varying stretch: fatigue builds up
out["actions"] = actions
for cond in ("", "sarcopenia", "fatigue", "reafferentation"):
    s = make_self(cond)
    ctrl = np.zeros((T, nu)); fat = np.zeros((T, 3, nmus))
    for t in range(T):
        base_v0.BaseV0.step(s, actions[t])
        ctrl[t] = s.robot.ctrl_desired
        assert s.robot.ctrl_normalized is False
        if cond == "fatigue":
            fat[t] = np.stack([s.muscle_fatigue.MA, s.muscle_fatigue.MR, s.muscle_fatigue.MF])
    out["ctrl_" + (cond or "none")] = ctrl
    if cond == "fatigue":
        out["fatigue_states"] = fat
        out["fatigue_F"], out["fatigue_R"], out["fatigue_r"] = s.muscle_fatigue.F, s.muscle_fatigue.R, s.muscle_fatigue.r
# normalize_act False / na == 0: BaseV0.step hands the action on unchanged and tells the robot whether to re-project
s = make_self("", normalize_act=False); base_v0.BaseV0.step(s, actions[0])
out["ctrl_raw"], out["ctrl_raw_isnormalized"] = s.robot.ctrl_desired, s.robot.ctrl_normalized
s = make_self("", normalize_act=True, na=0); base_v0.BaseV0.step(s, actions[0])
out["ctrl_na0"], out["ctrl_na0_isnormalized"] = s.robot.ctrl_desired, s.robot.ctrl_normalized

# ---- PoseEnvV0.get_obs_dict / get_reward_dict on random states (hand-sized: nq = nv = 23, na = 39)
n, nq, na = 512, 23, 39
qpos = rng.uniform(-1.2, 1.6, size=(n, nq)); qvel = rng.normal(size=(n, nq)) * 3; act = rng.uniform(0, 1, size=(n, na))
target = rng.uniform(-1.0, 1.5, size=(n, nq)); time = rng.uniform(0, 2, size=n)
near = rng.choice(n, 200, replace=False)       # worlds near / at the target exercise the bonus and solved thresholds
target[near] = qpos[near] + rng.normal(size=(200, nq)) * rng.choice([0.0, 0.03, 0.07, 0.1], size=(200, 1))
far = rng.choice(n, 40, replace=False); target[far] += 2.0    # beyond far_th = 2 pi: penalty and done
keys = ("time", "qpos", "qvel", "act", "pose_err")
obs = {k: [] for k in keys}
rkeys = ("pose", "bonus", "penalty", "act_reg", "sparse", "solved", "done", "dense")
rwd = {k: [] for k in rkeys}
wt = pose_v0.PoseEnvV0.DEFAULT_RWD_KEYS_AND_WEIGHTS
for i in range(n):
    sim = SimpleNamespace(data=SimpleNamespace(time=time[i], qpos=qpos[i], qvel=qvel[i], act=act[i]), model=SimpleNamespace(na=na))
    s = SimpleNamespace(dt=timestep * frame_skip, target_jnt_value=target[i], sim=sim, pose_thd=0.35, rwd_keys_wt=wt)
    od = pose_v0.PoseEnvV0.get_obs_dict(s, sim)
    s.obs_dict = od
    rd = pose_v0.PoseEnvV0.get_reward_dict(s, od)
    for k in keys: obs[k].append(np.asarray(od[k], dtype=np.float64))
    for k in rkeys: rwd[k].append(float(rd[k]))
out.update(pose_qpos=qpos, pose_qvel=qvel, pose_act=act, pose_target=target, pose_time=time, pose_thd=0.35,
           pose_weights=np.array([wt["pose"], wt["bonus"], wt["act_reg"], wt["penalty"]], dtype=np.float64),
           default_obs_keys=np.array(pose_v0.PoseEnvV0.DEFAULT_OBS_KEYS))
for k in keys: out["obs_" + k] = np.stack(obs[k])
for k in rkeys: out["rwd_" + k] = np.array(rwd[k])
np.savez_compressed(os.path.join(HERE, "env_kat.npz"), **out)
print("wrote env_kat.npz:", {k: np.shape(v) for k, v in out.items()})
