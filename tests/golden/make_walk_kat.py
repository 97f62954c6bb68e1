"""Generates tests/golden/walk_kat.npz by RUNNING the reference's own WalkEnvV0 observation / reward code
(/root/reference/myosuite/envs/myo/myobase/walk_v0.py: get_obs_dict :263-283, get_reward_dict :285-310 and the helpers
:352-478 they call) on seeded mjData-shaped inputs of the myoLegs sizes.  The file is loaded by path; mujoco / gym /
the MujocoEnv base class are absent from this image and replaced by inert stand-ins (see make_env_kat.py); quat2mat is
the reference's own (myosuite/utils/quat_math.py).  Run in the build container only:
    python tests/golden/make_walk_kat.py [/root/reference]"""
import importlib.util
import os
import sys
import types
from types import SimpleNamespace

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


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
gym = SimpleNamespace(utils=SimpleNamespace(seeding=SimpleNamespace(np_random=lambda seed=None: (np.random.default_rng(seed), seed)),
                                            EzPickle=type("EzPickle", (), {})))
_mod("myosuite"); _mod("myosuite.utils", gym=gym)
quat_math = _load("myosuite.utils.quat_math", "myosuite/utils/quat_math.py")
_mod("myosuite.envs", env_base=None); _mod("myosuite.envs.myo")
sys.modules["myosuite.envs"].env_base = _mod("myosuite.envs.env_base", MujocoEnv=type("MujocoEnv", (), {}))
_load("myosuite.envs.myo.fatigue", "myosuite/envs/myo/fatigue.py")
_load("myosuite.envs.myo.base_v0", "myosuite/envs/myo/base_v0.py")
_mod("myosuite.envs.myo.myobase")
walk_v0 = _load("myosuite.envs.myo.myobase.walk_v0", "myosuite/envs/myo/myobase/walk_v0.py")

from myosuite_mjx_b200 import mjcf   # only for the model's names / masses / keyframes (our own fixture of myolegs.xml)
mj = mjcf.Model.load(os.path.join(HERE, "..", "..", "myosuite_mjx_b200", "models", "legs.npz"))
names_b, names_j = [str(x) for x in mj.names_body], [str(x) for x in mj.names_jnt]
nq, nv, nu, nbody = int(mj.nq), int(mj.nv), int(mj.nu), int(mj.nbody)
model = SimpleNamespace(body_mass=np.asarray(mj.body_mass, dtype=np.float64), na=int(mj.na), jnt_qposadr=np.asarray(mj.jnt_qposadr),
                        body_name2id=lambda n: names_b.index(n), joint_name2id=lambda n: names_j.index(n))

rng = np.random.default_rng(20261105)
n = 128
key0 = np.asarray(mj.key_qpos, dtype=np.float64)[0]
qpos = key0 + rng.normal(size=(n, nq)) * 0.3
quat = key0[3:7] + rng.normal(size=(n, 4)) * rng.choice([0.02, 0.3, 1.0], size=(n, 1))
qpos[:, 3:7] = quat / np.linalg.norm(quat, axis=1, keepdims=True)
qvel = rng.normal(size=(n, nv)) * 2
act = rng.uniform(0, 1, size=(n, nu))
cvel = rng.normal(size=(n, nbody, 6)); cvel[:, :, 4] -= 1.0          # walking speed about 1 m/s along y
xipos = rng.normal(size=(n, nbody, 3)) * 0.2 + np.array([0, 0, 0.85])  # com height straddles min_height = 0.8
xpos = rng.normal(size=(n, nbody, 3)) * 0.3
xquat = rng.normal(size=(n, nbody, 4)); xquat /= np.linalg.norm(xquat, axis=2, keepdims=True)
alen = rng.uniform(0.05, 0.4, size=(n, nu)); avel = rng.normal(size=(n, nu)) * 60; afrc = -rng.uniform(0, 2.5e5, size=(n, nu))
steps = rng.integers(0, 400, size=n); time = steps * 0.01
# inputs are rounded to float32 (what the device holds), the reference then computes on them in float64
qpos, qvel, act, cvel, xipos, xpos, xquat, alen, avel, afrc = [a.astype(np.float32).astype(np.float64) for a in (qpos, qvel, act, cvel, xipos, xpos, xquat, alen, avel, afrc)]
wt = walk_v0.WalkEnvV0.DEFAULT_RWD_KEYS_AND_WEIGHTS
obs_keys = list(walk_v0.WalkEnvV0.DEFAULT_OBS_KEYS)
okeys = ("qpos_without_xy", "qvel", "com_vel", "torso_angle", "feet_heights", "height", "feet_rel_positions", "phase_var", "muscle_length",
         "muscle_velocity", "muscle_force", "act")
rkeys = ("vel_reward", "cyclic_hip", "ref_rot", "joint_angle_rew", "act_mag", "sparse", "solved", "done", "dense")
obs = {k: [] for k in okeys}; rwd = {k: [] for k in rkeys}
W = walk_v0.WalkEnvV0
for i in range(n):
    data = SimpleNamespace(time=time[i], qpos=qpos[i], qvel=qvel[i], act=act[i], cvel=cvel[i], xipos=xipos[i], body_xpos=xpos[i], body_xquat=xquat[i],
                           actuator_length=alen[i], actuator_velocity=avel[i], actuator_force=afrc[i])
    sim = SimpleNamespace(data=data, model=model)
    s = SimpleNamespace(sim=sim, dt=0.01, steps=int(steps[i]), hip_period=100, target_x_vel=0.0, target_y_vel=1.2, target_rot=None, init_qpos=key0,
                        min_height=0.8, max_rot=0.8, rwd_keys_wt=wt)
    for name in ("_get_com_velocity", "_get_torso_angle", "_get_feet_heights", "_get_height", "_get_feet_relative_position", "muscle_lengths",
                 "muscle_velocities", "muscle_forces", "_get_vel_reward", "_get_cyclic_rew", "_get_ref_rotation_rew", "_get_joint_angle_rew",
                 "_get_done", "_get_rot_condition", "_get_com", "_get_angle"):
        setattr(s, name, types.MethodType(getattr(W, name), s))
    od = W.get_obs_dict(s, sim)
    s.obs_dict = od
    rd = W.get_reward_dict(s, od)
    for k in okeys: obs[k].append(np.asarray(od[k], dtype=np.float64).ravel())
    for k in rkeys: rwd[k].append(float(rd[k]))
f32 = lambda a: a.astype(np.float32)
out = dict(qpos=f32(qpos), qvel=f32(qvel), act=f32(act), cvel=f32(cvel), xipos=f32(xipos), xpos=f32(xpos), xquat=f32(xquat), actuator_length=f32(alen),
           actuator_velocity=f32(avel), actuator_force=f32(afrc), steps=steps, time=time, default_obs_keys=np.array(obs_keys),
           weights=np.array([wt[k] for k in ("vel_reward", "done", "cyclic_hip", "ref_rot", "joint_angle_rew")], dtype=np.float64))
for k in okeys: out["obs_" + k] = np.stack(obs[k])
for k in rkeys: out["rwd_" + k] = np.array(rwd[k])
np.savez_compressed(os.path.join(HERE, "walk_kat.npz"), **out)
print("wrote walk_kat.npz; done fraction %.2f, solved fraction %.2f" % (out["rwd_done"].mean(), out["rwd_solved"].mean()))
print({k: v.shape for k, v in out.items() if k.startswith(("obs_", "rwd_"))})
