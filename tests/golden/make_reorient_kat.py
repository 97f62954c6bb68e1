"""Generates tests/golden/reorient_kat.npz by RUNNING the reference's own ProprioceptiveEnvV0.get_obs_dict /
get_reward_dict (/root/reference/myosuite/envs/myo/myobase/reorient_sar_v0.py:115-173, with calculate_cosine from
myosuite/utils/vector_math.py and euler2quat from myosuite/utils/quat_math.py) on seeded mjData-shaped inputs of the
myohand_sar sizes.  Third-party imports are replaced by inert stand-ins as in make_env_kat.py.  Build container only:
    python tests/golden/make_reorient_kat.py [/root/reference]"""
import importlib.util
import os
import sys
import types
from types import SimpleNamespace

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _mod(name, **attrs):
    m = types.ModuleType(name); m.__dict__.update(attrs); sys.modules[name] = m
    return m


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    m = importlib.util.module_from_spec(spec); sys.modules[name] = m; spec.loader.exec_module(m)
    return m


_mod("mujoco", mjtDyn=SimpleNamespace(mjDYN_MUSCLE=4))
gym = SimpleNamespace(utils=SimpleNamespace(seeding=SimpleNamespace(np_random=lambda seed=None: (np.random.default_rng(seed), seed)),
                                            EzPickle=type("EzPickle", (), {})))
_mod("myosuite"); _mod("myosuite.utils", gym=gym)
quat_math = _load("myosuite.utils.quat_math", "myosuite/utils/quat_math.py")
_load("myosuite.utils.vector_math", "myosuite/utils/vector_math.py")
_mod("myosuite.envs", env_base=None); _mod("myosuite.envs.myo")
sys.modules["myosuite.envs"].env_base = _mod("myosuite.envs.env_base", MujocoEnv=type("MujocoEnv", (), {}))
_load("myosuite.envs.myo.fatigue", "myosuite/envs/myo/fatigue.py")
_load("myosuite.envs.myo.base_v0", "myosuite/envs/myo/base_v0.py")
_mod("myosuite.envs.myo.myobase")
sar = _load("myosuite.envs.myo.myobase.reorient_sar_v0", "myosuite/envs/myo/myobase/reorient_sar_v0.py")

rng = np.random.default_rng(20261106)
n, nq, nv, nu, nbody, ngeom, nsite = 192, 29, 29, 39, 1, 4, 2       # only the rows the functions index are generated
obj_bid, eps_sid, top, bot, ttop, tbot, succ = 0, 0, 0, 1, 2, 3, 1
f32 = lambda a: a.astype(np.float32)
qpos = f32(rng.normal(size=(n, nq))); qvel = f32(rng.normal(size=(n, nv)) * 3); act = f32(rng.uniform(0, 1, size=(n, nu)))
xpos = f32(rng.normal(size=(n, nbody, 3))); site_xpos = f32(rng.normal(size=(n, nsite, 3))); geom_xpos = f32(rng.normal(size=(n, ngeom, 3)))
site_xpos[:, eps_sid] = xpos[:, obj_bid] + f32(rng.normal(size=(n, 3)) * rng.choice([0.01, 0.04, 0.2], size=(n, 1)))      # around the 0.075 drop radius
axis = rng.normal(size=(n, 3)); axis /= np.linalg.norm(axis, axis=1, keepdims=True)
geom_xpos[:, top] = geom_xpos[:, bot] + f32(0.07 * axis)
des = axis + rng.normal(size=(n, 3)) * rng.choice([0.05, 0.3, 1.5], size=(n, 1)); des /= np.linalg.norm(des, axis=1, keepdims=True)   # rot_align around 0.9 / 0.95
geom_xpos[:, ttop] = geom_xpos[:, tbot] + f32(0.07 * des)
alen = f32(rng.uniform(0.05, 0.4, size=(n, nu))); avel = f32(rng.normal(size=(n, nu))); afrc = f32(-rng.uniform(0, 300, size=(n, nu)))
time = rng.uniform(0, 1, n)
wt = sar.ProprioceptiveEnvV0.DEFAULT_RWD_KEYS_AND_WEIGHTS
okeys = ("hand_jnt", "obj_pos", "obj_vel", "obj_rot", "obj_des_rot", "obj_err_pos", "obj_err_rot", "act", "mlen", "mvel", "mforce")
rkeys = ("pos_align", "rot_align", "act_reg", "drop", "bonus", "sparse", "solved", "done", "dense")
obs = {k: [] for k in okeys}; rwd = {k: [] for k in rkeys}
d64 = lambda a: np.asarray(a, dtype=np.float64)
for i in range(n):
    data = SimpleNamespace(time=time[i], qpos=d64(qpos[i]), qvel=d64(qvel[i]), act=d64(act[i]), body_xpos=d64(xpos[i]), site_xpos=d64(site_xpos[i]),
                           geom_xpos=d64(geom_xpos[i]), actuator_length=d64(alen[i]), actuator_velocity=d64(avel[i]), actuator_force=d64(afrc[i]))
    model = SimpleNamespace(na=nu, site_rgba=np.zeros((nsite, 4)))
    sim = SimpleNamespace(data=data, model=model)
    s = SimpleNamespace(sim=sim, dt=0.01, obj_bid=obj_bid, eps_ball_sid=eps_sid, obj_t_gid=top, obj_b_gid=bot, tar_t_gid=ttop, tar_b_gid=tbot,
                        pen_length=0.07, tar_length=0.07, success_indicator_sid=succ, rwd_keys_wt=wt)
    od = sar.ProprioceptiveEnvV0.get_obs_dict(s, sim)
    s.obs_dict = od
    rd = sar.ProprioceptiveEnvV0.get_reward_dict(s, od)
    for k in okeys: obs[k].append(d64(od[k]).ravel())
    for k in rkeys: rwd[k].append(float(rd[k]))
# euler2quat known answers for the per-episode target orientation (reorient_sar_v0.py:251-257)
eul = np.stack([rng.uniform(-1, 1, 64), rng.uniform(-0.8, 1.2, 64), np.zeros(64)], axis=1)
out = dict(qpos=qpos, qvel=qvel, act=act, xpos=xpos, site_xpos=site_xpos, geom_xpos=geom_xpos, actuator_length=alen, actuator_velocity=avel,
           actuator_force=afrc, ids=np.array([obj_bid, eps_sid, top, bot, ttop, tbot]), euler=eul, euler_quat=quat_math.euler2quat(eul),
           weights=np.array([wt[k] for k in ("pos_align", "rot_align", "act_reg", "drop", "bonus")], dtype=np.float64),
           default_obs_keys=np.array(sar.ProprioceptiveEnvV0.DEFAULT_OBS_KEYS))
for k in okeys: out["obs_" + k] = np.stack(obs[k])
for k in rkeys: out["rwd_" + k] = np.array(rwd[k])
np.savez_compressed(os.path.join(HERE, "reorient_kat.npz"), **out)
print("wrote reorient_kat.npz; dropped %.2f, solved %.2f, bonus>0 %.2f" % (out["rwd_done"].mean(), out["rwd_solved"].mean(), (out["rwd_bonus"] > 0).mean()))
