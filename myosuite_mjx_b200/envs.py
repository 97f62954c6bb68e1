"""Fused pose-env layer (include/myostep.h "Fused env layer"): ctypes binding of myo_env_* and the batched
counterpart of the reference's PoseEnvV0 (myosuite/envs/myo/myobase/pose_v0.py) on BaseV0
(myosuite/envs/myo/base_v0.py).  Everything runs on the device: one env step = pre kernel (action -> ctrl, muscle
conditions), frame_skip fused mj_steps, post kernel (obs, reward terms, done, masked auto-reset).  No CPU fallback.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import lib as _lib

COND = {"": 0, None: 0, "none": 0, "sarcopenia": 1, "fatigue": 2, "reafferentation": 3}
OBS_KEY = {"time": 0, "qpos": 1, "qvel": 2, "act": 3, "pose_err": 4, "qpos_without_xy": 5, "com_vel": 6, "torso_angle": 7,
           "feet_heights": 8, "height": 9, "feet_rel_positions": 10, "phase_var": 11, "muscle_length": 12, "muscle_velocity": 13,
           "muscle_force": 14, "hand_jnt": 15, "obj_pos": 16, "obj_vel": 17, "obj_rot": 18, "obj_des_rot": 19, "obj_err_pos": 20,
           "obj_err_rot": 21, "mlen": 22, "mvel": 23, "mforce": 24}
MAX_OBS_KEYS = 16
E_FIELDS = ("action", "obs", "reward", "done", "rwd_terms", "target", "fatigue", "steps", "episode")
E_INT = ("steps", "episode")
RWD_TERMS = ("pose", "bonus", "penalty", "act_reg", "sparse", "solved", "done")
WALK_RWD_TERMS = ("vel_reward", "cyclic_hip", "ref_rot", "joint_angle_rew", "act_mag", "sparse", "solved", "done")
REORIENT_RWD_TERMS = ("pos_align", "rot_align", "act_reg", "drop", "bonus", "sparse", "solved", "done")
REORIENT_FIELDS = ("xpos", "site_xpos", "geom_xpos", "actuator_length", "actuator_velocity", "actuator_force")
WALK_FIELDS = ("xpos", "xquat", "xipos", "cvel", "actuator_length", "actuator_velocity", "actuator_force")   # what the walk obs reads

_i32p, _f32p = ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_float)


class EnvDesc(ctypes.Structure):
    """struct myo_env_desc"""
    _fields_ = [("frame_skip", ctypes.c_int32), ("normalize_act", ctypes.c_int32), ("muscle_condition", ctypes.c_int32),
                ("epl_id", ctypes.c_int32), ("eip_id", ctypes.c_int32), ("nobs_keys", ctypes.c_int32), ("obs_keys", ctypes.c_int32 * MAX_OBS_KEYS),
                ("ntarget", ctypes.c_int32), ("target_type", ctypes.c_int32), ("reset_type", ctypes.c_int32), ("auto_reset", ctypes.c_int32),
                ("horizon", ctypes.c_int32), ("world_offset", ctypes.c_int32), ("seed", ctypes.c_uint64),
                ("pose_thd", ctypes.c_float), ("far_th", ctypes.c_float), ("w_pose", ctypes.c_float), ("w_bonus", ctypes.c_float),
                ("w_act_reg", ctypes.c_float), ("w_penalty", ctypes.c_float), ("fatigue_F", ctypes.c_float), ("fatigue_R", ctypes.c_float),
                ("fatigue_r", ctypes.c_float),
                ("target_qposadr", _i32p), ("target_range", _f32p), ("target_value", _f32p), ("init_qpos", _f32p), ("init_qvel", _f32p),
                ("reset_lo", _f32p), ("reset_hi", _f32p), ("act_is_muscle", _i32p), ("ctrl_lo", _f32p), ("ctrl_hi", _f32p),
                ("tau_act", _f32p), ("tau_deact", _f32p),
                ("kind", ctypes.c_int32), ("foot_l_bid", ctypes.c_int32), ("foot_r_bid", ctypes.c_int32), ("pelvis_bid", ctypes.c_int32),
                ("torso_bid", ctypes.c_int32), ("hip_flex_adr", ctypes.c_int32 * 2), ("hip_angle_adr", ctypes.c_int32 * 4),
                ("min_height", ctypes.c_float), ("max_rot", ctypes.c_float), ("hip_period", ctypes.c_float), ("target_x_vel", ctypes.c_float),
                ("target_y_vel", ctypes.c_float), ("target_rot", ctypes.c_float * 4), ("reset_noise", ctypes.c_float),
                ("w_vel", ctypes.c_float), ("w_done", ctypes.c_float), ("w_cyclic", ctypes.c_float), ("w_ref_rot", ctypes.c_float),
                ("w_joint_angle", ctypes.c_float), ("body_mass", _f32p), ("alt_qpos", _f32p), ("alt_qvel", _f32p),
                ("obj_bid", ctypes.c_int32), ("eps_ball_sid", ctypes.c_int32), ("obj_top_gid", ctypes.c_int32), ("obj_bot_gid", ctypes.c_int32),
                ("pen_length", ctypes.c_float), ("tar_length", ctypes.c_float), ("target_axis", ctypes.c_float * 3),
                ("w_pos_align", ctypes.c_float), ("w_rot_align", ctypes.c_float), ("w_r_act_reg", ctypes.c_float), ("w_drop", ctypes.c_float),
                ("w_r_bonus", ctypes.c_float), ("randomize_shape", ctypes.c_int32), ("obj_gid", ctypes.c_int32)]


def _bind_protos(L):
    if getattr(L, "_env_protos", False):
        return
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
    L.myo_env_create.argtypes = [vp, ctypes.POINTER(EnvDesc), ctypes.POINTER(vp)]
    L.myo_env_destroy.argtypes = [vp]; L.myo_env_destroy.restype = None
    L.myo_env_field_count.argtypes = [vp, i32]; L.myo_env_field_count.restype = i64
    L.myo_env_bind.argtypes = [vp, i32, vp, i64]
    L.myo_env_seed.argtypes = [vp, ctypes.c_uint64]
    L.myo_env_reset.argtypes = [vp, vp, vp]
    L.myo_env_step.argtypes = [vp, vp]
    L.myo_env_forward.argtypes = [vp, vp]
    L._env_protos = True


def apply_sarcopenia(mj):
    """base_v0.py:61-66: a weaker muscle has half the peak force (actuator_gainprm[:, 2] *= 0.5); returns a copy of
    the host model to hand to put_model."""
    import copy
    m = copy.copy(mj)
    g = np.array(mj.actuator_gainprm, dtype=np.float64, copy=True)
    b = np.array(mj.actuator_biasprm, dtype=np.float64, copy=True)
    mus = np.asarray(mj.actuator_dyntype) == 4
    g[mus, 2] *= 0.5
    m.actuator_gainprm = g
    m.actuator_biasprm = b
    return m


def pose_config(mj, frame_skip=10, normalize_act=True, muscle_condition="", target_jnt_range=None, target_jnt_value=None,
                reset_type="init", obs_keys=("qpos", "qvel", "pose_err"), weights=None, pose_thd=0.35, far_th=4 * np.pi / 2,
                horizon=100, auto_reset=True, seed=0, world_offset=0, epl_id=-1, eip_id=-1, init_qpos=None, init_qvel=None):
    """Plain-dict env configuration (what myo_env_desc is filled from).
    Defaults are the reference's: pose_v0.py:15-21 (obs keys, reward weights), :55 (pose_thd), base_v0.py:34-38 ("act" is
    appended to obs_keys when the model has activations), frame_skip 10 (base_v0.py:28)."""
    nq, nv, nu, na = int(mj.nq), int(mj.nv), int(mj.nu), int(mj.na)
    obs_keys = list(obs_keys)
    if na > 0 and "act" not in obs_keys:
        obs_keys.append("act")
    w = dict(pose=1.0, bonus=4.0, act_reg=1.0, penalty=50.0)
    w.update(weights or {})
    jr = np.asarray(mj.jnt_range, dtype=np.float32).reshape(-1, 2)
    qadr = np.asarray(mj.jnt_qposadr, dtype=np.int32)
    lo, hi = np.array(mj.qpos0, dtype=np.float32), np.array(mj.qpos0, dtype=np.float32)
    for j in range(int(mj.njnt)):
        if int(mj.jnt_type[j]) in (2, 3):            # slide / hinge: one qpos each (pose_v0.py:248-251 draws over jnt_range)
            lo[qadr[j]], hi[qadr[j]] = jr[j, 0], jr[j, 1]
    if target_jnt_range is not None:
        tr = np.asarray(target_jnt_range, dtype=np.float32).reshape(-1, 2)
        target_type, tv = "generate", tr.mean(axis=1)
    else:
        tr = None
        target_type, tv = "fixed", np.asarray(target_jnt_value if target_jnt_value is not None else mj.qpos0, dtype=np.float32)
    ntarget = len(tv)
    if ntarget != nq:
        raise ValueError("pose targets cover every joint in the reference's pose envs (pose_err = target - qpos)")
    dynprm = np.asarray(mj.actuator_dynprm, dtype=np.float32).reshape(nu, -1) if nu else np.zeros((0, 3), np.float32)
    cr = np.asarray(mj.actuator_ctrlrange, dtype=np.float32).reshape(nu, 2) if nu else np.zeros((0, 2), np.float32)
    return dict(frame_skip=int(frame_skip), normalize_act=bool(normalize_act), cond=muscle_condition or "", epl=int(epl_id), eip=int(eip_id),
                obs_keys=obs_keys, target_qposadr=np.arange(nq, dtype=np.int32), target_type=target_type, target_range=tr, target_value=tv,
                reset_type=reset_type, auto_reset=bool(auto_reset), horizon=int(horizon), world_offset=int(world_offset), seed=int(seed),
                pose_thd=float(pose_thd), far_th=float(far_th), weights=(w["pose"], w["bonus"], w["act_reg"], w["penalty"]),
                init_qpos=np.asarray(init_qpos if init_qpos is not None else mj.qpos0, dtype=np.float32),
                init_qvel=np.asarray(init_qvel if init_qvel is not None else np.zeros(nv), dtype=np.float32),
                reset_lo=lo, reset_hi=hi, act_is_muscle=(np.asarray(mj.actuator_dyntype) == 4).astype(np.int32) if nu else np.zeros(0, np.int32),
                ctrl_lo=cr[:, 0].copy(), ctrl_hi=cr[:, 1].copy(), tau_act=dynprm[:, 0].copy(), tau_deact=dynprm[:, 1].copy(),
                fatigue_F=0.00912, fatigue_R=0.1 * 0.00094, fatigue_r=10.0 * 15, na=na, dt=float(mj.opt_timestep) * int(frame_skip))


def walk_config(mj, frame_skip=10, normalize_act=True, muscle_condition="", min_height=0.8, max_rot=0.8, hip_period=100, reset_type="init",
                target_x_vel=0.0, target_y_vel=1.2, target_rot=None, obs_keys=None, weights=None, horizon=1000, auto_reset=True, seed=0,
                world_offset=0, epl_id=-1, eip_id=-1):
    """WalkEnvV0 configuration (walk_v0.py:186-257; registered kwargs myobase/__init__.py:445-459).  Start states follow
    walk_v0.py:340-350: "init" -> keyframe 2, "random" -> keyframe 2 or 3 plus noise, anything else -> keyframe 0."""
    names_b, names_j = list(mj.names_body), list(mj.names_jnt)
    qadr = lambda n: int(mj.jnt_qposadr[names_j.index(n)])
    if obs_keys is None:
        obs_keys = ["qpos_without_xy", "qvel", "com_vel", "torso_angle", "feet_heights", "height", "feet_rel_positions", "phase_var",
                    "muscle_length", "muscle_velocity", "muscle_force"]          # walk_v0.py:186-198
    obs_keys = list(obs_keys)
    if int(mj.na) > 0 and "act" not in obs_keys:
        obs_keys.append("act")
    w = dict(vel_reward=5.0, done=-100.0, cyclic_hip=-10.0, ref_rot=10.0, joint_angle_rew=5.0)    # walk_v0.py:200-206
    w.update(weights or {})
    kq, kv = np.asarray(mj.key_qpos, dtype=np.float32), np.asarray(mj.key_qvel, dtype=np.float32)
    k0 = 2 if reset_type in ("init", "random") else 0
    base = pose_config(mj, frame_skip=frame_skip, normalize_act=normalize_act, muscle_condition=muscle_condition, obs_keys=("qpos",),
                       horizon=horizon, auto_reset=auto_reset, seed=seed, world_offset=world_offset, epl_id=epl_id, eip_id=eip_id,
                       target_jnt_value=np.zeros(int(mj.nq)), init_qpos=kq[k0], init_qvel=kv[k0])
    base.update(kind="walk", obs_keys=obs_keys, reset_type=reset_type, target_qposadr=np.zeros(0, np.int32), target_value=None, target_range=None,
                foot_l=names_b.index("talus_l"), foot_r=names_b.index("talus_r"), pelvis=names_b.index("pelvis"), torso=names_b.index("torso"),
                hip_flex=[qadr("hip_flexion_l"), qadr("hip_flexion_r")],
                hip_angle=[qadr(n) for n in ("hip_adduction_l", "hip_adduction_r", "hip_rotation_l", "hip_rotation_r")],
                min_height=float(min_height), max_rot=float(max_rot), hip_period=float(hip_period), target_x_vel=float(target_x_vel),
                target_y_vel=float(target_y_vel),
                target_rot=np.asarray(target_rot if target_rot is not None else kq[0][3:7], dtype=np.float32),   # init_qpos = key 0 (walk_v0.py:253)
                reset_noise=0.02, walk_weights=(w["vel_reward"], w["done"], w["cyclic_hip"], w["ref_rot"], w["joint_angle_rew"]),
                body_mass=np.asarray(mj.body_mass, dtype=np.float32), alt_qpos=kq[3], alt_qvel=kv[3])
    return base


def _quat2mat(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def apply_reorient_setup(mj, geometries8=False):
    """ProprioceptiveEnvV0._setup edits the model before the first step: the object is 25 % heavier
    (reorient_sar_v0.py:95); Geometries8EnvV0.reset then sets it to 1.2 at every reset (:247, the same value for every
    episode, so it is a model edit here too).  Returns a copy of the host model to hand to put_model."""
    import copy
    m = copy.copy(mj)
    mass = np.array(mj.body_mass, dtype=np.float64, copy=True)
    ob = list(map(str, mj.names_body)).index("Object")
    mass[ob] = 1.2 if geometries8 else mass[ob] * 1.25
    m.body_mass = mass
    return m


def reorient_config(mj, frame_skip=5, normalize_act=True, muscle_condition="", randomize_target=True, randomize_shape=True, obs_keys=None, weights=None, horizon=50,
                    auto_reset=True, seed=0, world_offset=0, epl_id=-1, eip_id=-1):
    """ProprioceptiveEnvV0 (reorient_sar_v0.py:67-173; registered with frame_skip 5 and a 50-step horizon, myobase/__init__.py:639-649).
    `randomize_target` draws a desired orientation per episode as Geometries8EnvV0.reset does (:251-257); `randomize_shape`
    draws the object's geometry per episode and per world (:178-249: capsule / ellipsoid / cylinder / box in two sizes, top and
    bottom markers moved along) through the batch's geom override buffers."""
    nb, ng, ns = [str(x) for x in mj.names_body], [str(x) for x in mj.names_geom], [str(x) for x in mj.names_site]
    if obs_keys is None:
        obs_keys = ["hand_jnt", "obj_pos", "obj_vel", "obj_rot", "obj_des_rot", "obj_err_pos", "obj_err_rot", "mlen", "mvel", "mforce"]   # :97-108
    obs_keys = list(obs_keys)
    if int(mj.na) > 0 and "act" not in obs_keys:
        obs_keys.append("act")
    w = dict(pos_align=1.0, rot_align=1.0, act_reg=5.0, drop=5.0, bonus=10.0)    # :37-43
    w.update(weights or {})
    gp = np.asarray(mj.geom_pos, dtype=np.float64)
    top, bot, ttop, tbot = (ng.index(n) for n in ("top", "bot", "t_top", "t_bot"))
    pen_length = float(np.linalg.norm(gp[top] - gp[bot])); tar_length = float(np.linalg.norm(gp[ttop] - gp[tbot]))
    axis = gp[ttop] - gp[tbot]
    tb = nb.index("target")
    des = _quat2mat(np.asarray(mj.body_quat, dtype=np.float64)[tb]) @ axis / tar_length      # the XML's target pose (a child of the world body)
    init_qpos = np.array(mj.qpos0, dtype=np.float32)
    init_qpos[:-6] = 0.0; init_qpos[0] = -1.5                                               # :110-111: fully open, palm up
    base = pose_config(mj, frame_skip=frame_skip, normalize_act=normalize_act, muscle_condition=muscle_condition, obs_keys=("qpos",), horizon=horizon,
                       auto_reset=auto_reset, seed=seed, world_offset=world_offset, epl_id=epl_id, eip_id=eip_id, target_jnt_value=np.zeros(int(mj.nq)),
                       init_qpos=init_qpos)
    base.update(kind="reorient", obs_keys=obs_keys, reset_type="init", target_qposadr=np.zeros(3, np.int32), target_range=None,
                target_value=des.astype(np.float32), target_type="generate" if randomize_target else "fixed",
                obj_bid=nb.index("Object"), eps_ball_sid=ns.index("eps_ball"), top_gid=top, bot_gid=bot, pen_length=pen_length, tar_length=tar_length,
                randomize_shape=bool(randomize_shape), obj_gid=ng.index("obj"),
                target_axis=axis.astype(np.float32), reorient_weights=(w["pos_align"], w["rot_align"], w["act_reg"], w["drop"], w["bonus"]))
    return base


class FusedEnv:
    """myo_env handle bound to a `mjx.Data` batch; env buffers are torch tensors in the [k, nworld] layout."""

    def __init__(self, data, cfg):
        L = _lib.lib()
        _bind_protos(L)
        self.data, self.cfg = data, cfg
        d = EnvDesc()
        keep = []

        def arr(v, dt):
            if v is None:
                return None
            a = np.ascontiguousarray(np.asarray(v, dtype=dt))
            keep.append(a)
            return a.ctypes.data_as(_i32p if dt == np.int32 else _f32p)
        d.frame_skip, d.normalize_act, d.muscle_condition = cfg["frame_skip"], int(cfg["normalize_act"]), COND[cfg["cond"]]
        d.epl_id, d.eip_id = cfg["epl"], cfg["eip"]
        d.nobs_keys = len(cfg["obs_keys"])
        for i, k in enumerate(cfg["obs_keys"]):
            d.obs_keys[i] = OBS_KEY[k]
        d.ntarget = len(cfg["target_qposadr"])
        d.target_type = 1 if cfg["target_type"] == "generate" else 0
        d.reset_type = 1 if cfg["reset_type"] == "random" else 0
        d.auto_reset, d.horizon, d.world_offset, d.seed = int(cfg["auto_reset"]), cfg["horizon"], cfg["world_offset"], cfg["seed"]
        d.pose_thd, d.far_th = cfg["pose_thd"], cfg["far_th"]
        d.w_pose, d.w_bonus, d.w_act_reg, d.w_penalty = cfg["weights"]
        d.fatigue_F, d.fatigue_R, d.fatigue_r = cfg["fatigue_F"], cfg["fatigue_R"], cfg["fatigue_r"]
        d.target_qposadr = arr(cfg["target_qposadr"], np.int32)
        d.target_range, d.target_value = arr(cfg["target_range"], np.float32), arr(cfg["target_value"], np.float32)
        d.init_qpos, d.init_qvel = arr(cfg["init_qpos"], np.float32), arr(cfg["init_qvel"], np.float32)
        d.reset_lo, d.reset_hi = arr(cfg["reset_lo"], np.float32), arr(cfg["reset_hi"], np.float32)
        d.act_is_muscle = arr(cfg["act_is_muscle"], np.int32)
        d.ctrl_lo, d.ctrl_hi = arr(cfg["ctrl_lo"], np.float32), arr(cfg["ctrl_hi"], np.float32)
        d.tau_act, d.tau_deact = arr(cfg["tau_act"], np.float32), arr(cfg["tau_deact"], np.float32)
        if cfg.get("kind") == "reorient":
            d.kind = 2
            d.obj_bid, d.eps_ball_sid, d.obj_top_gid, d.obj_bot_gid = cfg["obj_bid"], cfg["eps_ball_sid"], cfg["top_gid"], cfg["bot_gid"]
            d.pen_length, d.tar_length = cfg["pen_length"], cfg["tar_length"]
            for i in range(3):
                d.target_axis[i] = float(cfg["target_axis"][i])
            d.w_pos_align, d.w_rot_align, d.w_r_act_reg, d.w_drop, d.w_r_bonus = cfg["reorient_weights"]
            d.randomize_shape, d.obj_gid = int(cfg.get("randomize_shape", False)), int(cfg.get("obj_gid", -1))
        if cfg.get("kind") == "walk":
            d.kind = 1
            d.foot_l_bid, d.foot_r_bid, d.pelvis_bid, d.torso_bid = cfg["foot_l"], cfg["foot_r"], cfg["pelvis"], cfg["torso"]
            for i in range(2):
                d.hip_flex_adr[i] = cfg["hip_flex"][i]
            for i in range(4):
                d.hip_angle_adr[i] = cfg["hip_angle"][i]
                d.target_rot[i] = float(cfg["target_rot"][i])
            d.min_height, d.max_rot, d.hip_period = cfg["min_height"], cfg["max_rot"], cfg["hip_period"]
            d.target_x_vel, d.target_y_vel, d.reset_noise = cfg["target_x_vel"], cfg["target_y_vel"], cfg["reset_noise"]
            d.w_vel, d.w_done, d.w_cyclic, d.w_ref_rot, d.w_joint_angle = cfg["walk_weights"]
            d.body_mass, d.alt_qpos, d.alt_qvel = arr(cfg["body_mass"], np.float32), arr(cfg["alt_qpos"], np.float32), arr(cfg["alt_qvel"], np.float32)
        h = ctypes.c_void_p()
        _lib.check(L.myo_env_create(data._h, ctypes.byref(d), ctypes.byref(h)))
        self._h = h
        dev = data.qpos.device
        self._t = {}
        for fid, name in enumerate(E_FIELDS):
            cnt = L.myo_env_field_count(h, fid)
            if cnt == 0:
                continue
            t = torch.zeros((cnt, data.nworld), dtype=torch.int32 if name in E_INT else torch.float32, device=dev)
            _lib.check(L.myo_env_bind(h, fid, ctypes.c_void_p(t.data_ptr()), t.numel() * 4))
            self._t[name] = t

    def storage(self, name):
        return self._t[name]

    def __getattr__(self, name):
        t = self.__dict__.get("_t", {})
        if name in t:
            return t[name].t()      # logical [nworld, k] view
        raise AttributeError(name)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def seed(self, seed):
        _lib.check(_lib.lib().myo_env_seed(self._h, ctypes.c_uint64(int(seed) & ((1 << 64) - 1))))

    def reset(self, mask=None):
        if mask is not None:
            mask = mask.to(device=self.data.qpos.device, dtype=torch.int32).reshape(-1).contiguous()
        _lib.check(_lib.lib().myo_env_reset(self._h, ctypes.c_void_p(mask.data_ptr()) if mask is not None else None, self._stream()))

    def step(self, action=None):
        if action is not None:
            a = torch.as_tensor(action, dtype=torch.float32, device=self.data.qpos.device)
            self._t["action"].copy_(a.reshape(self.data.nworld, -1).t())
        _lib.check(_lib.lib().myo_env_step(self._h, self._stream()))

    def forward(self):
        """obs / reward terms / done of the current state (MujocoEnv.forward); no physics, no reset"""
        _lib.check(_lib.lib().myo_env_forward(self._h, self._stream()))

    def __del__(self):
        try:
            _lib.lib().myo_env_destroy(self._h)
        except Exception:
            pass
