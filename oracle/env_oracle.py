"""CPU restatement (numpy, float64) of the reference's pose-env layer.  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu legs, never by the product.

Follows (reference file:line):
    action -> ctrl            myosuite/envs/myo/base_v0.py:81-110; myosuite/robot/robot.py:771-783 (default_robot)
    fatigue 3CC-r             myosuite/envs/myo/fatigue.py:61-108
    obs_dict / reward_dict    myosuite/envs/myo/myobase/pose_v0.py:98-138
    reset / target sampling   myosuite/envs/myo/myobase/pose_v0.py:141-168,190-251
Pinned: tests/test_env_oracle.py checks every function here against tests/golden/env_kat.npz, which was produced by
running the reference's own BaseV0.step / CumulativeFatigue / PoseEnvV0 code (tests/golden/make_env_kat.py).
The uniform draws are the library's counter-based hash (csrc/env.cu envUniform), not numpy's generator: the reference
draws from np_random, whose stream a batched device RNG cannot reproduce; the distribution (uniform on the range) is
the same.
"""
import numpy as np

M64 = (1 << 64) - 1


def uniform(seed, gworld, episode, stream, idx):
    """envUniform of csrc/env.cu: splitmix64 finaliser of (seed, world, episode, stream, index) -> float32 in [0, 1)."""
    z = (int(seed) ^ (int(gworld) * 0xD1B54A32D192ED03) ^ (int(episode) * 0x8CB92BA72F3D8DD7) ^ (((int(stream) << 32) | int(idx)) * 0x9E3779B97F4A7C15)) & M64
    z = (z + 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    z ^= z >> 31
    return np.float32(z >> 40) * np.float32(1.0 / 16777216.0)


def action_to_ctrl(a, is_muscle, na, normalize_act, ctrl_lo, ctrl_hi, cond="", epl=-1, eip=-1, fatigue=None):
    """base_v0.py:81-110 + robot.py:771-783.  `a`: [..., nu]; returns ctrl_desired after the robot's re-projection."""
    x = np.array(a, dtype=np.float64)
    mus = np.asarray(is_muscle, dtype=bool)
    if na and normalize_act:
        x[..., mus] = 1.0 / (1.0 + np.exp(-5.0 * (x[..., mus] - 0.5)))
        robot_normalized = False
    else:
        robot_normalized = bool(normalize_act)
    if cond == "fatigue":
        x[..., mus] = fatigue.compute_act(x[..., mus])[0]
    elif cond == "reafferentation":
        x[..., epl] = x[..., eip].copy()
        x[..., eip] = 0
    if robot_normalized:
        x = 0.5 * (np.asarray(ctrl_lo) + np.asarray(ctrl_hi)) + x * (np.asarray(ctrl_hi) - np.asarray(ctrl_lo)) / 2.0
    return x


class Fatigue:
    """fatigue.py:9-108 (3CC-r); state arrays of shape [..., nmuscle]."""

    def __init__(self, tau_act, tau_deact, dt, shape=(), F=0.00912, R=0.1 * 0.00094, r=10 * 15):
        self.tauact, self.taudeact, self.dt, self.F, self.R, self.r = np.asarray(tau_act, float), np.asarray(tau_deact, float), dt, F, R, r
        n = len(self.tauact)
        self.MA, self.MR, self.MF = np.zeros(shape + (n,)), np.ones(shape + (n,)), np.zeros(shape + (n,))

    def reset(self, mask=None):
        m = slice(None) if mask is None else np.asarray(mask, bool)
        self.MA[m], self.MR[m], self.MF[m] = 0.0, 1.0, 0.0

    def compute_act(self, act):
        TL = np.asarray(act, float)
        MA, MR, MF, dt, F = self.MA, self.MR, self.MF, self.dt, self.F
        LD = 1 / self.tauact * (0.5 + 1.5 * MA)
        LR = (0.5 + 1.5 * MA) / self.taudeact
        C = np.where(MA < TL, np.where(MR > TL - MA, LD * (TL - MA), LD * MR), LR * (TL - MA))
        rR = np.where(MA >= TL, self.r * self.R, self.R)
        C = np.clip(C, np.maximum(-MA / dt + F * MA, (MR - 1) / dt + rR * MF), np.minimum((1 - MA) / dt + F * MA, MR / dt + rR * MF))
        dMA, dMR, dMF = (C - F * MA) * dt, (-C + rR * MF) * dt, (F * MA - rR * MF) * dt
        self.MA, self.MR, self.MF = MA + dMA, MR + dMR, MF + dMF
        return self.MA, self.MR, self.MF


def obs_dict(time, qpos, qvel, act, target, dt, target_qposadr=None):
    """pose_v0.py:106-116."""
    qpos = np.asarray(qpos, float)
    sel = qpos if target_qposadr is None else qpos[..., np.asarray(target_qposadr)]
    na = 0 if act is None else np.shape(act)[-1]
    return {"time": np.asarray(time, float)[..., None], "qpos": qpos, "qvel": np.asarray(qvel, float) * dt,
            "act": np.asarray(act, float) if na > 0 else np.zeros_like(qpos), "pose_err": np.asarray(target, float) - sel}


def reward_dict(od, na, pose_thd=0.35, weights=(1.0, 4.0, 1.0, 50.0)):
    """pose_v0.py:118-138; weights in DEFAULT_RWD_KEYS_AND_WEIGHTS order: pose, bonus, act_reg, penalty."""
    pose_dist = np.linalg.norm(od["pose_err"], axis=-1)
    act_mag = np.linalg.norm(od["act"], axis=-1)
    if na != 0:
        act_mag = act_mag / na
    far_th = 4 * np.pi / 2
    r = {"pose": -1.0 * pose_dist, "bonus": 1.0 * (pose_dist < pose_thd) + 1.0 * (pose_dist < 1.5 * pose_thd),
         "penalty": -1.0 * (pose_dist > far_th), "act_reg": -1.0 * act_mag, "sparse": -1.0 * pose_dist,
         "solved": pose_dist < pose_thd, "done": pose_dist > far_th}
    r["dense"] = weights[0] * r["pose"] + weights[1] * r["bonus"] + weights[2] * r["act_reg"] + weights[3] * r["penalty"]
    return r


class PoseEnvOracle:
    """Batched pose env on top of a physics stepper `physics_step(qpos, qvel, act, ctrl, time, n) -> (qpos, qvel, act, time)`
    (the mj_step oracle in tests); mirrors the library's myo_env_reset / myo_env_step including masked auto-reset."""

    def __init__(self, cfg, nworld, physics_step):
        self.c, self.nw, self.physics = cfg, nworld, physics_step
        c = cfg
        self.nq, self.nv, self.na, self.nu = len(c["init_qpos"]), len(c["init_qvel"]), c["na"], len(c["act_is_muscle"])
        self.qpos = np.zeros((nworld, self.nq)); self.qvel = np.zeros((nworld, self.nv)); self.act = np.zeros((nworld, self.na))
        self.ctrl = np.zeros((nworld, self.nu)); self.time = np.zeros(nworld)
        self.target = np.zeros((nworld, len(c["target_qposadr"])))
        self.steps = np.zeros(nworld, int); self.episode = np.zeros(nworld, int)
        self.fat = None
        if c.get("cond") == "fatigue":
            mus = np.asarray(c["act_is_muscle"], bool)
            self.fat = Fatigue(np.asarray(c["tau_act"])[mus], np.asarray(c["tau_deact"])[mus], c["dt"], (nworld,), c.get("fatigue_F", 0.00912),
                               c.get("fatigue_R", 0.1 * 0.00094), c.get("fatigue_r", 150))

    def _reset_world(self, w):
        c = self.c
        gw, ep = c.get("world_offset", 0) + w, self.episode[w]
        for k in range(self.target.shape[1]):
            if c["target_type"] == "generate":
                lo, hi = np.float32(c["target_range"][k][0]), np.float32(c["target_range"][k][1])
                self.target[w, k] = lo + (hi - lo) * uniform(c["seed"], gw, ep, 0, k)
            else:
                self.target[w, k] = c["target_value"][k]
        for j in range(self.nq):
            if c["reset_type"] == "random":
                lo, hi = np.float32(c["reset_lo"][j]), np.float32(c["reset_hi"][j])
                self.qpos[w, j] = lo + (hi - lo) * uniform(c["seed"], gw, ep, 1, j)
            else:
                self.qpos[w, j] = c["init_qpos"][j]
        self.qvel[w] = c["init_qvel"]; self.act[w] = 0; self.ctrl[w] = 0; self.time[w] = 0
        if self.fat is not None:
            m = np.zeros(self.nw, bool); m[w] = True
            self.fat.reset(m)
        self.steps[w] = 0; self.episode[w] += 1

    def _obs(self):
        c = self.c
        od = obs_dict(self.time, self.qpos, self.qvel, self.act if self.na else None, self.target, c["dt"], c["target_qposadr"])
        return od, np.concatenate([od[k] for k in c["obs_keys"]], axis=-1)

    def reset(self, mask=None):
        for w in range(self.nw):
            if mask is None or mask[w]:
                self._reset_world(w)
        return self._obs()[1]

    def step(self, action):
        c = self.c
        self.ctrl = action_to_ctrl(action, c["act_is_muscle"], self.na, c["normalize_act"], c["ctrl_lo"], c["ctrl_hi"], c.get("cond", ""),
                                   c.get("epl", -1), c.get("eip", -1), self.fat)
        self.qpos, self.qvel, self.act, self.time = self.physics(self.qpos, self.qvel, self.act, self.ctrl, self.time, c["frame_skip"])
        od, _ = self._obs()
        r = reward_dict(od, self.na, c["pose_thd"], c["weights"])
        self.steps += 1
        ended = r["done"] | ((c["horizon"] > 0) & (self.steps >= c["horizon"]))
        if c["auto_reset"]:
            for w in np.nonzero(ended)[0]:
                self._reset_world(w)
        return self._obs()[1], r, ended


# ---------------------------------------------------------------------------------------------------- walk env
def quat2mat00(q):
    """[0, 0] entry of myosuite/utils/quat_math.py:160-183 quat2mat."""
    q = np.asarray(q, float)
    Nq = np.sum(q * q, axis=-1)
    s = 2.0 / Nq
    return np.where(Nq > np.finfo(np.float64).eps, 1.0 - (q[..., 2] * (q[..., 2] * s) + q[..., 3] * (q[..., 3] * s)), 1.0)


def walk_obs_reward(c, qpos, qvel, act, cvel, xipos, xpos, xquat, alen, avel, afrc, steps):
    """walk_v0.py:263-310 with the helpers :352-478, batched over the leading axis.  `c`: envs.walk_config dict; `steps`
    is the reference's self.steps at evaluation time.  Returns (obs_dict, reward_dict)."""
    mass = np.asarray(c["body_mass"], float)[:, None]
    com_vel = (np.sum(mass * -np.asarray(cvel, float), axis=-2) / np.sum(mass))[..., 3:5]
    height = (np.sum(mass * np.asarray(xipos, float), axis=-2) / np.sum(mass))[..., 2]
    phase = (np.asarray(steps) / c["hip_period"]) % 1
    fl, fr, pe, to = c["foot_l"], c["foot_r"], c["pelvis"], c["torso"]
    od = {"qpos_without_xy": qpos[..., 2:], "qvel": qvel * c["dt"], "com_vel": com_vel, "torso_angle": xquat[..., to, :],
          "feet_heights": np.stack([xpos[..., fl, 2], xpos[..., fr, 2]], axis=-1), "height": height[..., None],
          "feet_rel_positions": np.concatenate([xpos[..., fl, :] - xpos[..., pe, :], xpos[..., fr, :] - xpos[..., pe, :]], axis=-1),
          "phase_var": phase[..., None], "muscle_length": alen, "muscle_velocity": np.clip(avel, -100, 100),
          "muscle_force": np.clip(afrc / 1000, -100, 100), "act": act}
    vel_reward = np.exp(-np.square(c["target_y_vel"] - com_vel[..., 1])) + np.exp(-np.square(c["target_x_vel"] - com_vel[..., 0]))
    des = np.stack([0.8 * np.cos(phase * 2 * np.pi + np.pi), 0.8 * np.cos(phase * 2 * np.pi)], axis=-1).astype(np.float32)
    cyclic = np.linalg.norm(des - qpos[..., c["hip_flex"]], axis=-1)
    ref_rot = np.exp(-np.linalg.norm(5.0 * (qpos[..., 3:7] - np.asarray(c["target_rot"], float)), axis=-1))
    joint_angle = np.exp(-5 * np.mean(np.abs(qpos[..., c["hip_angle"]]), axis=-1))
    na = act.shape[-1]
    act_mag = np.linalg.norm(act, axis=-1) / na if na else 0
    done = (height < c["min_height"]) | (np.abs(quat2mat00(qpos[..., 3:7])) > c["max_rot"])
    w = c["walk_weights"]
    r = {"vel_reward": vel_reward, "cyclic_hip": cyclic, "ref_rot": ref_rot, "joint_angle_rew": joint_angle, "act_mag": act_mag,
         "sparse": vel_reward, "solved": vel_reward >= 1.0, "done": done}
    r["dense"] = w[0] * vel_reward + w[1] * done + w[2] * cyclic + w[3] * ref_rot + w[4] * joint_angle
    return od, r


# ---------------------------------------------------------------------------------------------------- reorient env
def euler2quat(euler):
    """myosuite/utils/quat_math.py:77-93."""
    euler = np.asarray(euler, float)
    ai, aj, ak = euler[..., 2] / 2, -euler[..., 1] / 2, euler[..., 0] / 2
    si, sj, sk, ci, cj, ck = np.sin(ai), np.sin(aj), np.sin(ak), np.cos(ai), np.cos(aj), np.cos(ak)
    cc, cs, sc, ss = ci * ck, ci * sk, si * ck, si * sk
    return np.stack([cj * cc + sj * ss, cj * cs - sj * sc, -(cj * ss + sj * cc), cj * sc - sj * cs], axis=-1)


def reorient_des_rot(c, seed, gworld, episode):
    """obj_des_rot of a freshly reset world: reorient_sar_v0.py:251-257 with the library's counter-based draws."""
    if c["target_type"] != "generate":
        return np.asarray(c["target_value"], float)
    e0 = np.float32(-1.0) + np.float32(2.0) * uniform(seed, gworld, episode, 0, 0)
    e1 = np.float32(-0.8) + np.float32(2.0) * uniform(seed, gworld, episode, 0, 1)
    w, x, y, z = euler2quat(np.array([e0, e1, 0.0], float))
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return R @ np.asarray(c["target_axis"], float) / c["tar_length"]


def reorient_obs_reward(c, qpos, qvel, act, obj_pos, eps_pos, top_pos, bot_pos, des_rot, alen, avel, afrc):
    """reorient_sar_v0.py:115-173, batched over the leading axis.  `des_rot` is (geom_xpos[t_top] - geom_xpos[t_bot]) / tar_length."""
    obj_rot = (top_pos - bot_pos) / c["pen_length"]
    od = {"hand_jnt": qpos[..., :-6], "obj_pos": obj_pos, "obj_vel": qvel[..., -6:] * c["dt"], "obj_rot": obj_rot, "obj_des_rot": des_rot,
          "obj_err_pos": obj_pos - eps_pos, "obj_err_rot": obj_rot - des_rot, "act": act, "mlen": alen, "mvel": avel, "mforce": afrc}
    pos_align = np.linalg.norm(od["obj_err_pos"], axis=-1)
    rot_align = np.sum(obj_rot * des_rot, axis=-1) / (np.linalg.norm(obj_rot, axis=-1) * np.linalg.norm(des_rot, axis=-1))
    dropped = pos_align > 0.075
    na = act.shape[-1]
    act_mag = np.linalg.norm(act, axis=-1) / na if na else 0
    w = c["reorient_weights"]
    r = {"pos_align": -1.0 * pos_align, "rot_align": rot_align, "act_reg": -1.0 * act_mag, "drop": -1.0 * dropped,
         "bonus": 1.0 * (rot_align > 0.9) * (pos_align < 0.075) + 5.0 * (rot_align > 0.95) * (pos_align < 0.075),
         "sparse": -1.0 * pos_align + rot_align, "solved": (rot_align > 0.95) & (~dropped), "done": dropped}
    r["dense"] = w[0] * r["pos_align"] + w[1] * r["rot_align"] + w[2] * r["act_reg"] + w[3] * r["drop"] + w[4] * r["bonus"]
    return od, r
