"""brax-shaped env layer over the batched step: the API surface the reference's MJX env is written against.

    brax PipelineEnv(sys, n_frames, backend)        reference: myosuite/mjx/myodm_v0.py:43-49
    pipeline_init(q, qd) -> pipeline_state          reference: myosuite/mjx/myodm_v0.py:159
    pipeline_step(pipeline_state, action) -> state  reference: myosuite/mjx/myodm_v0.py:287  (n_frames x mjx.step, ctrl = action)
    env.reset(rng) -> State(pipeline_state, obs, reward, done, metrics)      reference: myodm_v0.py:152-173
    env.step(State, action) -> State                                          reference: myodm_v0.py:269-295

`pipeline_state` is an `mjx.Data` batch (leading world axis = the vmapped axis of brax's wrappers); the aliases
brax adds (`q`, `qd`, `x.pos`, `xd`) are exposed as `q`, `qd`.  Everything stays on the device; the n_frames substeps
run inside ONE kernel launch.  In-place semantics: the returned State shares the batch's tensors (clone the
pipeline_state to keep an old one).
"""
from __future__ import annotations

import dataclasses
from typing import Any, Dict, Optional

import numpy as np
import torch

from . import mjx


@dataclasses.dataclass
class State:
    """brax.envs.base.State counterpart."""
    pipeline_state: Any
    obs: torch.Tensor
    reward: torch.Tensor
    done: torch.Tensor
    metrics: Dict[str, torch.Tensor] = dataclasses.field(default_factory=dict)
    info: Dict[str, Any] = dataclasses.field(default_factory=dict)

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)


class PipelineEnv:
    """brax.envs.base.PipelineEnv counterpart bound to `nworld` batched worlds."""

    def __init__(self, sys, n_frames: int = 1, nworld: int = 1, backend: str = "mjx", device: int = 0, fields="core",
                 nconmax: int = 0, njmax: int = 0):
        if backend != "mjx":
            raise NotImplementedError("only backend='mjx' exists (reference: myosuite/mjx/myodm_v0.py:47)")
        self.sys = sys if isinstance(sys, mjx.Model) else mjx.put_model(sys, device=device)
        self._n_frames = int(n_frames)
        self.nworld = int(nworld)
        self._data = mjx.make_data(self.sys, self.nworld, fields=fields, nconmax=nconmax, njmax=njmax)
        self.device = self._data.qpos.device

    @property
    def dt(self) -> float:
        return float(self.sys.mj.opt_timestep) * self._n_frames

    @property
    def action_size(self) -> int:
        return self.sys.nu

    def _batched(self, v, n):
        t = torch.as_tensor(np.asarray(v, dtype=np.float32) if not isinstance(v, torch.Tensor) else v, dtype=torch.float32)
        t = t.to(self.device)
        return t.expand(self.nworld, n) if t.dim() == 1 else t.reshape(self.nworld, n)

    def pipeline_init(self, q, qd, act=None, ctrl=None):
        """mj_resetData + state write + mj_forward (brax pipeline_init; classic path: myosuite/robot/robot.py:976-979)."""
        d = self._data
        d.qpos.copy_(self._batched(q, self.sys.nq)); d.qvel.copy_(self._batched(qd, self.sys.nv))
        if self.sys.na:
            d.act.copy_(self._batched(act, self.sys.na)) if act is not None else d.act.zero_()
        if self.sys.nu:
            d.ctrl.copy_(self._batched(ctrl, self.sys.nu)) if ctrl is not None else d.ctrl.zero_()
        d.time.zero_(); d.qacc_warmstart.zero_(); d.status.zero_()
        return mjx.forward(self.sys, d)

    def pipeline_step(self, pipeline_state, action):
        """n_frames x mj_step with ctrl = action held (brax pipeline_step)."""
        pipeline_state.ctrl.copy_(self._batched(action, self.sys.nu))
        return mjx.step(self.sys, pipeline_state, self._n_frames)


class PoseEnv(PipelineEnv):
    """Joint-pose task (reference: myosuite/envs/myo/myobase/pose_v0.py on myosuite/envs/myo/base_v0.py) with the
    reset(rng) -> State / step(State, action) -> State shape of the reference's MJX TrackEnv (myodm_v0.py:152-173,
    269-295).  One env step is three launches on the device (envs.FusedEnv: action -> ctrl with the muscle squash and the
    muscle conditions, frame_skip fused mj_steps, obs / reward terms / done / masked auto-reset); nothing runs on the
    host.  Observation = concat of obs_keys (default qpos, qvel * dt, pose_err, act; pose_v0.py:15, base_v0.py:34-38)."""

    def __init__(self, sys, n_frames=10, nworld=1, reset_type="random", target_type="generate", normalize_act=True, pose_thd=0.35,
                 muscle_condition="", target_jnt_range=None, target_jnt_value=None, horizon=100, auto_reset=True, world_offset=0,
                 obs_keys=("qpos", "qvel", "pose_err"), weights=None, epl_id=-1, eip_id=-1, **kw):
        from . import envs
        if muscle_condition == "sarcopenia" and not isinstance(sys, mjx.Model):
            sys = envs.apply_sarcopenia(sys)
        super().__init__(sys, n_frames=n_frames, nworld=nworld, **kw)
        mj = self.sys.mj
        if target_type == "generate" and target_jnt_range is None:
            c0 = envs.pose_config(mj)
            target_jnt_range = np.stack([c0["reset_lo"], c0["reset_hi"]], axis=1)   # every joint over its full range
        if target_type == "fixed":
            target_jnt_range = None
        self.cfg = envs.pose_config(mj, frame_skip=n_frames, normalize_act=normalize_act, muscle_condition=muscle_condition,
                                    target_jnt_range=target_jnt_range, target_jnt_value=target_jnt_value, reset_type=reset_type,
                                    obs_keys=obs_keys, weights=weights, pose_thd=pose_thd, horizon=horizon, auto_reset=auto_reset,
                                    world_offset=world_offset, epl_id=epl_id, eip_id=eip_id)
        self.env = envs.FusedEnv(self._data, self.cfg)

    @property
    def target(self):
        return self.env.target

    def _metrics(self):
        terms = self.env.rwd_terms
        from .envs import RWD_TERMS
        return {k: terms[:, i] for i, k in enumerate(RWD_TERMS)}

    def reset(self, rng) -> State:
        self.env.seed(int(rng))
        self.env.storage("episode").zero_()
        self.env.reset()
        d = mjx.forward(self.sys, self._data)
        return State(d, self.env.obs, self.env.reward[:, 0], self.env.done[:, 0], self._metrics())

    def step(self, state: State, action) -> State:
        self.env.step(action)
        state.metrics.update(**self._metrics())
        return state.replace(pipeline_state=self._data, obs=self.env.obs, reward=self.env.reward[:, 0], done=self.env.done[:, 0])


class WalkEnv(PipelineEnv):
    """myoLegWalk (reference: myosuite/envs/myo/myobase/walk_v0.py WalkEnvV0) in the reset(rng) / step(State, action)
    shape.  The observation reads position- and velocity-dependent mjData fields of the stepped state, as the classic
    path does (dm_control's Physics.step ends with mj_step1): see myo_forward_posvel in include/myostep.h."""

    def __init__(self, sys, n_frames=10, nworld=1, fields=None, **kw):
        from . import envs
        base_kw = {k: kw.pop(k) for k in ("device", "nconmax", "njmax", "backend") if k in kw}
        if fields is None:
            fields = mjx.CORE_FIELDS + envs.WALK_FIELDS
        super().__init__(sys, n_frames=n_frames, nworld=nworld, fields=fields, **base_kw)
        self.cfg = envs.walk_config(self.sys.mj, frame_skip=n_frames, **kw)
        self.env = envs.FusedEnv(self._data, self.cfg)

    def _metrics(self):
        terms = self.env.rwd_terms
        from .envs import WALK_RWD_TERMS
        return {k: terms[:, i] for i, k in enumerate(WALK_RWD_TERMS)}

    def reset(self, rng) -> State:
        self.env.seed(int(rng))
        self.env.storage("episode").zero_()
        self.env.reset()
        return State(self._data, self.env.obs, self.env.reward[:, 0], self.env.done[:, 0], self._metrics())

    def step(self, state: State, action) -> State:
        self.env.step(action)
        state.metrics.update(**self._metrics())
        return state.replace(pipeline_state=self._data, obs=self.env.obs, reward=self.env.reward[:, 0], done=self.env.done[:, 0])


class ReorientEnv(WalkEnv):
    """myoHandReorient (reference: myosuite/envs/myo/myobase/reorient_sar_v0.py ProprioceptiveEnvV0 with the per-episode
    target orientation AND object shape of Geometries8EnvV0.reset, drawn per world) in the reset(rng) / step(State, action)
    shape."""

    def __init__(self, sys, n_frames=5, nworld=1, fields=None, **kw):
        from . import envs
        base_kw = {k: kw.pop(k) for k in ("device", "nconmax", "njmax", "backend") if k in kw}
        if not isinstance(sys, mjx.Model):
            sys = envs.apply_reorient_setup(sys, geometries8=kw.get("randomize_shape", True))
        if fields is None:
            fields = mjx.CORE_FIELDS + envs.REORIENT_FIELDS
            if kw.get("randomize_shape", True):
                fields = fields + ("geom_type_ovr", "geom_size_ovr", "geom_pos_ovr")
        PipelineEnv.__init__(self, sys, n_frames=n_frames, nworld=nworld, fields=fields, **base_kw)
        self.cfg = envs.reorient_config(self.sys.mj, frame_skip=n_frames, **kw)
        self.env = envs.FusedEnv(self._data, self.cfg)

    def _metrics(self):
        terms = self.env.rwd_terms
        from .envs import REORIENT_RWD_TERMS
        return {k: terms[:, i] for i, k in enumerate(REORIENT_RWD_TERMS)}
