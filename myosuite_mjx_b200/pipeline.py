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
    """Joint-pose task with the reset(rng)/step(state, action) shape of the reference's MJX TrackEnv, on the pose
    configs (reference: myosuite/envs/myo/myobase/pose_v0.py:98-138 reward terms, :248-251 random reset,
    myosuite/envs/myo/base_v0.py:85-91 muscle squash).  Observation = concat(qpos, qvel) as myodm_v0.py:297-304."""

    def __init__(self, sys, n_frames=10, nworld=1, reset_type="random", normalize_act=True, pose_thd=0.35, far_th=2 * np.pi, **kw):
        super().__init__(sys, n_frames=n_frames, nworld=nworld, **kw)
        mj = self.sys.mj
        self.reset_type, self.normalize_act, self.pose_thd, self.far_th = reset_type, normalize_act, pose_thd, far_th
        cr = torch.as_tensor(np.asarray(mj.actuator_ctrlrange, dtype=np.float32), device=self.device)
        self._ctrl_lo, self._ctrl_hi = cr[:, 0], cr[:, 1]
        self._is_muscle = torch.as_tensor(np.asarray(mj.actuator_dyntype) == 4, device=self.device)
        hinge = [j for j in range(mj.njnt) if mj.jnt_type[j] in (2, 3)]
        self._qadr = torch.as_tensor([int(mj.jnt_qposadr[j]) for j in hinge], device=self.device)
        rng_ = np.asarray(mj.jnt_range, dtype=np.float32)[hinge]
        lim = np.asarray(mj.jnt_limited)[hinge].astype(bool)
        lo = np.where(lim, rng_[:, 0], -0.5); hi = np.where(lim, rng_[:, 1], 0.5)
        self._jlo = torch.as_tensor(lo, dtype=torch.float32, device=self.device)
        self._jhi = torch.as_tensor(hi, dtype=torch.float32, device=self.device)
        self._qpos0 = torch.as_tensor(np.asarray(mj.qpos0, dtype=np.float32), device=self.device)
        self.target = self._qpos0[self._qadr].expand(self.nworld, -1).clone()

    def _generator(self, rng):
        if isinstance(rng, torch.Generator):
            return rng
        g = torch.Generator(device=self.device)
        g.manual_seed(int(rng))
        return g

    def _get_obs(self, d):
        return torch.cat((d.qpos, d.qvel), dim=1)

    def _uniform_pose(self, g):
        u = torch.rand((self.nworld, self._qadr.numel()), generator=g, device=self.device)
        return self._jlo + u * (self._jhi - self._jlo)

    def reset(self, rng) -> State:
        g = self._generator(rng)
        q = self._qpos0.expand(self.nworld, -1).clone()
        if self.reset_type == "random":
            q[:, self._qadr] = self._uniform_pose(g)
        self.target = self._uniform_pose(g)
        d = self.pipeline_init(q, torch.zeros((self.nworld, self.sys.nv), device=self.device))
        zero = torch.zeros(self.nworld, device=self.device)
        metrics = {"pose": zero.clone(), "bonus": zero.clone(), "act_reg": zero.clone(), "penalty": zero.clone()}
        return State(d, self._get_obs(d), zero.clone(), zero.clone(), metrics)

    def step(self, state: State, action) -> State:
        a = torch.as_tensor(action, dtype=torch.float32, device=self.device).reshape(self.nworld, self.sys.nu)
        if self.normalize_act:
            squashed = 1.0 / (1.0 + torch.exp(-5.0 * (a - 0.5)))                      # base_v0.py:85-91
            affine = (a.clamp(-1, 1) + 1) * (self._ctrl_hi - self._ctrl_lo) * 0.5 + self._ctrl_lo   # myodm_v0.py:273-275
            a = torch.where(self._is_muscle, squashed, affine)
        d = self.pipeline_step(state.pipeline_state, a)
        err = d.qpos[:, self._qadr] - self.target
        dist = torch.linalg.norm(err, dim=1)
        act_mag = torch.linalg.norm(d.act, dim=1) / max(1, self.sys.na) if self.sys.na else torch.zeros_like(dist)
        far = (dist > self.far_th).float()
        metrics = {"pose": -dist, "bonus": (dist < self.pose_thd).float() + (dist < 1.5 * self.pose_thd).float(),
                   "act_reg": -act_mag, "penalty": -far}
        reward = 1.0 * metrics["pose"] + 4.0 * metrics["bonus"] + 1.0 * metrics["act_reg"] + 50.0 * metrics["penalty"]
        state.metrics.update(**metrics)
        return state.replace(pipeline_state=d, obs=self._get_obs(d), reward=reward, done=far)
