"""MJX-shaped host API over libmyostep.so.

Mirrors what the reference calls on `mujoco.mjx` (third-party; call sites myosuite/mjx/play.py:10-11,41,47,49):

    model = mjx.put_model(mj_model)            ->  put_model(mj_model)
    data  = mjx.put_data(mj_model, mj_data)    ->  put_data(model, mj_data, nworld=...)
    data  = jax.jit(mjx.step)(model, data)     ->  step(model, data)          (vmapped: leading world axis)
    mj_d  = mjx.get_data(mj_model, data)       ->  get_data(model, data)

Field names are mjData's; every `Data` attribute is a torch CUDA tensor of logical shape `[nworld, ...]` (the vmapped
leading axis) that the kernel reads and writes in place.  PHYSICAL layout is struct-of-arrays with the world axis
fastest (`storage[k, world]`; the attribute is the transposed, strided view), which is what makes every warp access
one coalesced line.  Zero-copy: the storage is bound to the batch through `myo_batch_bind`; any DLPack producer of a
`[k, nworld]` buffer can be bound the same way, see `Data.bind`.  `step` steps in place by default (a donated
argument); `step(model, data, donate=False)` is the functional, non-aliasing `mjx.step` (two ping-ponged buffer sets).
Unsupported model features raise `NotImplementedError` at `put_model`, as MJX does.  No CPU fallback exists.
"""
from __future__ import annotations

import ctypes
from types import SimpleNamespace

import numpy as np
import torch

from . import desc as _desc
from . import lib as _lib
from . import mjcf as _mjcf

STATE_FIELDS = ("qpos", "qvel", "act", "ctrl", "time", "qacc_warmstart", "status")
# the subset a policy loop needs ("mode B" in DESIGN.md)
CORE_FIELDS = STATE_FIELDS + ("qacc", "actuator_force", "actuator_length", "actuator_velocity")
# optional inputs: bound only on request.  qfrc_applied reads as zero when unbound; the per-world MODEL overrides
# (include/myostep.h: the reference's envs rewrite mjModel per episode, reorient_sar_v0.py:207-245, base_v0.py:64-68) read as
# "the compiled model's value" when unbound and start at their keep values when bound
INPUT_FIELDS = ("qfrc_applied", "geom_type_ovr", "geom_size_ovr", "geom_pos_ovr", "act_gainscale")
_OVR_FILL = {"geom_type_ovr": -1, "geom_size_ovr": -1.0, "geom_pos_ovr": float("nan"), "act_gainscale": 1.0}

_SHAPES = {
    "xpos": ("nbody", 3), "xquat": ("nbody", 4), "xmat": ("nbody", 9), "xipos": ("nbody", 3), "ximat": ("nbody", 9),
    "xanchor": ("njnt", 3), "xaxis": ("njnt", 3), "geom_xpos": ("ngeom", 3), "geom_xmat": ("ngeom", 9),
    "site_xpos": ("nsite", 3), "site_xmat": ("nsite", 9), "subtree_com": ("nbody", 3), "cinert": ("nbody", 10),
    "crb": ("nbody", 10), "cdof": ("nv", 6), "cdof_dot": ("nv", 6), "cvel": ("nbody", 6), "ten_j": ("ntendon", "nv"),
    "actuator_moment": ("nu", "nv"), "qm": ("nv", "nv"), "qld": ("nv", "nv"), "contact_geom": (-1, 2), "contact_pos": (-1, 3),
    "contact_frame": (-1, 9), "contact_friction": (-1, 5), "contact_solref": (-1, 2), "contact_solimp": (-1, 5),
    "efc_j": (-1, "nv"), "wrap_obj": (-1, 2), "wrap_xpos": (-1, 6), "geom_size_ovr": ("ngeom", 3), "geom_pos_ovr": ("ngeom", 3),
}


class Model:
    """Device-resident compiled model (mjx.Model counterpart).  `.mj` is the host mjModel-shaped object."""

    def __init__(self, mj, device=0):
        self.mj = mj
        self.device = int(device)
        self._desc, self._keep = _desc.pack(mj)
        h = ctypes.c_void_p()
        rc = _lib.lib().myo_model_create(ctypes.byref(self._desc), self.device, ctypes.byref(h))
        if rc == -2:
            raise NotImplementedError(_lib.lib().myo_last_error().decode())
        _lib.check(rc)
        self._h = h
        for k in ("nq", "nv", "nu", "na", "nbody", "njnt", "ngeom", "nsite", "ntendon"):
            setattr(self, k, int(getattr(mj, k)))

    def __del__(self):
        try:
            _lib.lib().myo_model_destroy(self._h)
        except Exception:
            pass


class _Batch:
    """Owns one library batch and its buffer sets (two when the functional step ping-pongs).  `Data` objects are light views
    of one set; the batch does not reference them, so it is destroyed as soon as the last view goes."""

    def __init__(self, h):
        self.h, self.sets, self.bound = h, [{}, None], 0

    def __del__(self):
        try:
            _lib.lib().myo_batch_destroy(self.h)
        except Exception:
            pass


class Data:
    """A batch of `nworld` worlds (vmapped mjx.Data counterpart).  Attributes are torch tensors `[nworld, ...]`."""

    def __init__(self, model, nworld, fields="all", nconmax=0, njmax=0):
        L = _lib.lib()
        self.model = model
        self.nworld = int(nworld)
        h = ctypes.c_void_p()
        _lib.check(L.myo_batch_create(model._h, self.nworld, int(nconmax), int(njmax), ctypes.byref(h)))
        self._core, self._idx = _Batch(h), 0
        self.nconmax = L.myo_batch_nconmax(h)
        self.njmax = L.myo_batch_njmax(h)
        self.smem_bytes = L.myo_batch_smem_bytes(h)
        self.team = L.myo_batch_team(h)
        self.ctas_per_sm = L.myo_batch_ctas_per_sm(h)
        self.engine = "phase" if L.myo_batch_engine(h) == 1 else "mono"
        dev = torch.device("cuda", model.device)
        if fields == "all":
            names = [n.lower() for n in _desc.FIELD_NAMES if n.lower() not in INPUT_FIELDS]
        elif fields == "core":
            names = list(CORE_FIELDS)
        else:
            names = list(dict.fromkeys(list(STATE_FIELDS) + [f.lower() for f in fields]))
        for name in names:
            fid = _desc.FIELD_ID[name]
            cnt = L.myo_field_count(h, fid)
            if cnt == 0:
                continue
            dt = torch.int32 if L.myo_field_is_int(fid) else torch.float32
            self.bind(name, torch.full((cnt, self.nworld), _OVR_FILL.get(name, 0), dtype=dt, device=dev))

    @property
    def _h(self):
        return self._core.h

    @property
    def _t(self):
        return self._core.sets[self._idx]

    # -- zero-copy binding of caller-owned device memory (torch tensor or anything exporting __dlpack__)
    def bind(self, name, tensor):
        """`tensor`: contiguous CUDA storage `[k, nworld]` (world fastest), or its `[nworld, k]` transposed view."""
        if not isinstance(tensor, torch.Tensor):
            tensor = torch.from_dlpack(tensor)
        if tensor.dim() == 2 and not tensor.is_contiguous() and tensor.t().is_contiguous():
            tensor = tensor.t()
        if not tensor.is_cuda or not tensor.is_contiguous():
            raise ValueError("bound buffers must be contiguous [k, nworld] CUDA tensors")
        fid = _desc.FIELD_ID[name.lower()]
        self._t[name.lower()] = tensor
        if self._core.bound == self._idx:
            _lib.check(_lib.lib().myo_batch_bind(self._h, fid, ctypes.c_void_p(tensor.data_ptr()), tensor.numel() * tensor.element_size()))

    def _ensure_bound(self):
        """make this view's buffers the ones the library reads and writes (no-op unless the other set is bound)"""
        if self._core.bound != self._idx:
            L = _lib.lib()
            for k, t in self._t.items():
                _lib.check(L.myo_batch_bind(self._h, _desc.FIELD_ID[k], ctypes.c_void_p(t.data_ptr()), t.numel() * t.element_size()))
            self._core.bound = self._idx

    def _partner(self):
        """a view of the other buffer set of the ping-pong (allocated on first use): same batch, same fields, its own tensors"""
        other = self._idx ^ 1
        if self._core.sets[other] is None:
            self._core.sets[other] = {k: torch.zeros_like(v) for k, v in self._t.items()}
        p = Data.__new__(Data)
        p.__dict__.update({k: v for k, v in self.__dict__.items() if k != "_idx"})
        p.__dict__["_idx"] = other
        return p

    def set_pipelined(self, on=True):
        """Opt in to pipelined wide outputs (include/myostep.h, myo_batch_set_pipelined): the wide-output kernel of a step
        runs under the next step's kernel.  Reading a field through this object joins it first, so values read here are
        always complete; tensors fetched earlier and kept by the caller are only complete after `join()`."""
        if not on:
            self.join()
        _lib.check(_lib.lib().myo_batch_set_pipelined(self._h, 1 if on else 0))
        self.__dict__["_pipelined"] = bool(on)
        return self

    def set_count_aware(self, on=True):
        """Opt in to count-aware outputs (include/myostep.h, myo_batch_set_count_aware): `contact_*` / `efc_*` are written only
        up to `ncon` / `nefc`; the padding up to nconmax / njmax is left as it is (read those fields with `[:ncon]` / `[:nefc]`)."""
        _lib.check(_lib.lib().myo_batch_set_count_aware(self._h, 1 if on else 0))
        return self

    def join(self):
        if self.__dict__.get("_pipelined"):
            _lib.check(_lib.lib().myo_outputs_join(self._h, _stream()))

    def __getattr__(self, name):
        core = self.__dict__.get("_core")
        t = core.sets[self.__dict__.get("_idx", 0)] if core is not None else {}
        key = {"q": "qpos", "qd": "qvel"}.get(name, name.lower())   # brax State aliases (myodm_v0.py:201-202)
        if key in t:
            self.join()
            return self._view(key, t[key])
        raise AttributeError(name)

    def _view(self, key, t):
        """storage [k, nworld] -> logical [nworld, ...] strided view (no copy)"""
        shp = _SHAPES.get(key)
        if shp is None:
            return t.t()
        mj = self.model.mj
        dims = [d if isinstance(d, int) else int(getattr(mj, d)) for d in shp]
        if dims[0] == -1:
            dims[0] = t.shape[0] // dims[1]
        return t.view(*dims, self.nworld).permute(len(dims), *range(len(dims)))

    def storage(self, name):
        """the bound `[k, nworld]` buffer itself"""
        self.join()
        return self._t[name.lower()]

    def fields(self):
        return list(self._t)

    def clone(self):
        """an independent batch with copies of every bound buffer"""
        d = Data.__new__(Data)
        L = _lib.lib()
        h = ctypes.c_void_p()
        _lib.check(L.myo_batch_create(self.model._h, self.nworld, self.nconmax, self.njmax, ctypes.byref(h)))
        d.__dict__.update({k: v for k, v in self.__dict__.items() if k not in ("_core", "_idx")})
        d.__dict__["_core"], d.__dict__["_idx"] = _Batch(h), 0
        for k, v in self._t.items():
            d.bind(k, v.clone())
        return d


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def put_model(mj_model, device=0) -> Model:
    """mjx.put_model (reference: myosuite/mjx/play.py:10).  `mj_model`: mjcf.Model, a path to an MJCF file, or a
    real mujoco.MjModel."""
    if isinstance(mj_model, str):
        mj_model = _mjcf.from_xml_path(mj_model)
    from . import lengthrange as _lr
    if len(_lr.needs_length_range(mj_model)):      # MuJoCo does this inside mj_loadXML (mj_setLengthRange)
        _lr.set_length_range(mj_model, device=device)
    return Model(mj_model, device)


def make_data(model, nworld=1, fields="all", nconmax=0, njmax=0) -> Data:
    d = Data(model, nworld, fields, nconmax, njmax)
    d.storage("qpos").copy_(torch.as_tensor(np.asarray(model.mj.qpos0), dtype=torch.float32, device=d.qpos.device)[:, None].expand(-1, d.nworld))
    return d


def put_data(model, mj_data=None, nworld=1, fields="all", nconmax=0, njmax=0) -> Data:
    """mjx.put_data (reference: myosuite/mjx/play.py:11): host state (qpos, qvel, act, ctrl, time, qacc_warmstart;
    each `[n]` or `[nworld, n]`) -> device batch."""
    d = make_data(model, nworld, fields, nconmax, njmax)
    if mj_data is not None:
        for name in ("qpos", "qvel", "act", "ctrl", "qacc_warmstart", "time"):
            v = getattr(mj_data, name, None) if not isinstance(mj_data, dict) else mj_data.get(name)
            if v is None:
                continue
            t = torch.as_tensor(np.asarray(v, dtype=np.float32)).reshape(-1 if np.ndim(v) < 2 else (np.shape(v)[0], -1))
            getattr(d, name).copy_(t.to(d.qpos.device).expand_as(getattr(d, name)) if t.dim() == 1 else t.to(d.qpos.device))
    return d


def _mask_ptr(data, mask):
    if mask is None:
        return None, None
    mask = mask.to(device=data.qpos.device, dtype=torch.int32).reshape(-1).contiguous()
    return ctypes.c_void_p(mask.data_ptr()), mask


def forward(model, data, mask=None) -> Data:
    """mjx.forward / mj_forward (reference: sim.forward(), myosuite/robot/robot.py:979); `mask` restricts it to the
    flagged worlds."""
    data._ensure_bound()
    p, keep = _mask_ptr(data, mask)
    _lib.check(_lib.lib().myo_forward_masked(data._h, p, _stream()) if mask is not None else _lib.lib().myo_forward(data._h, _stream()))
    return data


def refresh(model, data, mask=None) -> Data:
    """mj_step1-style refresh of the position- and velocity-dependent fields from the current state (what dm_control's
    Physics.step leaves behind on the classic path, reference: myosuite/physics/mj_sim_scene.py:55); contacts, forces,
    accelerations and the warm start are left as the last full pass wrote them."""
    data._ensure_bound()
    p, keep = _mask_ptr(data, mask)
    _lib.check(_lib.lib().myo_forward_posvel(data._h, p, _stream()))
    return data


def step(model, data, nsubsteps=1, donate=True) -> Data:
    """mjx.step (reference: myosuite/mjx/play.py:47); `nsubsteps` fuses n mj_steps with ctrl held
    (reference: sim.advance(substeps=n_frames), myosuite/robot/robot.py:880-882).

    `donate=True` (default, the fast path): the batch is stepped in place and the same `Data` is returned -- what
    `jax.jit(mjx.step, donate_argnums=1)` does.  `donate=False` gives mjx.step's functional contract: `data` keeps the old
    state and every old field, the result is ANOTHER `Data` (reference: myodm_v0.py:278-287 reads pipeline_state0 after
    pipeline_step).  The two objects share one batch and ping-pong between two buffer sets (SURVEY.md 8b "Ownership"), so the
    result of step k reuses the buffers of step k-2: keep at most the last two states alive, or `clone()`."""
    if donate:
        data._ensure_bound()
        _lib.check(_lib.lib().myo_step(data._h, int(nsubsteps), _stream()))
        return data
    new = data._partner()
    for k in STATE_FIELDS + INPUT_FIELDS:
        if k in data._t:
            new._t[k].copy_(data._t[k])
    new._ensure_bound()
    _lib.check(_lib.lib().myo_step(new._h, int(nsubsteps), _stream()))
    return new


def reset(model, data, mask=None) -> Data:
    """mj_resetData + mj_forward on the masked worlds (reference: myosuite/robot/robot.py:976-979)."""
    data._ensure_bound()
    if mask is not None:
        mask = mask.to(device=data.qpos.device, dtype=torch.int32).reshape(-1).contiguous()
    _lib.check(_lib.lib().myo_reset(data._h, ctypes.c_void_p(mask.data_ptr()) if mask is not None else None, _stream()))
    return data


def get_data(model, data):
    """mjx.get_data (reference: myosuite/mjx/play.py:49): device batch -> host numpy arrays (one namespace,
    leading world axis kept)."""
    out = SimpleNamespace()
    for k in data.fields():
        setattr(out, k, getattr(data, k).detach().cpu().numpy())
    return out
