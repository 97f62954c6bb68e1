"""ctypes wrapper of the CPU oracle (TEST INFRASTRUCTURE; see mjstep_oracle.cpp header; PARITY UNPINNED).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from myosuite_mjx_b200 import desc as _desc  # noqa: E402  (struct layout only; no product compute)

LIB_PATH = os.path.join(_HERE, "libmjstep_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "mjstep_oracle.cpp")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = ctypes.CDLL(LIB_PATH)
        L.orc_create.restype = ctypes.c_void_p
        L.orc_create.argtypes = [ctypes.POINTER(_desc.ModelDesc), ctypes.c_int]
        L.orc_destroy.argtypes = [ctypes.c_void_p]
        L.orc_reset.argtypes = [ctypes.c_void_p]
        L.orc_forward.argtypes = [ctypes.c_void_p]
        L.orc_step.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_field.restype = ctypes.c_void_p
        L.orc_field.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int)]
        L.orc_scalar_size.argtypes = [ctypes.c_void_p]
        L.orc_wrap.restype = ctypes.c_double
        L.orc_wrap.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_double, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.orc_muscle.argtypes = [ctypes.c_double, ctypes.c_double, ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p,
                                 ctypes.c_void_p, ctypes.c_void_p]
        L.orc_muscle_dynamics.restype = ctypes.c_double
        L.orc_muscle_dynamics.argtypes = [ctypes.c_double, ctypes.c_double, ctypes.c_void_p]
        L.orc_rollout.argtypes = [ctypes.POINTER(_desc.ModelDesc), ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                  ctypes.c_uint64, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.orc_pool_create.restype = ctypes.c_void_p
        L.orc_pool_create.argtypes = [ctypes.POINTER(_desc.ModelDesc), ctypes.c_int, ctypes.c_void_p, ctypes.c_uint64]
        L.orc_pool_destroy.argtypes = [ctypes.c_void_p]
        L.orc_pool_advance.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.orc_pool_state.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


class Oracle:
    """One world.  Attribute access returns a live numpy view of the named mjData field."""

    def __init__(self, model, use_float=False):
        self._L = lib()
        self.model = model
        self._d, self._keep = _desc.pack(model)
        self._h = self._L.orc_create(ctypes.byref(self._d), int(use_float))
        self._ft = np.float32 if use_float else np.float64

    def __del__(self):
        try:
            self._L.orc_destroy(self._h)
        except Exception:
            pass

    def field(self, name):
        n, isint = ctypes.c_int64(0), ctypes.c_int(0)
        p = self._L.orc_field(self._h, name.encode(), ctypes.byref(n), ctypes.byref(isint))
        if not p:
            raise AttributeError(name)
        if n.value == 0:
            return np.zeros(0, dtype=np.int32 if isint.value else self._ft)
        ct = ctypes.c_int32 if isint.value else (ctypes.c_float if self._ft == np.float32 else ctypes.c_double)
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ct)), shape=(n.value,))

    def __getattr__(self, name):
        if name.startswith("_") or name in ("model",):
            raise AttributeError(name)
        if name == "ncon":
            return int(self.field("ncon_v")[0])
        if name == "nefc":
            return int(self.field("nefc_v")[0])
        return self.field(name)

    def set_state(self, qpos=None, qvel=None, act=None, ctrl=None, warmstart=None):
        for nm, v in (("qpos", qpos), ("qvel", qvel), ("act", act), ("ctrl", ctrl), ("qacc_warmstart", warmstart)):
            if v is not None:
                self.field(nm)[:] = np.asarray(v).ravel()

    def reset(self):
        self._L.orc_reset(self._h)

    def forward(self):
        self._L.orc_forward(self._h)

    def step(self, n=1):
        self._L.orc_step(self._h, int(n))


def wrap(x0, x1, gpos, gmat, radius, wtype, side=None):
    a = [np.ascontiguousarray(v, dtype=np.float64) for v in (x0, x1, gpos, np.asarray(gmat).ravel())]
    s = np.ascontiguousarray(side, dtype=np.float64) if side is not None else None
    out = np.zeros(6)
    w = lib().orc_wrap(a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, a[3].ctypes.data, float(radius), int(wtype),
                       s.ctypes.data if s is not None else None, out.ctypes.data)
    return w, out


def muscle(length, vel, lengthrange, acc0, prm):
    lr = np.ascontiguousarray(lengthrange, dtype=np.float64)
    p = np.ascontiguousarray(prm, dtype=np.float64)
    g, b = ctypes.c_double(0), ctypes.c_double(0)
    lib().orc_muscle(float(length), float(vel), lr.ctypes.data, float(acc0), p.ctypes.data, ctypes.byref(g), ctypes.byref(b))
    return g.value, b.value


def muscle_dynamics(ctrl, act, prm):
    p = np.ascontiguousarray(prm, dtype=np.float64)
    return lib().orc_muscle_dynamics(float(ctrl), float(act), p.ctypes.data)


def rollout(model, nworld, nsteps, hold, qpos_init=None, seed=1234, nthreads=1):
    d, keep = _desc.pack(model)
    qi = np.ascontiguousarray(qpos_init, dtype=np.float64) if qpos_init is not None else None
    qo = np.zeros((nworld, model.nq)); vo = np.zeros((nworld, model.nv))
    lib().orc_rollout(ctypes.byref(d), nworld, nsteps, hold, qi.ctypes.data if qi is not None else None, seed, nthreads,
                      qo.ctypes.data, vo.ctypes.data)
    return qo, vo


class Pool:
    """`nworld` worlds of one continuing random-ctrl rollout (state, activations and warm start carried across calls): the
    CPU baseline of bench.py."""

    def __init__(self, model, nworld, qpos_init=None, seed=1234):
        self._L = lib()
        self.model, self.nworld = model, nworld
        self._d, self._keep = _desc.pack(model)
        qi = np.ascontiguousarray(qpos_init, dtype=np.float64) if qpos_init is not None else None
        self._h = self._L.orc_pool_create(ctypes.byref(self._d), nworld, qi.ctypes.data if qi is not None else None, seed)

    def advance(self, nsteps, hold, nthreads=1):
        self._L.orc_pool_advance(self._h, int(nsteps), int(hold), int(nthreads))

    def state(self):
        q = np.zeros((self.nworld, self.model.nq)); v = np.zeros((self.nworld, self.model.nv))
        self._L.orc_pool_state(self._h, q.ctypes.data, v.ctypes.data)
        return q, v

    def __del__(self):
        try:
            self._L.orc_pool_destroy(self._h)
        except Exception:
            pass
